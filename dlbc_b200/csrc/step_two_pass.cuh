// step_two_pass.cuh -- kernel variant DLBC_KERNEL_TWO_PASS: (K1) density/psi kernel +
// (K2) fused pull + bounce-back + Shan-Chen force + BGK collide kernel, one thread per site.
//
// Replaces the reference's seven sweeps per timestep (main.d:179-195):
//   K1: advectFieldsKernel's gather (advection.d:61-89) -> densityField (density.d:89-96)
//       -> psiField (force.d:360-367), written as psi[c][site]
//   K2: advectFieldsKernel again (same gather) -> addShanChenForcePsi (force.d:238-309, psi of
//       the 18 neighbours read through L1/L2) -> precalculateWeightedVelocity (collision.d:156-196)
//       -> collideFieldsKernel (collision.d:90-119), written once as f_new.
// resetForce / distributeForce vanish: the force lives in registers and forceDistributed == 0
// when electrokinetics is off (force.d:179-197 then adds +0.0).
//
// Addressing: block-uniform 64-bit plane bases + 32-bit per-thread offsets (lattice.cuh).
//
// DRAM traffic per lattice-site update, C components: K1 reads 152 C, writes 8 C;
// K2 reads 152 C + 8 C (+1 mask byte), writes 152 C  ==> 472 C + 1 bytes (945 B at C = 2).
#pragma once
#include "lattice.cuh"

#ifndef DLBC_SOLID_COPY
#define DLBC_SOLID_COPY 1  // 1: the step kernels carry the populations of solid sites along (see fused_site)
#endif
#ifndef DLBC_MASK_BLOCKS
#define DLBC_MASK_BLOCKS 3  // resident 128-thread blocks per SM of the MASK instantiations: stand-alone fused kernel
#endif
#ifndef DLBC_MASK_BLOCKS_WAVE
#define DLBC_MASK_BLOCKS_WAVE 3  // ... and single-launch kernel
#endif
namespace dlbc {

// ---- K1 -------------------------------------------------------------------------------------
// psi of all components at one site of ghost-inclusive plane zg (and its periodic ghost copies)
// Mask-aware gather.  The wave variant (POL != POL_DEFAULT) steers bounce-back from the neighbourhood bits
// (porous 512^3: 4.53 -> 4.82 GLUPS); the stand-alone kernels measured 6 % slower with the bits than with
// the mask bytes (fused 21.7 -> 23.0 ms), so they keep the bytes.
template <bool MASK, int POL, bool SELF = false>
__device__ __forceinline__ void pull_site(const double *__restrict__ ps, const uint8_t *__restrict__ ms, uint32_t bits,
                                          const Geom &g, const Nbr &n, double (&f)[Q]) {
  if (MASK && POL != POL_DEFAULT) pull19_bits<POL, SELF>(ps, bits, g, n, f);
  else pull19<MASK, POL, SELF>(ps, ms, g, n, f);
}

// site's neighbourhood bits: from Fields::nb in the wave variant, else just the own-site flag from the bytes
template <bool MASK, int POL>
__device__ __forceinline__ uint32_t site_bits(const Geom &g, const Fields &fl, int zg, const Nbr &n) {
  if (!MASK) return 0u;
  if (POL != POL_DEFAULT) return fl.nb[(int64_t)zg * g.plane + n.sxy];
  return mask_plane(fl.mask, g, zg)[n.sxy] ? 1u : 0u;
}

// BATCH: gather the populations of ALL components before the first sum, so that C*19 loads are in
// flight per thread instead of 19 (the wave variant runs its psi items at 16 warps/SM, where the
// serial form spent two full memory latencies per site: profiles/r01_wave_before.md).
template <int C, int PSI, bool MASK, int POL = POL_DEFAULT, bool BATCH = false>
__device__ __forceinline__ void psi_site(const Geom &g, const Fields &fl, int x, int y, int zg) {
  const Nbr n = make_nbr(g, x, y);
  const uint8_t *ms = mask_plane(fl.mask, g, zg) + n.sxy;
  const uint32_t bits = site_bits<MASK, POL>(g, fl, zg, n);
  // psi of solid sites is never read (force.d:273 uses the own-site psi)
  if (MASK && (bits & 1u)) {
    if (DLBC_SOLID_COPY) {
      // ... but written all the same (0.0, no loads): the warp's stores then cover whole 32-byte sectors
#pragma unroll
      for (int c = 0; c < C; c++) {
        psi_plane(fl.psi, g, zg, c)[n.sxy] = 0.0;
        if (g.periodic_z_local) {
          if (zg == g.nz) psi_plane(fl.psi, g, 0, c)[n.sxy] = 0.0;
          if (zg == 1) psi_plane(fl.psi, g, g.nz + 1, c)[n.sxy] = 0.0;
        }
      }
    }
    return;
  }
  const double *ps = f_plane(fl.f_old, g, zg, 0) + n.sxy;
  double v[C];
  if (BATCH) {
    double f[C][Q];
#pragma unroll
    for (int c = 0; c < C; c++) pull_site<MASK, POL>(ps + c * g.fc, ms, bits, g, n, f[c]);
#pragma unroll
    for (int c = 0; c < C; c++) v[c] = psi_of<PSI>(density19(f[c]));
  } else {
#pragma unroll
    for (int c = 0; c < C; c++) {
      double f[Q];
      pull_site<MASK, POL>(ps + c * g.fc, ms, bits, g, n, f);
      v[c] = psi_of<PSI>(density19(f));
    }
  }
#pragma unroll
  for (int c = 0; c < C; c++) {
    psi_plane(fl.psi, g, zg, c)[n.sxy] = v[c];
    if (g.periodic_z_local) {
      if (zg == g.nz) psi_plane(fl.psi, g, 0, c)[n.sxy] = v[c];         // -> ghost plane 0
      if (zg == 1) psi_plane(fl.psi, g, g.nz + 1, c)[n.sxy] = v[c];     // -> ghost plane nz+1
    }
  }
}

template <int C, int PSI, bool MASK>
__global__ void __launch_bounds__(128) psi_kernel(Geom g, Fields fl, int z_begin) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= g.nx) return;
  psi_site<C, PSI, MASK>(g, fl, x, blockIdx.y, z_begin + blockIdx.z);  // ghost-inclusive plane index, 1..nz
}

// ---- Shan-Chen force for all components of one site (force.d:238-309) -----------------------
// Accumulation order per component f1 is the reference's: f2 outer, q = 1..18 inner, then the
// wall term q = 0..18; prefactor = ((w_q * psi1) * psi2) * cc.
// COHERENT: psi was written by other blocks of the SAME launch (step_wave.cuh): L1-bypassing loads
// (ld.global.cg) instead of the read-only (ld.global.nc) path.
template <int C, bool MASK, bool EXACT, bool COHERENT = false>
__device__ __forceinline__ void shan_chen_force(const Geom &g, const Fields &fl, const StepParams &P, int zg,
                                                const Nbr &n, uint32_t bits, const double (&psi1)[C], double (&F)[C][3]) {
  if (!COHERENT) {
    const uint8_t *ms = mask_plane(fl.mask, g, zg) + n.sxy;
    const double *pp = psi_plane(fl.psi, g, zg, 0) + n.sxy;
    const int64_t cstride = g.sites_g;
    sc_force<C, EXACT>(P, psi1, F, [&](int f2, int q) -> double {
      const int d = n.ox[1 + cx(q)] + n.oy[1 + cy(q)] + cz(q) * g.plane;
      if (MASK && ms[d]) return psi1[f2];
      return __ldg(pp + f2 * cstride + d);
    });
  } else {
    // psi was written by other blocks of this launch: L2 loads (ld.global.cg), 32-bit-offset addressing
    const double *pp[C];
#pragma unroll
    for (int c = 0; c < C; c++) pp[c] = psi_plane(fl.psi, g, zg, c) + n.sxy;
    sc_force<C, EXACT>(P, psi1, F, [&](int f2, int q) -> double {
      const int d = n.ox[1 + cx(q)] + n.oy[1 + cy(q)] + cz(q) * g.plane;
      if (MASK && ((bits >> q) & 1u)) return psi1[f2];
      return ld_global_cg(ptr_at(pp[f2], d));
    });
  }
}

// NOTE on ordering: the reference finishes ALL f2 for a given f1 (then the wall term of f1)
// before moving to the next f1; since F[f1] only ever receives f1's own terms, interleaving the
// f1 loops as above leaves each F[f1]'s operation sequence unchanged.

// BITS: `bits` holds the whole neighbourhood (wave variant); otherwise the mask bytes are read
template <int C, bool BITS = false>
__device__ __forceinline__ void wall_force(const Geom &g, const Fields &fl, const StepParams &P, int zg,
                                           const Nbr &n, uint32_t bits, const double (&psi1)[C], double (&F)[C][3]) {
  const uint8_t *mp = mask_plane(fl.mask, g, zg);
#pragma unroll
  for (int f1 = 0; f1 < C; f1++) {
    const double wc = P.gwc[f1];
    if (wc == 0.0) continue;
    const double *__restrict__ rw = psi_plane(fl.rho_wall, g, zg, f1);  // same [c][zg][y][x] layout as psi
#pragma unroll
    for (int q = 0; q < Q; q++) {
      const int nb = n.at(cx(q), cy(q)) + cz(q) * g.plane;
      if (BITS ? ((bits >> q) & 1u) != 0u : mp[nb] != 0) {
        const double pref = weight(q) * psi1[f1] * rw[nb] * wc;
        if (cx(q) == 1) F[f1][0] -= pref; else if (cx(q) == -1) F[f1][0] += pref;
        if (cy(q) == 1) F[f1][1] -= pref; else if (cy(q) == -1) F[f1][1] += pref;
        if (cz(q) == 1) F[f1][2] -= pref; else if (cz(q) == -1) F[f1][2] += pref;
      }
    }
  }
}

// BDist2 common velocity (collision.d:156-196): u = (sum_f m_f / tau_f) / (sum_f rho_f)
template <int C, bool EXACT>
__device__ __forceinline__ void common_velocity(const StepParams &P, const double (&m)[C][3],
                                                const double (&rho)[C], double (&u)[3]) {
  double total = 0.0;
#pragma unroll
  for (int c = 0; c < C; c++) total += rho[c];
  const double inv_total = EXACT ? 0.0 : 1.0 / total;
#pragma unroll
  for (int d = 0; d < 3; d++) {
    double wv = 0.0;
#pragma unroll
    for (int c = 0; c < C; c++) {
      // x / tau == x * (1/tau) bit-for-bit when tau is a power of two
      wv += (EXACT && !P.tau_pow2[c]) ? m[c][d] / P.tau[c] : m[c][d] * P.inv_tau[c];
    }
    u[d] = div_by<EXACT>(wv, total, inv_total);
  }
}

// ---- K2 -------------------------------------------------------------------------------------
// Store one component's post-collision populations of site (plane zg, sxy) and, when the slab is
// its own z neighbour, fill next step's ghost planes with the populations that will cross the
// periodic boundary (only c_z = +1 upward, c_z = -1 downward).
template <int POL = POL_DEFAULT>
__device__ __forceinline__ void store19(const Geom &g, double *__restrict__ f_new, int zg, int c, int sxy,
                                        const double (&f)[Q]) {
  double *__restrict__ out = f_plane(f_new, g, zg, c) + sxy;
  const uint64_t pol = make_policy<POL>();
#pragma unroll
  for (int q = 0; q < Q; q++) {
    if (POL == POL_DEFAULT) store_pop<POL>(out + qslot(q) * g.plane, f[q], pol);
    else store_pop<POL>(ptr_at(out, qslot(q) * g.plane), f[q], pol);
  }
  if (g.periodic_z_local) {
    if (zg == g.nz) {
      double *__restrict__ gh = f_plane(f_new, g, 0, c) + sxy;
#pragma unroll
      for (int q = 0; q < Q; q++) if (cz(q) == 1) gh[qslot(q) * g.plane] = f[q];
    }
    if (zg == 1) {
      double *__restrict__ gh = f_plane(f_new, g, g.nz + 1, c) + sxy;
#pragma unroll
      for (int q = 0; q < Q; q++) if (cz(q) == -1) gh[qslot(q) * g.plane] = f[q];
    }
  }
}

// One site of K2 in two halves, so that a caller can put a wait between the gather of the populations (which
// depend on nothing written in this launch) and the rest (step_wave.cuh waits for psi there):
//   fused_gather: f[c][q] <- streamed populations; returns false when there is nothing to do for this site
//   fused_finish: moments, Shan-Chen force from psi, collision, store
template <int C, bool MASK, int POL = POL_DEFAULT>
__device__ __forceinline__ bool fused_gather(const Geom &g, const Fields &fl, int x, int y, int zg, Nbr &n, uint32_t &bits,
                                             double (&f)[C][Q]) {
  n = make_nbr(g, x, y);
  const uint8_t *ms = mask_plane(fl.mask, g, zg) + n.sxy;
  bits = site_bits<MASK, POL>(g, fl, zg, n);
  // Solid sites: advection copies them (advection.d:83-85) and they never collide; both population buffers hold
  // their (static) values since upload, so there is nothing to do -- except that the fluid sites of a warp then
  // write 32-byte sectors only in part, which the L2 has to fill from DRAM first (porous workload, 30 % solid:
  // 23.1 GB read per 33.5 M sites against 12.1 GB without a mask, profiles/r02/r02d_porous_*).  CARRY: the
  // solid lanes go through the same instructions as their fluid neighbours -- they gather their OWN populations
  // (an address select in pull19*), collide with omega = 0 (collide_site KEEP: f comes back bit for bit) and
  // store them, so every sector is written whole: porous 512^3 4.80 -> 5.59 GLUPS (single launch: reads 23.1 ->
  // 13.4 GB, r02e_porous_*), 4.30 -> 4.71 GLUPS (two kernels).  (A separate copy branch for
  // the solid lanes -- the warp runs both paths one after the other -- measured 10 % SLOWER than not writing.)
  constexpr bool CARRY = MASK && DLBC_SOLID_COPY;
  if (MASK && (bits & 1u) && !CARRY) return false;
  const double *ps = f_plane(fl.f_old, g, zg, 0) + n.sxy;
#pragma unroll
  for (int c = 0; c < C; c++) pull_site<MASK, POL, CARRY>(ps + c * g.fc, ms, bits, g, n, f[c]);
  return true;
}

template <int C, int EQ, bool SC, bool MASK, bool EXACT, bool COHERENT = false, int POL = POL_DEFAULT>
__device__ __forceinline__ void fused_finish(const Geom &g, const Fields &fl, const StepParams &P, int zg, const Nbr &n,
                                             uint32_t bits, double (&f)[C][Q]) {
  constexpr bool CARRY = MASK && DLBC_SOLID_COPY;
  const bool solid = MASK && (bits & 1u);
  double rho[C], m[C][3], F[C][3];
#pragma unroll
  for (int c = 0; c < C; c++) {
    rho[c] = density19(f[c]);
    momentum19(f[c], m[c]);
    F[c][0] = 0.0; F[c][1] = 0.0; F[c][2] = 0.0;
  }
  if (SC) {
    double psi1[C];
#pragma unroll
    for (int c = 0; c < C; c++) {
      const double *a = psi_plane(fl.psi, g, zg, c) + n.sxy;
      psi1[c] = COHERENT ? ld_global_cg(a) : *a;
    }
    shan_chen_force<C, MASK, EXACT, COHERENT>(g, fl, P, zg, n, bits, psi1, F);
    if (MASK && P.any_gwc) wall_force<C, POL != POL_DEFAULT>(g, fl, P, zg, n, bits, psi1, F);
  }
  double u[3] = {0.0, 0.0, 0.0};
  if (EQ == 2) common_velocity<C, EXACT>(P, m, rho, u);
#pragma unroll
  for (int c = 0; c < C; c++) {
    collide_site<EQ, EXACT, CARRY>(f[c], rho[c], m[c], F[c], u, P, c, solid);
    store19<POL>(g, fl.f_new, zg, c, n.sxy, f[c]);
  }
}

template <int C, int EQ, bool SC, bool MASK, bool EXACT, bool COHERENT = false, int POL = POL_DEFAULT>
__device__ __forceinline__ void fused_site(const Geom &g, const Fields &fl, const StepParams &P, int x, int y, int zg) {
  Nbr n;
  uint32_t bits;
  double f[C][Q];
  if (fused_gather<C, MASK, POL>(g, fl, x, y, zg, n, bits, f))
    fused_finish<C, EQ, SC, MASK, EXACT, COHERENT, POL>(g, fl, P, zg, n, bits, f);
}

template <int C, int EQ, bool SC, bool MASK, bool EXACT>
__global__ void __launch_bounds__(128, (MASK && SC) ? DLBC_MASK_BLOCKS : 4) fused_kernel(Geom g, Fields fl, StepParams P, int z_begin) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= g.nx) return;
  fused_site<C, EQ, SC, MASK, EXACT>(g, fl, P, x, blockIdx.y, z_begin + blockIdx.z);
}

// ---- L.force[comp] for output (lb/io.d:86-90) -----------------------------------------------
// Not on the per-step path: the step kernels keep the force in registers.  Recomputed from the
// psi field of the last step; zero on solid sites (resetForce, force.d:165-170, and the
// isCollidable test at force.d:258).  out: SoA vector field [3][sites_g].
template <int C, bool MASK>
__global__ void __launch_bounds__(128) force_kernel(Geom g, Fields fl, StepParams P, int comp,
                                                    double *__restrict__ out, int z_begin) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= g.nx) return;
  const int y = blockIdx.y;
  const int zg = z_begin + blockIdx.z;
  const Nbr n = make_nbr(g, x, y);
  double F[C][3];
#pragma unroll
  for (int c = 0; c < C; c++) { F[c][0] = 0.0; F[c][1] = 0.0; F[c][2] = 0.0; }
  const uint32_t bits = site_bits<MASK, POL_DEFAULT>(g, fl, zg, n);
  if (!(MASK && (bits & 1u))) {
    double psi1[C];
#pragma unroll
    for (int c = 0; c < C; c++) psi1[c] = psi_plane(fl.psi, g, zg, c)[n.sxy];
    shan_chen_force<C, MASK, true>(g, fl, P, zg, n, bits, psi1, F);
    if (MASK && P.any_gwc) wall_force<C>(g, fl, P, zg, n, bits, psi1, F);
  }
  const int64_t s = (int64_t)zg * g.plane + n.sxy;
#pragma unroll
  for (int c = 0; c < C; c++) {
    if (c == comp) {
#pragma unroll
      for (int d = 0; d < 3; d++) out[(int64_t)d * g.sites_g + s] = F[c][d];
    }
  }
}

}  // namespace dlbc

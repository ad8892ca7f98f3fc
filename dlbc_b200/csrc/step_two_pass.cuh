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
// DRAM traffic per lattice-site update, C components: K1 reads 152 C, writes 8 C;
// K2 reads 152 C + 8 C (+1 mask byte), writes 152 C  ==> 472 C + 1 bytes (945 B at C = 2).
#pragma once
#include "lattice.cuh"

namespace dlbc {

// ---- K1 -------------------------------------------------------------------------------------
template <int C, int PSI, bool MASK>
__global__ void __launch_bounds__(128) psi_kernel(Geom g, Fields fl, int z_begin) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= g.nx) return;
  const int y = blockIdx.y;
  const int zg = z_begin + blockIdx.z;  // ghost-inclusive plane index, 1..nz
  const Nbr n = make_nbr(g, x, y, zg);
  const int64_t s = n.at(0, 0, 0);
  if (MASK && fl.mask[s]) return;  // psi of solid sites is never read (force.d:273 uses the own-site psi)
#pragma unroll
  for (int c = 0; c < C; c++) {
    double f[Q];
    pull19<MASK>(fl.f_old + (int64_t)c * Q * g.sites_g, g.sites_g, fl.mask, n, f);
    const double v = psi_of<PSI>(density19(f));
    double *p = fl.psi + (int64_t)c * g.sites_g;
    p[s] = v;
    if (g.periodic_z_local) {
      if (zg == g.nz) p[s - (int64_t)g.nz * g.plane] = v;  // -> ghost plane 0
      if (zg == 1) p[s + (int64_t)g.nz * g.plane] = v;     // -> ghost plane nz+1
    }
  }
}

// ---- Shan-Chen force for all components of one site (force.d:238-309) -----------------------
// Accumulation order per component f1 is the reference's: f2 outer, q = 1..18 inner, then the
// wall term q = 0..18; prefactor = ((w_q * psi1) * psi2) * cc.
template <int C, bool MASK>
__device__ __forceinline__ void shan_chen_force(const Geom &g, const Fields &fl, const StepParams &P,
                                                const Nbr &n, const double (&psi1)[C], double (&F)[C][3]) {
  const int64_t s = n.at(0, 0, 0);
#pragma unroll
  for (int f2 = 0; f2 < C; f2++) {
    bool any = false;
#pragma unroll
    for (int f1 = 0; f1 < C; f1++) any |= (P.gcc[f1 * MAXC + f2] != 0.0);
    if (!any) continue;
    const double *__restrict__ psi2p = fl.psi + (int64_t)f2 * g.sites_g;
#pragma unroll
    for (int q = 1; q < Q; q++) {
      int64_t nb = n.at(cx(q), cy(q), cz(q));
      if (MASK) { if (fl.mask[nb]) nb = s; }
      const double psi2 = (MASK && nb == s) ? psi1[f2] : __ldg(psi2p + nb);
#pragma unroll
      for (int f1 = 0; f1 < C; f1++) {
        const double cc = P.gcc[f1 * MAXC + f2];
        if (cc == 0.0) continue;
        const double pref = weight(q) * psi1[f1] * psi2 * cc;
        if (cx(q) == 1) F[f1][0] -= pref; else if (cx(q) == -1) F[f1][0] += pref;
        if (cy(q) == 1) F[f1][1] -= pref; else if (cy(q) == -1) F[f1][1] += pref;
        if (cz(q) == 1) F[f1][2] -= pref; else if (cz(q) == -1) F[f1][2] += pref;
      }
    }
  }
}

// NOTE on ordering: the reference finishes ALL f2 for a given f1 (then the wall term of f1)
// before moving to the next f1; since F[f1] only ever receives f1's own terms, interleaving the
// f1 loops as above leaves each F[f1]'s operation sequence unchanged.

template <int C>
__device__ __forceinline__ void wall_force(const Geom &g, const Fields &fl, const StepParams &P,
                                           const Nbr &n, const double (&psi1)[C], double (&F)[C][3]) {
#pragma unroll
  for (int f1 = 0; f1 < C; f1++) {
    const double wc = P.gwc[f1];
    if (wc == 0.0) continue;
    const double *__restrict__ rw = fl.rho_wall + (int64_t)f1 * g.sites_g;
#pragma unroll
    for (int q = 0; q < Q; q++) {
      const int64_t nb = n.at(cx(q), cy(q), cz(q));
      if (fl.mask[nb]) {
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
template <int C, int EQ, bool SC, bool MASK, bool EXACT>
__global__ void __launch_bounds__(128, (MASK && SC) ? 3 : 4) fused_kernel(Geom g, Fields fl, StepParams P, int z_begin) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= g.nx) return;
  const int y = blockIdx.y;
  const int zg = z_begin + blockIdx.z;
  const Nbr n = make_nbr(g, x, y, zg);
  const int64_t s = n.at(0, 0, 0);
  // Solid sites: advection copies them (advection.d:83-85) and they never collide; both
  // population buffers hold their (static) values since upload, so there is nothing to do.
  if (MASK && fl.mask[s]) return;

  double f[C][Q], rho[C], m[C][3], F[C][3];
#pragma unroll
  for (int c = 0; c < C; c++) {
    pull19<MASK>(fl.f_old + (int64_t)c * Q * g.sites_g, g.sites_g, fl.mask, n, f[c]);
    rho[c] = density19(f[c]);
    momentum19(f[c], m[c]);
    F[c][0] = 0.0; F[c][1] = 0.0; F[c][2] = 0.0;
  }
  if (SC) {
    double psi1[C];
#pragma unroll
    for (int c = 0; c < C; c++) psi1[c] = fl.psi[(int64_t)c * g.sites_g + s];
    shan_chen_force<C, MASK>(g, fl, P, n, psi1, F);
    if (MASK && P.any_gwc) wall_force<C>(g, fl, P, n, psi1, F);
  }
  double u[3] = {0.0, 0.0, 0.0};
  if (EQ == 2) common_velocity<C, EXACT>(P, m, rho, u);
#pragma unroll
  for (int c = 0; c < C; c++) {
    collide_site<EQ, EXACT>(f[c], rho[c], m[c], F[c], u, P, c);
    double *__restrict__ out = fl.f_new + (int64_t)c * Q * g.sites_g;
#pragma unroll
    for (int q = 0; q < Q; q++) {
      out[(int64_t)q * g.sites_g + s] = f[c][q];
      // the slab is its own z neighbour: fill next step's ghost planes with the populations that
      // will cross the periodic boundary (only c_z = +1 upward, c_z = -1 downward)
      if (g.periodic_z_local) {
        if (cz(q) == 1 && zg == g.nz) out[(int64_t)q * g.sites_g + s - (int64_t)g.nz * g.plane] = f[c][q];
        if (cz(q) == -1 && zg == 1) out[(int64_t)q * g.sites_g + s + (int64_t)g.nz * g.plane] = f[c][q];
      }
    }
  }
}

// ---- L.force[comp] for output (lb/io.d:86-90) -----------------------------------------------
// Not on the per-step path: the step kernels keep the force in registers.  Recomputed from the
// psi field of the last step; zero on solid sites (resetForce, force.d:165-170, and the
// isCollidable test at force.d:258).
template <int C, bool MASK>
__global__ void __launch_bounds__(128) force_kernel(Geom g, Fields fl, StepParams P, int comp,
                                                    double *__restrict__ out, int z_begin) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= g.nx) return;
  const int y = blockIdx.y;
  const int zg = z_begin + blockIdx.z;
  const Nbr n = make_nbr(g, x, y, zg);
  const int64_t s = n.at(0, 0, 0);
  double F[C][3];
#pragma unroll
  for (int c = 0; c < C; c++) { F[c][0] = 0.0; F[c][1] = 0.0; F[c][2] = 0.0; }
  if (!(MASK && fl.mask[s])) {
    double psi1[C];
#pragma unroll
    for (int c = 0; c < C; c++) psi1[c] = fl.psi[(int64_t)c * g.sites_g + s];
    shan_chen_force<C, MASK>(g, fl, P, n, psi1, F);
    if (MASK && P.any_gwc) wall_force<C>(g, fl, P, n, psi1, F);
  }
#pragma unroll
  for (int c = 0; c < C; c++) {
    if (c == comp) {
#pragma unroll
      for (int d = 0; d < 3; d++) out[(int64_t)d * g.sites_g + s] = F[c][d];
    }
  }
}

}  // namespace dlbc

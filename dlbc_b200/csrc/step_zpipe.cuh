// step_zpipe.cuh -- kernel variant DLBC_KERNEL_ZPIPE: the single-pass z-marching timestep of
// step_zmarch.cuh with the population loads turned into an asynchronous shared-memory pipeline.
//
// Why: the z-march kernel already moves only ~630 B per site update (ncu, profiles/), but runs at
// ~57 % of the HBM roofline because one 512-thread block per SM alternates between a load phase and a
// compute phase separated by __syncthreads: while the block computes, the SM has no loads in
// flight.  Here every thread issues the 19 C pulled populations of plane z+2 with per-thread
// cp.async (LDGSTS, 8 bytes each, periodic wrap and bounce-back folded into the source address
// exactly as advection.d:61-89 prescribes) BEFORE it does the arithmetic of plane z, so HBM works
// through the whole iteration:
//
//   registers  : populations of plane z (the plane being collided)
//   stage[z+1] : populations of plane z+1 (landed; psi(z+1) is computed from them, then they move
//                to the registers at the end of the iteration)
//   stage[z]   : free once plane z sits in registers -> landing zone of plane z+2
//   rim stage  : pulled populations of the tile's one-site rim (psi of the neighbours' sites)
//   psi ring   : 4 slots of (TX+2) x (TY+2) psi values per component
//
// All staged values are thread-private (a thread only reads what it copied itself), so the only
// block-level synchronisation is the one __syncthreads per plane that publishes the psi ring.
// Multi-rank slabs: psi of the two ghost planes (0 and nz+1) is not computable locally (it needs
// the neighbour's plane nz-1 / 2); it is read from the global psi array, into which the neighbour's
// boundary psi planes have been received (dlbc_cuda.cu: step_once).
//
// Shared memory (TX x TY = 32 x 8, C = 2): 2 x 77.8 KB stages + 25.5 KB rim stage + 21.8 KB ring
// = 203 KB, one 256-thread block per SM.  DRAM traffic as step_zmarch.cuh: 304 C + 1 bytes per site
// update plus the rim share that misses L2.  Results are bit-identical to the other variants: the
// same per-site functions in the same order.
#pragma once
#include "lattice.cuh"
#include "step_two_pass.cuh"
#include "step_zmarch.cuh"

namespace dlbc {

__device__ __forceinline__ void cp_async8(double *smem_dst, const double *gmem_src) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }

// Issue the asynchronous pull (advection.d:61-89) of all C components of site n in plane w into
// dst[(c * Q + q) * pitch].  Solid sites are skipped: their staged values are never read.
template <int C, bool MASK>
__device__ __forceinline__ void pull_async(const Geom &g, const Fields &fl, const Nbr &n, int w, double *dst, int pitch) {
  const uint8_t *ms = mask_plane(fl.mask, g, w) + n.sxy;
  if (MASK && ms[0]) return;
  const double *ps = f_plane(fl.f_old, g, w, 0) + n.sxy;
#pragma unroll
  for (int q = 0; q < Q; q++) {
    const int d = n.ox[1 - cx(q)] + n.oy[1 - cy(q)];
    int src = q * g.plane - cz(q) * g.fz + d;
    if (MASK && q > 0) {
      if (ms[d - cz(q) * g.plane]) src = bounce(q) * g.plane;
    }
#pragma unroll
    for (int c = 0; c < C; c++) cp_async8(dst + (c * Q + q) * pitch, ps + c * g.fc + src);
  }
}

template <int C, int PSI>
__device__ __forceinline__ void psi_from_stage(const double *src, int pitch, double (&psi)[C]) {
#pragma unroll
  for (int c = 0; c < C; c++) {
    double f[Q];
#pragma unroll
    for (int q = 0; q < Q; q++) f[q] = src[(c * Q + q) * pitch];
    psi[c] = psi_of<PSI>(density19(f));
  }
}

template <int C, int EQ, int PSI, bool MASK, bool EXACT, int TX, int TY>
__global__ void __launch_bounds__(TX *TY, 1) zpipe_kernel(Geom g, Fields fl, StepParams P, ZMarch zm) {
  constexpr int NT = TX * TY;
  constexpr int RW = TX + 2;
  constexpr int RP = RW * (TY + 2);
  constexpr int NRIM = 2 * RW + 2 * TY;
  static_assert(NRIM <= NT, "one rim cell per thread");
  extern __shared__ double smem[];
  double *stage = smem;                          // [2][C*Q][NT]
  double *rimst = smem + 2 * C * Q * NT;         // [C*Q][NRIM]
  double *ring = rimst + C * Q * NRIM;           // [4][C][RP]

  const int tid = threadIdx.y * TX + threadIdx.x;
  const int x0 = blockIdx.x * TX, y0 = blockIdx.y * TY;
  const int x = x0 + threadIdx.x, y = y0 + threadIdx.y;
  const bool inside = x < g.nx && y < g.ny;
  const Nbr n = make_nbr(g, wrap_coord(x, g.nx), wrap_coord(y, g.ny));
  const int cell = (threadIdx.y + 1) * RW + threadIdx.x + 1;

  // this thread's rim cell (threads 0..NRIM-1)
  const bool has_rim = tid < NRIM;
  int rcell = 0;
  Nbr rn = n;
  if (has_rim) {
    int lx, ly;
    if (tid < RW) { lx = tid - 1; ly = -1; }
    else if (tid < 2 * RW) { lx = tid - RW - 1; ly = TY; }
    else { const int k = tid - 2 * RW; ly = k >> 1; lx = (k & 1) ? TX : -1; }
    rcell = (ly + 1) * RW + lx + 1;
    rn = make_nbr(g, wrap_coord(x0 + lx, g.nx), wrap_coord(y0 + ly, g.ny));
  }

  const int zs = zm.z_begin + blockIdx.z * zm.z_chunk;
  const int ze = min(zs + zm.z_chunk, zm.z_end);
  if (zs >= ze) return;

  // planes whose psi this rank can compute from its own populations; the others (ghost planes of a
  // multi-rank slab) are read from the global psi array
  auto local_plane = [&](int zg) { return g.periodic_z_local || (zg >= 1 && zg <= g.nz); };
  auto stage_of = [&](int zg) { return stage + (zg & 1) * (C * Q * NT) + tid; };

  auto issue_main = [&](int zg) {
    if (local_plane(zg)) pull_async<C, MASK>(g, fl, n, wrap_plane(g, zg), stage_of(zg), NT);
  };
  auto issue_rim = [&](int zg) {
    if (has_rim && local_plane(zg)) pull_async<C, MASK>(g, fl, rn, wrap_plane(g, zg), rimst + tid, NRIM);
  };
  // psi(zg) of the own cell and the rim cell -> ring slot zg & 3 (staged populations have landed)
  auto publish_psi = [&](int zg) {
    double *slot = ring + (zg & 3) * C * RP;
    double v[C];
    if (local_plane(zg)) {
      const int w = wrap_plane(g, zg);
      const uint8_t *mp = mask_plane(fl.mask, g, w);
      if (MASK && mp[n.sxy]) {
#pragma unroll
        for (int c = 0; c < C; c++) v[c] = 0.0;  // never read (force.d:273 substitutes the reader's own psi)
      } else {
        psi_from_stage<C, PSI>(stage_of(zg), NT, v);
      }
#pragma unroll
      for (int c = 0; c < C; c++) slot[c * RP + cell] = v[c];
      if (has_rim) {
        if (MASK && mp[rn.sxy]) {
#pragma unroll
          for (int c = 0; c < C; c++) v[c] = 0.0;
        } else {
          psi_from_stage<C, PSI>(rimst + tid, NRIM, v);
        }
#pragma unroll
        for (int c = 0; c < C; c++) slot[c * RP + rcell] = v[c];
      }
    } else {
#pragma unroll
      for (int c = 0; c < C; c++) {
        const double *pg = psi_plane(fl.psi, g, zg, c);
        slot[c * RP + cell] = pg[n.sxy];
        if (has_rim) slot[c * RP + rcell] = pg[rn.sxy];
      }
    }
  };

  double f[C][Q];

  // ---- prologue: psi(zs-1), psi(zs) into the ring; plane zs into the registers; plane zs+1 in flight
  issue_main(zs - 1);
  issue_rim(zs - 1);
  cp_async_wait_all();
  publish_psi(zs - 1);
  issue_main(zs);
  issue_rim(zs);
  cp_async_wait_all();
  publish_psi(zs);
  {
    const double *s = stage_of(zs);
#pragma unroll
    for (int c = 0; c < C; c++)
#pragma unroll
      for (int q = 0; q < Q; q++) f[c][q] = s[(c * Q + q) * NT];
  }
  issue_main(zs + 1);
  issue_rim(zs + 1);

  for (int zg = zs; zg < ze; zg++) {
    cp_async_wait_all();                 // plane zg+1 (own site and rim) has landed
    if (zg + 2 <= ze) issue_main(zg + 2);  // stage[zg & 1] is free: plane zg lives in the registers
    publish_psi(zg + 1);
    if (zg + 2 <= ze) issue_rim(zg + 2);   // the rim stage was consumed just above
    __syncthreads();                     // ring slots zg-1, zg, zg+1 complete (4 slots: no WAR hazard)

    const uint8_t *mp = mask_plane(fl.mask, g, zg);
    if (inside && !(MASK && mp[n.sxy])) {
      double rho[C], m[C][3], F[C][3], psi1[C];
      const double *r0 = ring + (zg & 3) * C * RP;
#pragma unroll
      for (int c = 0; c < C; c++) {
        rho[c] = density19(f[c]);
        momentum19(f[c], m[c]);
        F[c][0] = 0.0; F[c][1] = 0.0; F[c][2] = 0.0;
        psi1[c] = r0[c * RP + cell];
      }
      sc_force<C, EXACT>(P, psi1, F, [&](int f2, int q) -> double {
        if (MASK) { if (mp[n.at(cx(q), cy(q)) + cz(q) * g.plane]) return psi1[f2]; }
        return ring[(((zg + cz(q)) & 3) * C + f2) * RP + cell + cy(q) * RW + cx(q)];
      });
      if (MASK && P.any_gwc) wall_force<C>(g, fl, P, zg, n, 0u, psi1, F);
      double u[3] = {0.0, 0.0, 0.0};
      if (EQ == 2) common_velocity<C, EXACT>(P, m, rho, u);
#pragma unroll
      for (int c = 0; c < C; c++) {
        collide_site<EQ, EXACT>(f[c], rho[c], m[c], F[c], u, P, c);
        store19(g, fl.f_new, zg, c, n.sxy, f[c]);
      }
    }
    if (zg + 1 < ze) {
      const double *s = stage_of(zg + 1);
#pragma unroll
      for (int c = 0; c < C; c++)
#pragma unroll
        for (int q = 0; q < Q; q++) f[c][q] = s[(c * Q + q) * NT];
    }
  }
}

}  // namespace dlbc

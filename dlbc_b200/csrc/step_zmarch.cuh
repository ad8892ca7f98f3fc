// step_zmarch.cuh -- kernel variant DLBC_KERNEL_ZMARCH: the whole Shan-Chen timestep in ONE pass
// over the populations (hard part H1, option c).
//
// Why: psi at the 18 neighbours of a site depends on the POST-advection density there, i.e. on
// populations two sites away, so the two-pass design reads every population twice
// (472 C + 1 = 945 B per site update at C = 2).  Here a thread block owns a TX x TY column of
// sites and marches along z.  Per plane step it
//   A. pulls the 19 C populations of its sites at plane z+1 (advection.d:61-89), sums them
//      (density.d:37-39), applies psi (force.d:320-332) and stores psi(z+1) in a 4-slot ring of
//      (TX+2) x (TY+2) psi planes in shared memory; the ring's one-site rim (the neighbours'
//      sites) is computed the same way by the first 2(TX+TY)+4 threads -- the only redundant
//      loads: (TX+2)(TY+2)/(TX TY) of one plane, mostly L2 hits because the neighbouring blocks
//      pull the same lines at the same time;
//   B. exchanges the freshly pulled populations (plane z+1) with the ones it parked in a
//      thread-private shared-memory stash one step earlier (plane z), so the register file only
//      ever holds ONE plane's 19 C values;
//   C. after one __syncthreads computes the Shan-Chen force of plane z from ring slots z-1, z,
//      z+1 (force.d:238-309), the BDist2 common velocity (collision.d:156-196), collides
//      (collision.d:90-119) and writes the populations once.
// DRAM traffic per site update: 152 C read + 152 C written + 1 mask byte + the rim share
// ==> ~609 B at C = 2 plus (rim misses).  Results are bit-identical to the two-pass variant: the
// same per-site functions in the same order.
//
// Shared memory: stash C*19*TX*TY doubles + ring 4*C*(TX+2)*(TY+2) doubles
// (64 x 8, C = 2: 152 KB + 41 KB; 64 x 4: 76 KB + 25 KB, two blocks per SM).
#pragma once
#include "lattice.cuh"
#include "step_two_pass.cuh"

namespace dlbc {

struct ZMarch {
  int z_begin;    // first plane this launch updates (ghost-inclusive index, >= 1)
  int z_end;      // one past the last plane
  int z_chunk;    // planes per block
};

__device__ __forceinline__ int wrap_plane(const Geom &g, int zg) {
  // in-process periodicity: psi of ghost plane 0 is psi of plane nz, of ghost nz+1 that of plane 1
  if (!g.periodic_z_local) return zg;
  return zg < 1 ? zg + g.nz : (zg > g.nz ? zg - g.nz : zg);
}

__device__ __forceinline__ int wrap_coord(int i, int n) {
  i %= n;
  return i < 0 ? i + n : i;
}

// psi of all components at site (x, y, plane w) from freshly pulled populations; returns them in f
template <int C, int PSI, bool MASK>
__device__ __forceinline__ bool pull_site(const Geom &g, const Fields &fl, int x, int y, int w, double (&f)[C][Q],
                                          double (&psi)[C]) {
  const Nbr n = make_nbr(g, x, y);
  const uint8_t *ms = mask_plane(fl.mask, g, w) + n.sxy;
  if (MASK && ms[0]) {
#pragma unroll
    for (int c = 0; c < C; c++) psi[c] = 0.0;  // never read: bounced links use the reader's own psi (force.d:273)
    return false;
  }
#pragma unroll
  for (int c = 0; c < C; c++) pull19<MASK>(f_plane(fl.f_old, g, w, c) + n.sxy, ms, g, n, f[c]);
#pragma unroll
  for (int c = 0; c < C; c++) psi[c] = psi_of<PSI>(density19(f[c]));
  return true;
}

template <int C, int EQ, int PSI, bool MASK, bool EXACT, int TX, int TY, int MINB>
__global__ void __launch_bounds__(TX *TY, MINB) zmarch_kernel(Geom g, Fields fl, StepParams P, ZMarch zm) {
  constexpr int NT = TX * TY;
  constexpr int RW = TX + 2;                 // ring row pitch
  constexpr int RP = RW * (TY + 2);          // cells per ring plane
  constexpr int NRIM = 2 * RW + 2 * TY;      // rim cells
  extern __shared__ double smem[];
  double *stash = smem;                      // [C][Q][NT]
  double *ring = smem + C * Q * NT;          // [4][C][RP]

  const int tid = threadIdx.y * TX + threadIdx.x;
  const int x0 = blockIdx.x * TX, y0 = blockIdx.y * TY;
  const int x = x0 + threadIdx.x, y = y0 + threadIdx.y;
  const bool inside = x < g.nx && y < g.ny;
  // sites past a ragged edge still fill their ring cell with the periodic image: cell (lx, ly) always
  // holds psi of site ((x0+lx) mod nx, (y0+ly) mod ny), which is what the last valid site's +1
  // neighbour is
  const int xs = wrap_coord(x, g.nx), ys = wrap_coord(y, g.ny);
  const int cell = (threadIdx.y + 1) * RW + threadIdx.x + 1;

  const int zs = zm.z_begin + blockIdx.z * zm.z_chunk;
  const int ze = min(zs + zm.z_chunk, zm.z_end);
  if (zs >= ze) return;

  double f[C][Q];
  double psi[C];

  // psi of one ring plane's rim cells (strided over the block)
  auto rim = [&](int zg) {
    const int w = wrap_plane(g, zg);
    double *slot = ring + (zg & 3) * C * RP;
    for (int h = tid; h < NRIM; h += NT) {
      int lx, ly;
      if (h < RW) { lx = h - 1; ly = -1; }
      else if (h < 2 * RW) { lx = h - RW - 1; ly = TY; }
      else { const int k = h - 2 * RW; ly = k >> 1; lx = (k & 1) ? TX : -1; }
      double fr[C][Q], pr[C];
      pull_site<C, PSI, MASK>(g, fl, wrap_coord(x0 + lx, g.nx), wrap_coord(y0 + ly, g.ny), w, fr, pr);
#pragma unroll
      for (int c = 0; c < C; c++) slot[c * RP + (ly + 1) * RW + lx + 1] = pr[c];
    }
  };

  // ---- prime: psi(zs-1) into the ring; populations of plane zs into the stash, psi(zs) into the ring
  {
    pull_site<C, PSI, MASK>(g, fl, xs, ys, wrap_plane(g, zs - 1), f, psi);
    double *slot = ring + ((zs - 1) & 3) * C * RP;
#pragma unroll
    for (int c = 0; c < C; c++) slot[c * RP + cell] = psi[c];
  }
  rim(zs - 1);
  rim(zs);
  {
    pull_site<C, PSI, MASK>(g, fl, xs, ys, wrap_plane(g, zs), f, psi);
    double *slot = ring + (zs & 3) * C * RP;
#pragma unroll
    for (int c = 0; c < C; c++) {
      slot[c * RP + cell] = psi[c];
#pragma unroll
      for (int q = 0; q < Q; q++) stash[(c * Q + q) * NT + tid] = f[c][q];
    }
  }

  for (int zg = zs; zg < ze; zg++) {
    // ---- A: plane zg+1 -> psi ring (rim first: f is dead here, so the rim's 19 C temporaries do
    // not add to the register peak); B: swap with the stash (registers now hold plane zg)
    rim(zg + 1);
    pull_site<C, PSI, MASK>(g, fl, xs, ys, wrap_plane(g, zg + 1), f, psi);
    {
      double *slot = ring + ((zg + 1) & 3) * C * RP;
#pragma unroll
      for (int c = 0; c < C; c++) {
        slot[c * RP + cell] = psi[c];
#pragma unroll
        for (int q = 0; q < Q; q++) {
          double *p = stash + (c * Q + q) * NT + tid;
          const double t = *p;
          *p = f[c][q];
          f[c][q] = t;
        }
      }
    }
    __syncthreads();  // ring slots zg-1, zg, zg+1 complete (4 slots: the next A writes slot zg+2)

    // ---- C: force + collide plane zg, write once
    if (!inside) continue;
    const Nbr n = make_nbr(g, xs, ys);
    const uint8_t *mp = mask_plane(fl.mask, g, zg);
    if (MASK && mp[n.sxy]) continue;  // solid sites keep their populations in both buffers
    double rho[C], m[C][3], F[C][3], psi1[C];
    const double *r0 = ring + (zg & 3) * C * RP;
#pragma unroll
    for (int c = 0; c < C; c++) {
      rho[c] = density19(f[c]);
      momentum19(f[c], m[c]);
      F[c][0] = 0.0; F[c][1] = 0.0; F[c][2] = 0.0;
      psi1[c] = r0[c * RP + cell];
    }
    // addShanChenForcePsi (force.d:238-309): neighbours' psi from the shared-memory ring
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
}

}  // namespace dlbc

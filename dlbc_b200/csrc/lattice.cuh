// lattice.cuh -- device-side description of one rank's z-slab and indexing helpers.
#pragma once
#include <cstdint>
#include "lb_site.cuh"

namespace dlbc {

// Device layout (DESIGN.md "Data layout in HBM"):
//   populations  f[c][q][zg][y][x]   fp64, x fastest, zg in [0, nz+2): planes 0 and nz+1 are ghosts
//   psi          psi[c][zg][y][x]    fp64, same ghosts
//   mask         m[zg][y][x]         uint8 (0 = None, 1 = Solid), same ghosts
// x and y are periodic inside the kernels (the slab owns all of x and y); z neighbours come from
// the ghost planes, which are filled by the step kernels themselves when the rank is its own
// neighbour (nranks == 1) and by NCCL send/recv otherwise.
struct Geom {
  int nx, ny, nz;        // local physical extents (nz = this slab's planes)
  int periodic_z_local;  // nranks == 1: the slab is its own z neighbour
  int64_t plane;         // nx * ny
  int64_t sites_g;       // (nz + 2) * plane  == stride between consecutive (c, q) arrays
};

struct Fields {
  const double *f_old;   // read
  double *f_new;         // written
  double *psi;           // [C][sites_g]
  const double *rho_wall;// [C][sites_g] static own-site densities (only when gwc != 0 and solids exist)
  const uint8_t *mask;   // [sites_g]
};

__device__ __forceinline__ int wrap_inc(int i, int n) { return i + 1 == n ? 0 : i + 1; }
__device__ __forceinline__ int wrap_dec(int i, int n) { return i == 0 ? n - 1 : i - 1; }

// Offsets of the 3x3 rows (dy, dz in {-1,0,1}) around site (x, y, zg) and the wrapped x columns.
struct Nbr {
  int64_t row[3][3];  // [dz+1][dy+1] -> index of (x=0, y+dy, zg+dz)
  int xc[3];          // [dx+1] -> wrapped x
  __device__ __forceinline__ int64_t at(int dx, int dy, int dz) const { return row[dz + 1][dy + 1] + xc[dx + 1]; }
};

__device__ __forceinline__ Nbr make_nbr(const Geom &g, int x, int y, int zg) {
  Nbr n;
  const int ys[3] = {wrap_dec(y, g.ny), y, wrap_inc(y, g.ny)};
#pragma unroll
  for (int dz = 0; dz < 3; dz++)
#pragma unroll
    for (int dy = 0; dy < 3; dy++) n.row[dz][dy] = ((int64_t)(zg + dz - 1) * g.ny + ys[dy]) * g.nx;
  n.xc[0] = wrap_dec(x, g.nx);
  n.xc[1] = x;
  n.xc[2] = wrap_inc(x, g.nx);
  return n;
}

// lb/advection.d:61-89: pull streaming with half-way bounce-back for one component at one
// (advectable) site: f[q] = bounce-back(mask[p - c_q]) ? old[p][bounce q] : old[p - c_q][q].
template <bool MASK>
__device__ __forceinline__ void pull19(const double *__restrict__ fc, int64_t qstride,
                                       const uint8_t *__restrict__ mask, const Nbr &n, double (&f)[Q]) {
  const int64_t self = n.at(0, 0, 0);
#pragma unroll
  for (int q = 0; q < Q; q++) {
    int64_t src = n.at(-cx(q), -cy(q), -cz(q));
    int qq = q;
    if (MASK && q > 0) {
      if (mask[src]) { src = self; qq = bounce(q); }
    }
    f[q] = __ldg(fc + (int64_t)qq * qstride + src);
  }
}

}  // namespace dlbc

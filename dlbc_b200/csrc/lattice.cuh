// lattice.cuh -- device-side description of one rank's z-slab and indexing helpers.
#pragma once
#include <cstdint>
#include "lb_site.cuh"

namespace dlbc {

// Device layout (DESIGN.md "Data layout in HBM"), plane-major structure of arrays:
//   populations  f[zg][c][s][y][x]   fp64, x fastest; zg in [0, nz+2): planes 0 and nz+1 are ghosts;
//                                    s = qslot(q), the storage slot of direction q (lb_site.cuh)
//   psi          psi[c][zg][y][x]    fp64, same ghosts
//   mask         m[zg][y][x]         uint8 (0 = None, 1 = Solid), same ghosts
// One z-plane of ALL C*19 population arrays is one contiguous block of C*19*nx*ny doubles, so
//   * everything a site update touches (planes zg-1, zg, zg+1) lies within 3*C*19*nx*ny elements
//     of the plane's base pointer: per-thread addressing is 32-bit offsets from a block-uniform
//     64-bit base (3 integer instructions per access instead of ~10 with per-array 64-bit strides);
//   * a boundary plane of one (c, q) array is still one contiguous run of nx*ny doubles for the
//     NCCL ghost exchange, and a whole ghost plane is one contiguous copy.
// x and y are periodic inside the kernels (the slab owns all of x and y); z neighbours come from
// the ghost planes, which are filled by the step kernels themselves when the rank is its own
// neighbour (nranks == 1) and by NCCL send/recv otherwise.
struct Geom {
  int nx, ny, nz;        // local physical extents (nz = this slab's planes)
  int periodic_z_local;  // nranks == 1: the slab is its own z neighbour
  int plane;             // nx * ny                      == stride between consecutive q arrays
  int fc;                // 19 * plane                   == stride between components within a plane block
  int fz;                // ncomp * 19 * plane           == stride between z planes of the populations
  int ncomp;
  int64_t sites_g;       // (nz + 2) * plane             == stride between the components of psi
};

struct Fields {
  const double *f_old;   // read
  double *f_new;         // written
  double *psi;           // [C][sites_g]
  const double *rho_wall;// [C][sites_g] static own-site densities (only when gwc != 0 and solids exist)
  const uint8_t *mask;   // [sites_g]
  const uint32_t *nb;    // [sites_g] bit q: site p + c_q is Solid (bit 0: p itself); only when solids exist
};

__host__ __device__ __forceinline__ int64_t f_index(const Geom &g, int c, int q, int zg, int64_t sxy) {
  return (int64_t)zg * g.fz + (int64_t)c * g.fc + (int64_t)qslot_rt(q) * g.plane + sxy;
}

__device__ __forceinline__ int wrap_inc(int i, int n) { return i + 1 == n ? 0 : i + 1; }
__device__ __forceinline__ int wrap_dec(int i, int n) { return i == 0 ? n - 1 : i - 1; }

// In-plane neighbourhood of site (x, y): 32-bit element offsets with the periodic wrap folded in.
struct Nbr {
  int sxy;    // y * nx + x
  int ox[3];  // [dx+1]: offset to x+dx  (-1 / +1, or +-(nx-1) across the periodic boundary)
  int oy[3];  // [dy+1]: offset to y+dy  (-nx / +nx, or +-(ny-1)*nx)
  __device__ __forceinline__ int at(int dx, int dy) const { return sxy + ox[dx + 1] + oy[dy + 1]; }
};

__device__ __forceinline__ Nbr make_nbr(const Geom &g, int x, int y) {
  Nbr n;
  n.sxy = y * g.nx + x;
  n.ox[0] = x == 0 ? g.nx - 1 : -1;
  n.ox[1] = 0;
  n.ox[2] = x == g.nx - 1 ? 1 - g.nx : 1;
  n.oy[0] = y == 0 ? (g.ny - 1) * g.nx : -g.nx;
  n.oy[1] = 0;
  n.oy[2] = y == g.ny - 1 ? -(g.ny - 1) * g.nx : g.nx;
  return n;
}

// Block-uniform base pointers of ghost-inclusive plane zg.
__device__ __forceinline__ const double *f_plane(const double *f, const Geom &g, int zg, int c) {
  return f + (int64_t)zg * g.fz + (int64_t)c * g.fc;
}
__device__ __forceinline__ double *f_plane(double *f, const Geom &g, int zg, int c) {
  return f + (int64_t)zg * g.fz + (int64_t)c * g.fc;
}
__device__ __forceinline__ const uint8_t *mask_plane(const uint8_t *m, const Geom &g, int zg) {
  return m + (int64_t)zg * g.plane;
}
__device__ __forceinline__ const double *psi_plane(const double *p, const Geom &g, int zg, int c) {
  return p + (int64_t)c * g.sites_g + (int64_t)zg * g.plane;
}
__device__ __forceinline__ double *psi_plane(double *p, const Geom &g, int zg, int c) {
  return p + (int64_t)c * g.sites_g + (int64_t)zg * g.plane;
}

// lb/advection.d:61-89: pull streaming with half-way bounce-back for one component at one
// (advectable) site: f[q] = bounce-back(mask[p - c_q]) ? old[p][bounce q] : old[p - c_q][q].
// ps = f_plane(f_old, g, zg, c) + n.sxy  (per-thread pointer to the site's q = 0 entry: every
// access is then ps[int offset], one IMAD.WIDE), ms = mask_plane(mask, g, zg) + n.sxy.
// base + off elements as ONE IMAD.WIDE.  Written as PTX because nvcc otherwise folds the per-thread base
// pointer back into (block base + 64-bit site index) and rebuilds every address from a sign-extended
// 64-bit index sum: shift, add, add-with-carry, LEA, LEA.HI.X -- 5-6 integer instructions per load, ~45 %
// of the fused path's instruction stream (profiles/README.md).  All offsets a site update needs lie within
// +-1 plane block of its own entry, hence 32 bits (see Geom).
__device__ __forceinline__ const double *ptr_at(const double *base, int off) {
  const double *r;
  asm("mad.wide.s32 %0, %1, 8, %2;" : "=l"(r) : "r"(off), "l"(base));
  return r;
}
__device__ __forceinline__ double *ptr_at(double *base, int off) {
  double *r;
  asm("mad.wide.s32 %0, %1, 8, %2;" : "=l"(r) : "r"(off), "l"(base));
  return r;
}
__device__ __forceinline__ const uint8_t *ptr_at(const uint8_t *base, int off) {
  const uint8_t *r;
  asm("mad.wide.s32 %0, %1, 1, %2;" : "=l"(r) : "r"(off), "l"(base));
  return r;
}

// Loads through a ptr_at() address: the address-space inference is lost across the asm, so say .global
// explicitly (a generic LD costs an extra address-space check).  ld_global_cg is the L1-bypassing coherent
// load for psi written by other blocks of the SAME launch (step_wave.cuh: always served by L2, where the
// writers' release made it visible), ld_global_nc the read-only path.
__device__ __forceinline__ double ld_global_cg(const double *a) {
  double v;
  asm volatile("ld.global.cg.f64 %0, [%1];" : "=d"(v) : "l"(a) : "memory");
  return v;
}
__device__ __forceinline__ double ld_global_nc(const double *a) {
  double v;
  asm("ld.global.nc.f64 %0, [%1];" : "=d"(v) : "l"(a));
  return v;
}
__device__ __forceinline__ int ld_mask(const uint8_t *a) {
  int v;
  asm("ld.global.nc.u8 %0, [%1];" : "=r"(v) : "l"(a));
  return v;
}

// L2 eviction-priority hints (createpolicy + ld/st .L2::cache_hint).  Used by the wave variant
// (step_wave.cuh), whose psi items read populations that the fused items read again shortly
// afterwards: first touch KEEP (evict_last), second touch and the freshly written populations
// STREAM (evict_first), so that the written lines do not push the waiting ones out of L2.
enum { POL_DEFAULT = 0, POL_KEEP = 1, POL_STREAM = 2 };

template <int POL>
__device__ __forceinline__ uint64_t make_policy() {
  uint64_t p = 0;
  if (POL == POL_KEEP) asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;\n" : "=l"(p));
  if (POL == POL_STREAM) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;\n" : "=l"(p));
  return p;
}
template <int POL>
__device__ __forceinline__ double load_pop(const double *a, uint64_t pol) {
  if (POL == POL_DEFAULT) return ld_global_nc(a);
  double v;
  asm volatile("ld.global.nc.L2::cache_hint.f64 %0, [%1], %2;\n" : "=d"(v) : "l"(a), "l"(pol));
  return v;
}
template <int POL>
__device__ __forceinline__ void store_pop(double *a, double v, uint64_t pol) {
  if (POL == POL_DEFAULT) { asm volatile("st.global.f64 [%0], %1;" ::"l"(a), "d"(v) : "memory"); return; }
  asm volatile("st.global.L2::cache_hint.f64 [%0], %1, %2;\n" ::"l"(a), "d"(v), "l"(pol) : "memory");
}

// lb/advection.d:61-89: pull streaming with half-way bounce-back for one component at one
// (advectable) site: f[q] = bounce-back(mask[p - c_q]) ? old[p][bounce q] : old[p - c_q][q].
// ps = f_plane(f_old, g, zg, c) + n.sxy  (per-thread pointer to the site's q = 0 entry: every
// access is then ps[int offset], one IMAD.WIDE), ms = mask_plane(mask, g, zg) + n.sxy.
// SELF (with MASK): a solid site reads its OWN populations, direction for direction (see pull19_bits).
template <bool MASK, int POL = POL_DEFAULT, bool SELF = false>
__device__ __forceinline__ void pull19(const double *__restrict__ ps, const uint8_t *__restrict__ ms, const Geom &g,
                                       const Nbr &n, double (&f)[Q]) {
  const uint64_t pol = make_policy<POL>();
  if (POL == POL_DEFAULT) {
    // stand-alone kernels: the compiler's own address arithmetic measured faster there
    const bool self = MASK && SELF && ms[0] != 0;
#pragma unroll
    for (int q = 0; q < Q; q++) {
      const int d = n.ox[1 - cx(q)] + n.oy[1 - cy(q)];  // in-plane offset to the upstream site p - c_q
      int src = qslot(q) * g.plane - cz(q) * g.fz + d;
      if (MASK && q > 0) {
        if (ms[d - cz(q) * g.plane]) src = qslot(bounce(q)) * g.plane;
      }
      if (self) src = qslot(q) * g.plane;
      f[q] = load_pop<POL>(ps + src, pol);
    }
  } else {
    // wave variant: explicit 32-bit-offset addressing (ptr_at), -9 % instructions, +2 % throughput
#pragma unroll
    for (int q = 0; q < Q; q++) {
      const int d = n.ox[1 - cx(q)] + n.oy[1 - cy(q)];
      int src = qslot(q) * g.plane - cz(q) * g.fz + d;
      if (MASK && q > 0) {
        if (ld_mask(ptr_at(ms, d - cz(q) * g.plane))) src = qslot(bounce(q)) * g.plane;
      }
      f[q] = load_pop<POL>(ptr_at(ps, src), pol);
    }
  }
}

// The same with the neighbourhood's solid flags in a register (Fields::nb: bit q = site p + c_q is Solid)
// instead of 18 byte loads per component whose results steer the addresses of the population loads:
// one 4-byte load per site serves every component, the streaming step and the force (geometry is static,
// the bits are built once at mask upload).  bounce-back of direction q looks at p - c_q, i.e. bit bounce(q).
// SELF: a solid site (bit 0) reads its OWN populations, direction for direction, so that the caller can carry them
// along to the new buffer (fused_site) -- an address select, not a branch.
template <int POL = POL_DEFAULT, bool SELF = false>
__device__ __forceinline__ void pull19_bits(const double *__restrict__ ps, uint32_t bits, const Geom &g, const Nbr &n,
                                            double (&f)[Q]) {
  const uint64_t pol = make_policy<POL>();
#pragma unroll
  for (int q = 0; q < Q; q++) {
    const int d = n.ox[1 - cx(q)] + n.oy[1 - cy(q)];
    int src = qslot(q) * g.plane - cz(q) * g.fz + d;
    if (q > 0 && ((bits >> bounce(q)) & 1u)) src = qslot(bounce(q)) * g.plane;
    if (SELF && (bits & 1u)) src = qslot(q) * g.plane;
    if (POL == POL_DEFAULT) f[q] = load_pop<POL>(ps + src, pol);
    else f[q] = load_pop<POL>(ptr_at(ps, src), pol);
  }
}

}  // namespace dlbc

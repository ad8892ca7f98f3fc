// field_io.cu -- layout conversion between the reference's host arrays (AoS f[x][y][z][q], z
// fastest, src/dlbc/fields/field.d:114-125) and the device's SoA f[zg][c][q][y][x] (x fastest),
// derived output fields (lb/io.d:56-101) and global reductions (density.d:136-180,
// momentum.d:129-180).  None of this is on the per-step hot path.
#include "field_io.h"

namespace dlbc {

constexpr int TILE = 16;

// ---- populations: dense AoS chunk [xc][ny][nz][19]  <->  SoA [zg][c][q][y][x0 + xi] ---------
// One block per (z-tile, y, x-tile): reads are contiguous runs of TILE*19 doubles per x-layer,
// writes are contiguous runs of TILE doubles along x; the transposition goes through shared
// memory (row pitch TILE*19+1 doubles: odd, so the strided phase is bank-conflict free).
__global__ void __launch_bounds__(256) aos_to_soa_kernel(Geom g, const double *__restrict__ aos, int xc, int x0, int c,
                                                         double *__restrict__ soa_a, double *__restrict__ soa_b) {
  __shared__ double tile[TILE][TILE * Q + 1];
  const int z0 = blockIdx.x * TILE, y = blockIdx.y, xi0 = blockIdx.z * TILE;
  const int nzt = min(TILE, g.nz - z0), nxt = min(TILE, xc - xi0);
  const int run = nzt * Q;
  for (int xi = 0; xi < nxt; xi++) {
    const double *src = aos + (((int64_t)(xi0 + xi) * g.ny + y) * g.nz + z0) * Q;
    for (int i = threadIdx.x; i < run; i += blockDim.x) tile[xi][i] = src[i];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < run * TILE; i += blockDim.x) {
    const int xi = i % TILE, zq = i / TILE;
    if (xi >= nxt) continue;
    const int zz = zq / Q, q = zq % Q;
    const int64_t dst = f_index(g, c, q, z0 + zz + 1, (int64_t)y * g.nx + x0 + xi0 + xi);
    const double v = tile[xi][zz * Q + q];
    soa_a[dst] = v;
    if (soa_b) soa_b[dst] = v;
  }
}

__global__ void __launch_bounds__(256) soa_to_aos_kernel(Geom g, const double *__restrict__ soa, int xc, int x0, int c,
                                                         double *__restrict__ aos) {
  __shared__ double tile[TILE][TILE * Q + 1];
  const int z0 = blockIdx.x * TILE, y = blockIdx.y, xi0 = blockIdx.z * TILE;
  const int nzt = min(TILE, g.nz - z0), nxt = min(TILE, xc - xi0);
  const int run = nzt * Q;
  for (int i = threadIdx.x; i < run * TILE; i += blockDim.x) {
    const int xi = i % TILE, zq = i / TILE;
    if (xi >= nxt) continue;
    const int zz = zq / Q, q = zq % Q;
    tile[xi][zz * Q + q] = soa[f_index(g, c, q, z0 + zz + 1, (int64_t)y * g.nx + x0 + xi0 + xi)];
  }
  __syncthreads();
  for (int xi = 0; xi < nxt; xi++) {
    double *dst = aos + (((int64_t)(xi0 + xi) * g.ny + y) * g.nz + z0) * Q;
    for (int i = threadIdx.x; i < run; i += blockDim.x) dst[i] = tile[xi][i];
  }
}

cudaError_t launch_aos_to_soa(const Geom &g, const double *aos, int xc, int x0, int c, double *soa_a, double *soa_b,
                              cudaStream_t s) {
  dim3 grid((g.nz + TILE - 1) / TILE, g.ny, (xc + TILE - 1) / TILE);
  aos_to_soa_kernel<<<grid, 256, 0, s>>>(g, aos, xc, x0, c, soa_a, soa_b);
  return cudaGetLastError();
}

cudaError_t launch_soa_to_aos(const Geom &g, const double *soa, int xc, int x0, int c, double *aos, cudaStream_t s) {
  dim3 grid((g.nz + TILE - 1) / TILE, g.ny, (xc + TILE - 1) / TILE);
  soa_to_aos_kernel<<<grid, 256, 0, s>>>(g, soa, xc, x0, c, aos);
  return cudaGetLastError();
}

// ---- mask: dense int32 [nx][ny][nz] <-> uint8 [zg][y][x] ------------------------------------
// flags[0] |= 1 when a value other than 0/1 is seen (enum Mask, mask.d:84-93); flags[1] counts
// solid sites (countSites, mask.d:105-116).
__global__ void mask_in_kernel(Geom g, const int32_t *__restrict__ in, uint8_t *__restrict__ out, int *flags) {
  __shared__ int32_t tile[32][33];
  const int x0 = blockIdx.x * 32, y = blockIdx.y, z0 = blockIdx.z * 32;
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {  // j: x within tile, threadIdx.x: z
    const int x = x0 + j, z = z0 + threadIdx.x;
    if (x < g.nx && z < g.nz) tile[j][threadIdx.x] = in[((int64_t)x * g.ny + y) * g.nz + z];
  }
  __syncthreads();
  int solid = 0;
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {  // j: z within tile, threadIdx.x: x
    const int x = x0 + threadIdx.x, z = z0 + j;
    if (x < g.nx && z < g.nz) {
      const int32_t v = tile[threadIdx.x][j];
      if (v != 0 && v != 1) atomicOr(&flags[0], 1);
      solid += (v == 1);
      out[((int64_t)(z + 1) * g.ny + y) * g.nx + x] = (uint8_t)(v == 1);
    }
  }
  if (solid) atomicAdd(&flags[1], solid);
}

__global__ void mask_out_kernel(Geom g, const uint8_t *__restrict__ in, int32_t *__restrict__ out) {
  __shared__ int32_t tile[32][33];
  const int x0 = blockIdx.x * 32, y = blockIdx.y, z0 = blockIdx.z * 32;
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int x = x0 + threadIdx.x, z = z0 + j;
    if (x < g.nx && z < g.nz) tile[j][threadIdx.x] = in[((int64_t)(z + 1) * g.ny + y) * g.nx + x];
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int x = x0 + j, z = z0 + threadIdx.x;
    if (x < g.nx && z < g.nz) out[((int64_t)x * g.ny + y) * g.nz + z] = tile[threadIdx.x][j];
  }
}

cudaError_t launch_mask_in(const Geom &g, const int32_t *in, uint8_t *out, int *flags, cudaStream_t s) {
  dim3 grid((g.nx + 31) / 32, g.ny, (g.nz + 31) / 32);
  mask_in_kernel<<<grid, dim3(32, 8), 0, s>>>(g, in, out, flags);
  return cudaGetLastError();
}
cudaError_t launch_mask_out(const Geom &g, const uint8_t *in, int32_t *out, cudaStream_t s) {
  dim3 grid((g.nx + 31) / 32, g.ny, (g.nz + 31) / 32);
  mask_out_kernel<<<grid, dim3(32, 8), 0, s>>>(g, in, out);
  return cudaGetLastError();
}

// ---- ghost planes when the slab is its own z neighbour ---------------------------------------
// For fields laid out [narrays][zg][y][x] (psi, wall densities).  The populations' ghost planes
// are whole contiguous plane blocks in the plane-major layout and are copied with cudaMemcpyAsync.
__global__ void ghost_fill_kernel(Geom g, double *__restrict__ f, int narrays) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.plane) return;
  for (int a = blockIdx.y; a < narrays; a += gridDim.y) {
    double *p = f + (int64_t)a * g.sites_g;
    p[i] = p[(int64_t)g.nz * g.plane + i];                           // ghost 0      <- plane nz
    p[(int64_t)(g.nz + 1) * g.plane + i] = p[g.plane + i];           // ghost nz + 1 <- plane 1
  }
}
__global__ void ghost_fill_mask_kernel(Geom g, uint8_t *__restrict__ m) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.plane) return;
  m[i] = m[(int64_t)g.nz * g.plane + i];
  m[(int64_t)(g.nz + 1) * g.plane + i] = m[g.plane + i];
}
// Fields::nb: bit q of nb[site] = mask[site + c_q] is Solid (bit 0 = the site itself), for the physical
// planes 1..nz; x and y wrap periodically, z neighbours come from the (already valid) ghost planes.
// Built once per mask upload: the geometry is static (the reference re-reads and re-exchanges the mask
// every step, lattice.d:200-214).
__global__ void neighbour_bits_kernel(Geom g, const uint8_t *__restrict__ m, uint32_t *__restrict__ nb) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= g.nx) return;
  const int y = blockIdx.y, zg = 1 + blockIdx.z;
  const Nbr n = make_nbr(g, x, y);
  const uint8_t *ms = m + (int64_t)zg * g.plane + n.sxy;
  uint32_t bits = 0;
#pragma unroll
  for (int q = 0; q < Q; q++) {
    if (ms[n.ox[1 + cx(q)] + n.oy[1 + cy(q)] + cz(q) * g.plane]) bits |= 1u << q;
  }
  nb[(int64_t)zg * g.plane + n.sxy] = bits;
}
cudaError_t launch_neighbour_bits(const Geom &g, const uint8_t *m, uint32_t *nb, cudaStream_t s) {
  const unsigned bx = g.nx >= 128 ? 128 : ((g.nx + 31) / 32) * 32;
  neighbour_bits_kernel<<<dim3((g.nx + bx - 1) / bx, g.ny, g.nz), bx, 0, s>>>(g, m, nb);
  return cudaGetLastError();
}

// Solid sites are static: the step kernels skip them, relying on BOTH population buffers holding their
// values (the reference copies them every step, advection.d:83-85).  After a mask upload that follows the
// populations, the sites that just became solid must therefore be copied into the other buffer once.
__global__ void sync_solid_kernel(Geom g, const uint8_t *__restrict__ m, const double *__restrict__ src,
                                  double *__restrict__ dst) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= g.nx) return;
  const int y = blockIdx.y, zg = blockIdx.z;  // ghost planes included
  const int64_t sxy = (int64_t)y * g.nx + x;
  if (!m[(int64_t)zg * g.plane + sxy]) return;
  for (int cq = 0; cq < g.ncomp * Q; cq++) {
    const int64_t i = (int64_t)zg * g.fz + (int64_t)cq * g.plane + sxy;
    dst[i] = src[i];
  }
}
cudaError_t launch_sync_solid(const Geom &g, const uint8_t *m, const double *src, double *dst, cudaStream_t s) {
  const unsigned bx = g.nx >= 128 ? 128 : ((g.nx + 31) / 32) * 32;
  sync_solid_kernel<<<dim3((g.nx + bx - 1) / bx, g.ny, g.nz + 2), bx, 0, s>>>(g, m, src, dst);
  return cudaGetLastError();
}

cudaError_t launch_ghost_fill(const Geom &g, double *f, int narrays, cudaStream_t s) {
  dim3 grid((unsigned)((g.plane + 255) / 256), (unsigned)min(narrays, 65535));
  ghost_fill_kernel<<<grid, 256, 0, s>>>(g, f, narrays);
  return cudaGetLastError();
}
cudaError_t launch_ghost_fill_mask(const Geom &g, uint8_t *m, cudaStream_t s) {
  ghost_fill_mask_kernel<<<(unsigned)((g.plane + 255) / 256), 256, 0, s>>>(g, m);
  return cudaGetLastError();
}

// ---- derived outputs -------------------------------------------------------------------------
// kind 0: density of comp c1 (densityField, density.d:89-96: every site, solid included)
// kind 1: colour c1 - c2 (colourField, density.d:284-295: zero on solid sites)
// kind 2: raw velocity of c1 (rawVelocityField, velocity.d:106-117: zero on solid sites)
// kind 3: own-site density into an SoA array (no transposition) for rho_wall
// Output (kinds 0-2): dense [nx][ny][nz][K] (K = 1 or 3), written through a 32x32 (x,z) tile.
__device__ __forceinline__ void load_site(const double *__restrict__ f, const Geom &g, int c, int zg, int sxy,
                                          double (&v)[Q]) {
  const double *__restrict__ p = f_plane(f, g, zg, c) + sxy;
#pragma unroll
  for (int q = 0; q < Q; q++) v[q] = p[qslot(q) * g.plane];
}

template <int KIND>
__global__ void output_kernel(Geom g, const double *__restrict__ f, const uint8_t *__restrict__ mask, int c1, int c2,
                              double *__restrict__ out) {
  constexpr int K = KIND == 2 ? 3 : 1;
  __shared__ double tile[K][32][33];
  const int x0 = blockIdx.x * 32, y = blockIdx.y, z0 = blockIdx.z * 32;
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {  // j: z in tile; threadIdx.x: x
    const int x = x0 + threadIdx.x, z = z0 + j;
    if (x >= g.nx || z >= g.nz) continue;
    const int64_t s = ((int64_t)(z + 1) * g.ny + y) * g.nx + x;
    const bool fluid = mask[s] == 0;
    double fa[Q];
    load_site(f, g, c1, z + 1, y * g.nx + x, fa);
    if (KIND == 0) {
      tile[0][j][threadIdx.x] = density19(fa);
    } else if (KIND == 1) {
      double fb[Q];
      load_site(f, g, c2, z + 1, y * g.nx + x, fb);
      tile[0][j][threadIdx.x] = fluid ? density19(fa) - density19(fb) : 0.0;
    } else {
      double m[3];
      momentum19(fa, m);
      const double rho = density19(fa);
#pragma unroll
      for (int d = 0; d < 3; d++) tile[d][j][threadIdx.x] = fluid ? m[d] / rho : 0.0;
    }
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {  // j: x in tile; threadIdx.x: z
    const int x = x0 + j, z = z0 + threadIdx.x;
    if (x >= g.nx || z >= g.nz) continue;
    const int64_t o = (((int64_t)x * g.ny + y) * g.nz + z) * K;
#pragma unroll
    for (int d = 0; d < K; d++) out[o + d] = tile[d][threadIdx.x][j];
  }
}

// kind 3: transpose an SoA vector field v[3][zg][y][x] into the dense host order [nx][ny][nz][3]
__global__ void vector_out_kernel(Geom g, const double *__restrict__ v, double *__restrict__ out) {
  __shared__ double tile[3][32][33];
  const int x0 = blockIdx.x * 32, y = blockIdx.y, z0 = blockIdx.z * 32;
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int x = x0 + threadIdx.x, z = z0 + j;
    if (x >= g.nx || z >= g.nz) continue;
    const int64_t s = ((int64_t)(z + 1) * g.ny + y) * g.nx + x;
#pragma unroll
    for (int d = 0; d < 3; d++) tile[d][j][threadIdx.x] = v[(int64_t)d * g.sites_g + s];
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int x = x0 + j, z = z0 + threadIdx.x;
    if (x >= g.nx || z >= g.nz) continue;
    const int64_t o = (((int64_t)x * g.ny + y) * g.nz + z) * 3;
#pragma unroll
    for (int d = 0; d < 3; d++) out[o + d] = tile[d][threadIdx.x][j];
  }
}

cudaError_t launch_output(const Geom &g, int kind, const double *f, const uint8_t *mask, int c1, int c2, double *out,
                          cudaStream_t s) {
  dim3 grid((g.nx + 31) / 32, g.ny, (g.nz + 31) / 32), block(32, 8);
  if (kind == 3) {
    vector_out_kernel<<<grid, block, 0, s>>>(g, f, out);
    return cudaGetLastError();
  }
  if (kind == 0) output_kernel<0><<<grid, block, 0, s>>>(g, f, mask, c1, c2, out);
  else if (kind == 1) output_kernel<1><<<grid, block, 0, s>>>(g, f, mask, c1, c2, out);
  else if (kind == 2) output_kernel<2><<<grid, block, 0, s>>>(g, f, mask, c1, c2, out);
  else return cudaErrorInvalidValue;
  return cudaGetLastError();
}

// ---- the per-site device functions, one site at a time ---------------------------------------------------
// For the reference's in-source known-answer tests (density.d:50-61, velocity.d:59-74, force.d:370-384,
// eqdist.d:317-360): evaluates the functions of lb_site.cuh -- the ones every step kernel is built from, in their
// EXACT instantiation -- on n sites given as pop[n][19] (+ dv[n][3] for the equilibrium).
//   fn 0: density19 -> out[n];  1: raw velocity (momentum19 / rho, rho = in2[i] if given else density19) -> out[n][3];
//   2: psi_of<form>(in[i]) -> out[n];  3: equilibrium distribution of `form` with shift dv = in2[i] -> out[n][19]
__global__ void site_function_kernel(int fn, int form, int n, const double *__restrict__ in, const double *__restrict__ in2,
                                     double *__restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (fn == 2) {
    out[i] = form == 0 ? psi_of<0>(in[i]) : psi_of<1>(in[i]);
    return;
  }
  double f[Q];
  for (int q = 0; q < Q; q++) f[q] = in[(int64_t)i * Q + q];
  if (fn == 0) {
    out[i] = density19(f);
  } else if (fn == 1) {
    double m[3];
    momentum19(f, m);
    const double rho = in2 ? in2[i] : density19(f);
    for (int d = 0; d < 3; d++) out[(int64_t)i * 3 + d] = m[d] / rho;
  } else {
    double m[3], dv[3], eq[Q];
    momentum19(f, m);
    for (int d = 0; d < 3; d++) dv[d] = in2[(int64_t)i * 3 + d];
    const double rho = density19(f);
    if (form == 0) eqdist_poly<0, true>(m, dv, rho, eq);
    else if (form == 1) eqdist_poly<1, true>(m, dv, rho, eq);
    else eqdist_bdist2(dv, rho, eq);
    for (int q = 0; q < Q; q++) out[(int64_t)i * Q + q] = eq[q];
  }
}
cudaError_t launch_site_function(int fn, int form, int n, const double *in, const double *in2, double *out, cudaStream_t s) {
  site_function_kernel<<<(n + 127) / 128, 128, 0, s>>>(fn, form, n, in, in2, out);
  return cudaGetLastError();
}

// ---- more derived outputs ----------------------------------------------------------
// pressure (density.d:306-326): P = css * sum_i (rho_i + sum_j 0.5 * gcc[i][j] * psi(rho_i) * psi(rho_j)),
// accumulated in the reference's order; computed on every site from the densities found there.
template <int PSI>
__global__ void pressure_out_kernel(Geom g, const double *__restrict__ f, StepParams P, double *__restrict__ out) {
  __shared__ double tile[32][33];
  const int x0 = blockIdx.x * 32, y = blockIdx.y, z0 = blockIdx.z * 32;
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int x = x0 + threadIdx.x, z = z0 + j;
    if (x >= g.nx || z >= g.nz) continue;
    double rho[MAXC];
    for (int c = 0; c < P.ncomp; c++) {
      double fa[Q];
      load_site(f, g, c, z + 1, y * g.nx + x, fa);
      rho[c] = density19(fa);
    }
    double pr = 0.0;
    for (int i = 0; i < P.ncomp; i++) {
      pr += rho[i];
      for (int k = 0; k < P.ncomp; k++) pr += 0.5 * P.gcc[i * MAXC + k] * psi_of<PSI>(rho[i]) * psi_of<PSI>(rho[k]);
    }
    tile[j][threadIdx.x] = pr * (1.0 / 3.0);
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int x = x0 + j, z = z0 + threadIdx.x;
    if (x >= g.nx || z >= g.nz) continue;
    out[((int64_t)x * g.ny + y) * g.nz + z] = tile[threadIdx.x][j];
  }
}
cudaError_t launch_pressure(const Geom &g, const double *f, const StepParams &P, int psi_form, double *out, cudaStream_t s) {
  dim3 grid((g.nx + 31) / 32, g.ny, (g.nz + 31) / 32), block(32, 8);
  if (psi_form == 0) pressure_out_kernel<0><<<grid, block, 0, s>>>(g, f, P, out);
  else pressure_out_kernel<1><<<grid, block, 0, s>>>(g, f, P, out);
  return cudaGetLastError();
}

// corrected velocity (velocity.d:177-185): rawVelocity + tau * F / rho per dimension, zero on solid sites (as
// velocityField, velocity.d:106-117); F = the SoA force field [3][sites_g] of the last step
__global__ void corrected_velocity_kernel(Geom g, const double *__restrict__ f, const uint8_t *__restrict__ mask, int c,
                                          double tau, const double *__restrict__ F, double *__restrict__ out) {
  __shared__ double tile[3][32][33];
  const int x0 = blockIdx.x * 32, y = blockIdx.y, z0 = blockIdx.z * 32;
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int x = x0 + threadIdx.x, z = z0 + j;
    if (x >= g.nx || z >= g.nz) continue;
    const int64_t s = ((int64_t)(z + 1) * g.ny + y) * g.nx + x;
    const bool fluid = mask[s] == 0;
    double fa[Q], m[3];
    load_site(f, g, c, z + 1, y * g.nx + x, fa);
    momentum19(fa, m);
    const double rho = density19(fa);
#pragma unroll
    for (int d = 0; d < 3; d++) {
      double v = m[d] / rho;
      v += tau * F[(int64_t)d * g.sites_g + s] / rho;
      tile[d][j][threadIdx.x] = fluid ? v : 0.0;
    }
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int x = x0 + j, z = z0 + threadIdx.x;
    if (x >= g.nx || z >= g.nz) continue;
    const int64_t o = (((int64_t)x * g.ny + y) * g.nz + z) * 3;
#pragma unroll
    for (int d = 0; d < 3; d++) out[o + d] = tile[d][threadIdx.x][j];
  }
}
cudaError_t launch_corrected_velocity(const Geom &g, const double *f, const uint8_t *mask, int c, double tau, const double *F,
                                      double *out, cudaStream_t s) {
  dim3 grid((g.nx + 31) / 32, g.ny, (g.nz + 31) / 32), block(32, 8);
  corrected_velocity_kernel<<<grid, block, 0, s>>>(g, f, mask, c, tau, F, out);
  return cudaGetLastError();
}

// volumeAveragedDensity (laplace.d:236-272): sum and count over the fluid sites of the cube [lo, hi] (global
// coordinates, inclusive, not wrapped) that lie in this slab, in the reference's order (x, y, z; z fastest).
// out[2 c] = sum, out[2 c + 1] = number of sites.  A handful of sites: one thread.
__global__ void cube_density_kernel(Geom g, const double *__restrict__ f, const uint8_t *__restrict__ mask, int ncomp,
                                    int64_t z_first, int64_t lx, int64_t hx, int64_t ly, int64_t hy, int64_t lz, int64_t hz,
                                    double *__restrict__ out) {
  if (blockIdx.x != 0 || threadIdx.x != 0) return;
  for (int c = 0; c < ncomp; c++) {
    double sum = 0.0, cnt = 0.0;
    for (int64_t x = lx < 0 ? 0 : lx; x <= hx && x < g.nx; x++)
      for (int64_t y = ly < 0 ? 0 : ly; y <= hy && y < g.ny; y++)
        for (int64_t z = lz < z_first ? z_first : lz; z <= hz && z < z_first + g.nz; z++) {
          const int zg = (int)(z - z_first) + 1;
          const int sxy = (int)(y * g.nx + x);
          if (mask[(int64_t)zg * g.plane + sxy]) continue;
          double fa[Q];
          load_site(f, g, c, zg, sxy, fa);
          sum += density19(fa);
          cnt += 1.0;
        }
    out[2 * c] = sum;
    out[2 * c + 1] = cnt;
  }
}
cudaError_t launch_cube_density(const Geom &g, const double *f, const uint8_t *mask, int ncomp, int64_t z_first,
                                const int64_t lo[3], const int64_t hi[3], double *out, cudaStream_t s) {
  cube_density_kernel<<<1, 32, 0, s>>>(g, f, mask, ncomp, z_first, lo[0], hi[0], lo[1], hi[1], lo[2], hi[2], out);
  return cudaGetLastError();
}

// ---- initial conditions and geometry on the device -----------------------------------
// Mt19937 (Phobos std.random: the standard 32-bit Mersenne Twister, Knuth seeding; random.d:19-26) as ONE block:
// the 624-word state is regenerated in the three dependent sections of the reference algorithm (words 0..226,
// 227..453, 454..622, then 623), each section in parallel; tempered outputs skip .. skip + n - 1 are stored.
constexpr int MT_N = 624, MT_M = 397;
__global__ void __launch_bounds__(256) mt19937_kernel(uint32_t seed, int64_t skip, int64_t n, uint32_t *__restrict__ out) {
  __shared__ uint32_t mt[MT_N];
  if (threadIdx.x == 0) {
    uint32_t x = seed;
    mt[0] = x;
    for (int i = 1; i < MT_N; i++) {
      x = 1812433253u * (x ^ (x >> 30)) + (uint32_t)i;
      mt[i] = x;
    }
  }
  __syncthreads();
  const int64_t end = skip + n;
  for (int64_t base = 0; base < end; base += MT_N) {
    // twist: mt[k] = mt[k + M (mod N)] ^ (y >> 1) ^ (y & 1 ? 0x9908b0df : 0), y = upper bit of mt[k] | lower 31 of mt[k + 1]
    for (int sec = 0; sec < 4; sec++) {
      const int lo = sec == 0 ? 0 : (sec == 1 ? MT_N - MT_M : (sec == 2 ? 2 * (MT_N - MT_M) : MT_N - 1));
      const int hi = sec == 0 ? MT_N - MT_M : (sec == 1 ? 2 * (MT_N - MT_M) : (sec == 2 ? MT_N - 1 : MT_N));
      const int k = lo + (int)threadIdx.x;
      uint32_t v = 0;
      if (k < hi) {
        const uint32_t y = (mt[k] & 0x80000000u) | (mt[k + 1 == MT_N ? 0 : k + 1] & 0x7fffffffu);
        v = mt[k + MT_M < MT_N ? k + MT_M : k + MT_M - MT_N] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
      }
      __syncthreads();  // every input of the section is read before any of its words is replaced
      if (k < hi) mt[k] = v;
      __syncthreads();
    }
    for (int i = threadIdx.x; i < MT_N; i += blockDim.x) {
      const int64_t idx = base + i;
      if (idx < skip || idx >= end) continue;
      uint32_t y = mt[i];
      y ^= y >> 11;
      y ^= (y << 7) & 0x9d2c5680u;
      y ^= (y << 15) & 0xefc60000u;
      y ^= y >> 18;
      out[idx - skip] = y;
    }
    __syncthreads();
  }
}
cudaError_t launch_mt19937(uint32_t seed, int64_t skip, int64_t n, uint32_t *out, cudaStream_t s) {
  static_assert(MT_N - MT_M <= 256 && MT_N - 1 - 2 * (MT_N - MT_M) <= 256, "a section per pass of the block");
  mt19937_kernel<<<1, 256, 0, s>>>(seed, skip, n, out);
  return cudaGetLastError();
}

// initEqDist / initEqDistRandom / initEqDistPerturb / initEqDistPerturbFrac (fields/init.d:52-82) for the physical
// sites of one component, straight into the SoA buffers.  `raw` = the Mt19937 outputs of the reference's draw
// order: one per site of the rank's FULL array incl. its halo of 2, x slowest, z fastest (byElementForward);
// uniform(a, b) of Phobos for doubles = a + (nextafter(b, -inf) - a) * raw / 4294967295.0 (dlbc_b200/rng.py).
// kind 0: rho * eq; 1: rho * u * eq, u in [0, 2); 2: (rho + u) * eq, u in [-p, p); 3: rho * (1 + u) * eq
struct InitParams {
  int kind;
  double density, a, b_below;  // uniform(a, b): b_below = nextafter(b, -inf)
  double eq[Q];                // eqDistUnity (eqdist.d:310-315)
};
__global__ void __launch_bounds__(128) init_populations_kernel(Geom g, InitParams ip, const uint32_t *__restrict__ raw, int c,
                                                               double *__restrict__ fa, double *__restrict__ fb) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= g.nx) return;
  const int y = blockIdx.y, zg = blockIdx.z + 1;
  double rho = ip.density;
  if (ip.kind != 0) {
    const int64_t idx = ((int64_t)(x + 2) * (g.ny + 4) + (y + 2)) * (g.nz + 4) + (zg + 1);
    const double u = ip.a + (ip.b_below - ip.a) * (double)raw[idx] / 4294967295.0;
    rho = ip.kind == 1 ? ip.density * u : (ip.kind == 2 ? ip.density + u : ip.density * (1.0 + u));
  }
  const int64_t base = (int64_t)zg * g.fz + (int64_t)c * g.fc + (int64_t)y * g.nx + x;
#pragma unroll
  for (int q = 0; q < Q; q++) {
    const double v = rho * ip.eq[q];
    fa[base + (int64_t)qslot(q) * g.plane] = v;
    fb[base + (int64_t)qslot(q) * g.plane] = v;
  }
}
cudaError_t launch_init_populations(const Geom &g, int kind, double density, double a, double b_below, const double *eq,
                                    const uint32_t *raw, int c, double *fa, double *fb, cudaStream_t s) {
  InitParams ip;
  ip.kind = kind; ip.density = density; ip.a = a; ip.b_below = b_below;
  for (int q = 0; q < Q; q++) ip.eq[q] = eq[q];
  const unsigned bx = g.nx >= 128 ? 128 : ((g.nx + 31) / 32) * 32;
  init_populations_kernel<<<dim3((g.nx + bx - 1) / bx, g.ny, g.nz), bx, 0, s>>>(g, ip, raw, c, fa, fb);
  return cudaGetLastError();
}

// initEqDistSphere / TwoSpheres / Cylinder (fields/init.d:84-259): rho = symmetricLinearTransition(distance to the
// nearest centre - radius, thickness, d1, d2) (init.d:346-366), distance = sqrt(sum_i (gn_i - centre_i + 0.5)^2)
// accumulated over i = 0, 1, 2 (the skipped axis of a cylinder contributes 0.0); or (kind 1) the value of a 1-D
// profile along `axis` (lamellae, init.d:261-313: the host evaluates the profile, gn[axis] numbers).
struct ProfileParams {
  int kind;         // 0 spheres / cylinder, 1 axis profile
  int ncentres, skip_axis, axis;
  double centre[2][3], radius, width, negval, posval;
  int64_t z_first;  // global z of local plane 1
  double eq[Q];
};
__device__ __forceinline__ double symmetric_linear_transition(double pos, double width, double negval, double posval) {
  if (width == 0.0) return pos < 0.0 ? negval : posval;
  if (pos < -0.5 * width) return negval;
  if (pos > 0.5 * width) return posval;
  return negval * (0.5 - pos / width) + posval * (0.5 + pos / width);
}
__global__ void __launch_bounds__(128) init_profile_kernel(Geom g, ProfileParams pp, const double *__restrict__ profile, int c,
                                                           double *__restrict__ fa, double *__restrict__ fb) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= g.nx) return;
  const int y = blockIdx.y, zg = blockIdx.z + 1;
  const int64_t p[3] = {x, y, pp.z_first + zg - 1};
  double rho;
  if (pp.kind == 1) {
    rho = profile[p[pp.axis]];
  } else {
    double rel = 0.0;
    for (int k = 0; k < pp.ncentres; k++) {
      double acc = 0.0;
      for (int d = 0; d < 3; d++) {
        const double off = d == pp.skip_axis ? 0.0 : (double)p[d] - pp.centre[k][d] + 0.5;
        acc = acc + off * off;
      }
      const double r = sqrt(acc) - pp.radius;
      rel = k == 0 ? r : (r < rel ? r : rel);
    }
    rho = symmetric_linear_transition(rel, pp.width, pp.negval, pp.posval);
  }
  const int64_t base = (int64_t)zg * g.fz + (int64_t)c * g.fc + (int64_t)y * g.nx + x;
#pragma unroll
  for (int q = 0; q < Q; q++) {
    const double v = rho * pp.eq[q];
    fa[base + (int64_t)qslot(q) * g.plane] = v;
    fb[base + (int64_t)qslot(q) * g.plane] = v;
  }
}
cudaError_t launch_init_profile(const Geom &g, int kind, int ncentres, const double *centres, int skip_axis, int axis, double radius,
                                double width, double negval, double posval, const double *profile, int64_t z_first,
                                const double *eq, int c, double *fa, double *fb, cudaStream_t s) {
  ProfileParams pp;
  pp.kind = kind; pp.ncentres = ncentres; pp.skip_axis = skip_axis; pp.axis = axis;
  for (int k = 0; k < 2; k++)
    for (int d = 0; d < 3; d++) pp.centre[k][d] = centres && k < ncentres ? centres[3 * k + d] : 0.0;
  pp.radius = radius; pp.width = width; pp.negval = negval; pp.posval = posval; pp.z_first = z_first;
  for (int q = 0; q < Q; q++) pp.eq[q] = eq[q];
  const unsigned bx = g.nx >= 128 ? 128 : ((g.nx + 31) / 32) * 32;
  init_profile_kernel<<<dim3((g.nx + bx - 1) / bx, g.ny, g.nz), bx, 0, s>>>(g, pp, profile, c, fa, fb);
  return cudaGetLastError();
}

// initEqDistWall (fields/init.d:469-477): populations of solid sites := (1, 0, ..., 0), every component, both
// buffers, ghost planes included (their mask is valid after the upload / generation of the mask)
__global__ void init_wall_kernel(Geom g, const uint8_t *__restrict__ m, double *__restrict__ fa, double *__restrict__ fb) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= g.nx) return;
  const int y = blockIdx.y, zg = blockIdx.z;
  const int64_t sxy = (int64_t)y * g.nx + x;
  if (!m[(int64_t)zg * g.plane + sxy]) return;
  for (int c = 0; c < g.ncomp; c++)
    for (int q = 0; q < Q; q++) {
      const int64_t i = (int64_t)zg * g.fz + (int64_t)c * g.fc + (int64_t)qslot_rt(q) * g.plane + sxy;
      const double v = q == 0 ? 1.0 : 0.0;
      fa[i] = v;
      fb[i] = v;
    }
}
cudaError_t launch_init_wall(const Geom &g, const uint8_t *m, double *fa, double *fb, cudaStream_t s) {
  const unsigned bx = g.nx >= 128 ? 128 : ((g.nx + 31) / 32) * 32;
  init_wall_kernel<<<dim3((g.nx + bx - 1) / bx, g.ny, g.nz + 2), bx, 0, s>>>(g, m, fa, fb);
  return cudaGetLastError();
}

// initMask (lb/mask.d:134-172; fields/init.d:377-467): kind 0 None, 1 Box, 2 Tube (open along `axis`),
// 3 Plates (normal to `axis`, `wall_offset` sites in), 4 Walls (per dimension and side, `thickness` sites).
// Physical planes only, global coordinates (z_first = global z of local plane 1); counts the solid sites.
struct MaskParams {
  int kind, axis, wall_offset, thickness;
  int has_wall[3][2];
  int64_t gn[3], z_first;
};
__global__ void init_mask_kernel(Geom g, MaskParams mp, uint8_t *__restrict__ m, int *__restrict__ nsolid) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= g.nx) return;
  const int y = blockIdx.y, zg = blockIdx.z + 1;
  const int64_t p[3] = {x, y, mp.z_first + zg - 1};
  bool solid = false;
  for (int d = 0; d < 3; d++) {
    const bool lo = p[d] == 0, hi = p[d] == mp.gn[d] - 1;
    if (mp.kind == 1) solid |= lo || hi;
    else if (mp.kind == 2) solid |= d != mp.axis && (lo || hi);
    else if (mp.kind == 3) solid |= d == mp.axis && (p[d] == mp.wall_offset || p[d] == mp.gn[d] - 1 - mp.wall_offset);
    else if (mp.kind == 4) solid |= (mp.has_wall[d][0] && p[d] < mp.thickness) || (mp.has_wall[d][1] && p[d] >= mp.gn[d] - mp.thickness);
  }
  m[(int64_t)zg * g.plane + (int64_t)y * g.nx + x] = solid ? 1 : 0;
  if (solid) atomicAdd(nsolid, 1);
}
cudaError_t launch_init_mask(const Geom &g, int kind, int axis, int wall_offset, int thickness, const int has_wall[6],
                             const int64_t gn[3], int64_t z_first, uint8_t *m, int *nsolid, cudaStream_t s) {
  MaskParams mp;
  mp.kind = kind; mp.axis = axis; mp.wall_offset = wall_offset; mp.thickness = thickness;
  for (int d = 0; d < 3; d++) { mp.has_wall[d][0] = has_wall[2 * d]; mp.has_wall[d][1] = has_wall[2 * d + 1]; mp.gn[d] = gn[d]; }
  mp.z_first = z_first;
  const unsigned bx = g.nx >= 128 ? 128 : ((g.nx + 31) / 32) * 32;
  init_mask_kernel<<<dim3((g.nx + bx - 1) / bx, g.ny, g.nz), bx, 0, s>>>(g, mp, m, nsolid);
  return cudaGetLastError();
}

// Own-site densities of every site incl. ghosts, SoA (for the wall term, force.d:298: the density
// "on the wall" is the sum of the solid site's own, never-advected populations).
__global__ void site_density_kernel(Geom g, const double *__restrict__ f, int c, double *__restrict__ rho) {
  const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= g.sites_g) return;
  double v[Q];
  load_site(f, g, c, (int)(s / g.plane), (int)(s % g.plane), v);
  rho[s] = density19(v);
}
cudaError_t launch_site_density(const Geom &g, const double *f, int c, double *rho, cudaStream_t s) {
  site_density_kernel<<<(unsigned)((g.sites_g + 255) / 256), 256, 0, s>>>(g, f, c, rho);
  return cudaGetLastError();
}

// ---- reductions: mass (1 value) and momentum (3 values) over physical fluid sites ---------------
// Deterministic: fixed grid, each block reduces a fixed range in a fixed tree order, then one
// block reduces the partials.  partial layout: [4][REDUCE_BLOCKS].
__device__ __forceinline__ double block_sum(double v, double *sh) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) sh[w] = v;
  __syncthreads();
  double r = 0.0;
  if (w == 0) {
    r = l < (int)(blockDim.x >> 5) ? sh[l] : 0.0;
    for (int o = 16; o > 0; o >>= 1) r += __shfl_down_sync(0xffffffffu, r, o);
  }
  __syncthreads();
  return r;
}

__global__ void __launch_bounds__(256) reduce_sites_kernel(Geom g, const double *__restrict__ fp, int c,
                                                           const uint8_t *__restrict__ mask, double *__restrict__ partial) {
  __shared__ double sh[8];
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  const int64_t n = (int64_t)g.nz * g.plane;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t s = i + g.plane;  // skip ghost plane 0
    if (mask[s]) continue;          // isFluid only (density.d:142, momentum.d:135)
    double f[Q], m[3];
    load_site(fp, g, c, (int)(s / g.plane), (int)(s % g.plane), f);
    momentum19(f, m);
    acc[0] += density19(f);
    acc[1] += m[0]; acc[2] += m[1]; acc[3] += m[2];
  }
  for (int k = 0; k < 4; k++) {
    const double r = block_sum(acc[k], sh);
    if (threadIdx.x == 0) partial[k * REDUCE_BLOCKS + blockIdx.x] = r;
  }
}

__global__ void __launch_bounds__(256) reduce_final_kernel(double *__restrict__ partial) {
  __shared__ double sh[8];
  for (int k = 0; k < 4; k++) {
    double v = 0.0;
    for (int i = threadIdx.x; i < REDUCE_BLOCKS; i += blockDim.x) v += partial[k * REDUCE_BLOCKS + i];
    const double r = block_sum(v, sh);
    if (threadIdx.x == 0) partial[4 * REDUCE_BLOCKS + k] = r;
  }
}

cudaError_t launch_reduce(const Geom &g, const double *f, int c, const uint8_t *mask, double *partial, cudaStream_t s) {
  reduce_sites_kernel<<<REDUCE_BLOCKS, 256, 0, s>>>(g, f, c, mask, partial);
  reduce_final_kernel<<<1, 256, 0, s>>>(partial);
  return cudaGetLastError();
}

}  // namespace dlbc

// no_alloc.cu -- dlbc_cuda_mass_momentum_no_alloc: the mass / momentum reduction of field_io.cu
// (reduce_sites_kernel + reduce_final_kernel, behind dlbc_cuda_total_mass / dlbc_cuda_total_momentum) for callers
// that cannot call cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that forbid it.
// It has no counterpart in the reference.
//
// The existing entry points keep their per-block partial sums in ctx->partial, which dlbc_cuda_create
// cudaMallocs.  Here the partials live in ONE file-scope __device__ array of fixed size (below): every GPU that
// loads the library holds DLBC_NO_ALLOC_SCRATCH_BYTES of it until the process exits, nothing is allocated or freed
// by a call, and there is no context -- the caller passes device pointers to the populations, the mask and the
// result.  Consequences, all documented in include/dlbc_cuda.h:
//   * the array is shared by every call on a device: two calls that overlap on the GPU (different streams, no
//     dependency between them) race on it;
//   * it keeps the last call's contents.  Nothing has to start at zero: every block of the first kernel writes its
//     four partials unconditionally (a block without sites writes 0.0), and the second kernel reads exactly the
//     partials the first one of the same call wrote.
//
// Same grid, same per-thread site ranges, same tree order as field_io.cu: on the same data the sums are
// bit-identical to dlbc_cuda_total_mass / dlbc_cuda_total_momentum of a single-rank context.  This unit is
// compiled with -fmad=false like field_io.cu.
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#include "dlbc_cuda.h"
#include "field_io.h"
#include "lb_site.cuh"

namespace dlbc {
namespace {

constexpr int NA_MAXC = DLBC_CUDA_NO_ALLOC_MAX_COMPONENTS;
constexpr int NA_DOUBLES = NA_MAXC * 4 * REDUCE_BLOCKS;  // partial[c][k][block]
static_assert(NA_DOUBLES * sizeof(double) == DLBC_CUDA_NO_ALLOC_SCRATCH_BYTES, "header and scratch array disagree");

// In HOST code this name is a host-side stand-in for the symbol, not a device address: the launches below get the
// address cudaGetSymbolAddress returns for the current device (scratch_for_current_device).
static __device__ __align__(256) double na_scratch[NA_DOUBLES];

// reduction tree of field_io.cu (block_sum): warp shuffles, one value per warp through shared memory, warp 0
__device__ __forceinline__ double na_block_sum(double v, double *sh) {
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

// blockIdx.y = component.  pops[c][q][site], mask[site] (nonzero: not a fluid site; may be null), site = (z, y, x)
// with x fastest: the order in which reduce_sites_kernel walks the physical planes of the device layout.
__global__ void __launch_bounds__(256) na_reduce_sites_kernel(int64_t sites, const double *__restrict__ pops,
                                                              const uint8_t *__restrict__ mask, double *__restrict__ partial) {
  __shared__ double sh[8];
  const int c = blockIdx.y;
  const double *__restrict__ p = pops + (int64_t)c * Q * sites;
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < sites; i += (int64_t)gridDim.x * blockDim.x) {
    if (mask && mask[i]) continue;
    double f[Q], m[3];
#pragma unroll
    for (int q = 0; q < Q; q++) f[q] = p[q * sites + i];
    momentum19(f, m);
    acc[0] += density19(f);
    acc[1] += m[0]; acc[2] += m[1]; acc[3] += m[2];
  }
  for (int k = 0; k < 4; k++) {
    const double r = na_block_sum(acc[k], sh);
    if (threadIdx.x == 0) partial[(c * 4 + k) * REDUCE_BLOCKS + blockIdx.x] = r;
  }
}

// blockIdx.x = component: out[c][0] = mass, out[c][1..3] = momentum
__global__ void __launch_bounds__(256) na_reduce_final_kernel(const double *__restrict__ partial, double *__restrict__ out) {
  __shared__ double sh[8];
  const int c = blockIdx.x;
  for (int k = 0; k < 4; k++) {
    double v = 0.0;
    for (int i = threadIdx.x; i < REDUCE_BLOCKS; i += blockDim.x) v += partial[(c * 4 + k) * REDUCE_BLOCKS + i];
    const double r = na_block_sum(v, sh);
    if (threadIdx.x == 0) out[c * 4 + k] = r;
  }
}

int fail(char *err, size_t cap, int code, const char *what, const char *detail) {
  if (err && cap) snprintf(err, cap, "dlbc_cuda_mass_momentum_no_alloc: %s%s%s", what, detail ? ": " : "", detail ? detail : "");
  return code;
}

// device address of na_scratch on the CURRENT device, looked up once per device (plain table: no allocation)
cudaError_t scratch_for_current_device(double **out) {
  static double *addr[64] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
  if (!addr[dev]) {
    double *p = nullptr;
    e = cudaGetSymbolAddress((void **)&p, na_scratch);
    if (e != cudaSuccess) return e;
    addr[dev] = p;
  }
  *out = addr[dev];
  return cudaSuccess;
}

}  // namespace
}  // namespace dlbc

using namespace dlbc;

extern "C" int dlbc_cuda_mass_momentum_no_alloc(int ncomp, int64_t nx, int64_t ny, int64_t nz, const double *d_pops,
                                                const uint8_t *d_mask, double *d_out, void *stream, char *err, size_t err_cap) {
  if (err && err_cap) err[0] = 0;
  if (ncomp < 1 || ncomp > NA_MAXC) {
    char d[160];
    snprintf(d, sizeof d, "ncomp = %d, the static scratch array holds the partial sums of at most %d components (%d bytes)",
             ncomp, NA_MAXC, (int)DLBC_CUDA_NO_ALLOC_SCRATCH_BYTES);
    return fail(err, err_cap, 1, "size beyond the maximum", d);
  }
  if (nx < 1 || ny < 1 || nz < 1) return fail(err, err_cap, 1, "bad argument", "nx, ny, nz must be positive");
  if (nx > DLBC_CUDA_NO_ALLOC_MAX_SITES / ny || nx * ny > DLBC_CUDA_NO_ALLOC_MAX_SITES / nz) {
    char d[160];
    snprintf(d, sizeof d, "%lld x %lld x %lld sites, at most %lld per call", (long long)nx, (long long)ny, (long long)nz,
             (long long)DLBC_CUDA_NO_ALLOC_MAX_SITES);
    return fail(err, err_cap, 1, "size beyond the maximum", d);
  }
  if (!d_pops || !d_out) return fail(err, err_cap, 1, "bad argument", "null device pointer");
  double *scratch = nullptr;
  cudaError_t e = scratch_for_current_device(&scratch);
  if (e != cudaSuccess) return fail(err, err_cap, 1000 + (int)e, "cudaGetSymbolAddress(scratch)", cudaGetErrorString(e));
  const int64_t sites = nx * ny * nz;
  cudaStream_t s = (cudaStream_t)stream;
  na_reduce_sites_kernel<<<dim3(REDUCE_BLOCKS, ncomp), 256, 0, s>>>(sites, d_pops, d_mask, scratch);
  na_reduce_final_kernel<<<ncomp, 256, 0, s>>>(scratch, d_out);
  e = cudaGetLastError();
  if (e != cudaSuccess) return fail(err, err_cap, 1000 + (int)e, "kernel launch", cudaGetErrorString(e));
  return 0;
}

// ubench_tma.cu -- how fast can ONE SM's producer warps feed shared memory with many SMALL TMA copies?
// (design input for step_tma.cuh: the wave items need ~40 row copies of 0.5-2 KB each).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_tma scripts/ubench_tma.cu
//   ./ubench_tma
// Each CTA (one per SM) runs NW producer warps over a ring of NS stages; a stage is K copies of S bytes on
// one mbarrier.  No consumer: a stage is reused as soon as its copies have landed.  Reports aggregate GB/s and
// ns per copy per SM for linear bulk copies (cp.async.bulk) and 2-D tensor-map boxes (cp.async.bulk.tensor),
// issued by every lane (the compiler serialises them) or by lane 0 only, source L2-resident or DRAM-resident.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n.reg .pred P1;\nLAB_WAIT:\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n@P1 bra DONE;\nbra LAB_WAIT;\nDONE:\n}\n" ::"r"(bar),
      "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(dst), "l"(src),
               "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void tensor2d_g2s(uint32_t dst, const CUtensorMap *map, int x, int y, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n" ::"r"(dst),
               "l"(map), "r"(bar), "r"(x), "r"(y) : "memory");
}

struct Args {
  const double *src;
  long long span_doubles;  // source region the copies walk through (power of two)
  int S;                   // bytes per copy
  int K;                   // copies per stage
  int NS;                  // stages per CTA
  int NW;                  // producer warps
  int iters;               // stages filled per CTA
  int mode;                // 0 linear all lanes, 1 linear lane 0, 2 tensor all lanes, 3 tensor lane 0
  int bx, by;              // tensor box (doubles, rows)
  int ncols;               // tensor row length (doubles)
  int shift;               // 1: box starts at odd element offsets too (8-byte granular), 0: 16-byte aligned starts
};

extern __shared__ __align__(128) unsigned char smem[];

__global__ void __launch_bounds__(128, 1) feed(Args a, const __grid_constant__ CUtensorMap map) {
  __shared__ unsigned long long full[32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < a.NS; i++) mbar_init(smem_u32(&full[i]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  }
  __syncthreads();
  if (warp >= a.NW) return;
  const long long mask = a.span_doubles - 1;
  const int sd = a.S / 8;
  for (int n = warp; n < a.iters; n += a.NW) {
    const int st = n % a.NS;
    const uint32_t bar = smem_u32(&full[st]);
    const uint32_t base = smem_u32(smem) + st * a.K * a.S;
    if (n >= a.NS) mbar_wait(bar, ((n / a.NS) - 1) & 1);
    if (lane == 0) mbar_arrive_tx(bar, (uint32_t)(a.K * a.S));
    __syncwarp();
    const long long item = (long long)n * gridDim.x + blockIdx.x;  // neighbouring SMs read neighbouring regions
    if (a.mode == 0 || a.mode == 2) {
      for (int j = lane; j < a.K; j += 32) {
        if (a.mode == 0) {
          const long long off = ((item * a.K + j) * sd) & mask;
          bulk_g2s(base + j * a.S, a.src + off, a.S, bar);
        } else {
          const long long row = ((item * a.K + j) * a.by) % (a.span_doubles / a.ncols - a.by);
          tensor2d_g2s(base + j * a.S, &map, 64 * (j & 3) + ((j % 3) - 1) * a.shift + 8, (int)row, bar);
        }
      }
    } else if (lane == 0) {
      for (int j = 0; j < a.K; j++) {
        if (a.mode == 1) {
          const long long off = ((item * a.K + j) * sd) & mask;
          bulk_g2s(base + j * a.S, a.src + off, a.S, bar);
        } else {
          const long long row = ((item * a.K + j) * a.by) % (a.span_doubles / a.ncols - a.by);
          tensor2d_g2s(base + j * a.S, &map, 64 * (j & 3) + ((j % 3) - 1) * a.shift + 8, (int)row, bar);
        }
      }
    }
    __syncwarp();
  }
  // drain: wait for the last use of every stage this warp filled
  for (int n = warp; n < a.iters; n += a.NW)
    if (n + a.NS >= a.iters) mbar_wait(smem_u32(&full[n % a.NS]), (n / a.NS) & 1);
}

typedef CUresult (*encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                              const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                              CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  const long long big = 1ll << 30;  // doubles: 8 GB
  double *buf;
  CK(cudaMalloc(&buf, big * 8));
  CK(cudaMemset(buf, 0, big * 8));
  encode_fn encode = nullptr;
  cudaDriverEntryPointQueryResult qres;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void **)&encode, cudaEnableDefault, &qres));
  CK(cudaFuncSetAttribute(feed, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int ncols = 512;
  printf("%-28s %6s %3s %3s %3s %9s %10s %9s\n", "mode", "S", "K", "NS", "NW", "src", "GB/s", "ns/copy/SM");
  const int mode0 = getenv("UB_MODE0") ? atoi(getenv("UB_MODE0")) : 0;
  const int shift = getenv("UB_SHIFT") ? atoi(getenv("UB_SHIFT")) : 1;
  const int quick = getenv("UB_QUICK") ? 1 : 0;
  for (int dram = 0; dram < 2; dram++)
    for (int mode = mode0; mode < 4; mode++)
      for (int cfg = 0; cfg < 4; cfg++)
        for (int NW = 1; NW <= 4; NW *= 2) {
          const int Ss[4] = {512, 1024, 2048, 4096};
          const int bys[4] = {1, 2, 4, 8};
          Args a;
          a.src = buf;
          a.span_doubles = dram ? big : (1ll << 22);  // 8 GB vs 32 MB
          a.S = Ss[cfg]; a.bx = 64; a.by = bys[cfg];
          a.K = 38;
          a.NS = (190 * 1024) / (a.K * a.S);
          if (a.NS > 32) a.NS = 32;
          if (a.NS < NW) continue;  // a warp per stage at most (each stage's barrier has ONE producer at a time)
          a.NW = NW;
          a.iters = (int)((dram ? 96ll : 48ll) * 1024 * 1024 / (a.K * a.S));  // per SM
          a.mode = mode; a.ncols = ncols; a.shift = shift;
          if (quick) a.iters = 64;
          CUtensorMap map;
          cuuint64_t gdim[2] = {(cuuint64_t)ncols, (cuuint64_t)(a.span_doubles / ncols)};
          cuuint64_t gstr[1] = {(cuuint64_t)ncols * 8};
          cuuint32_t box[2] = {(cuuint32_t)a.bx, (cuuint32_t)a.by};
          cuuint32_t estr[2] = {1, 1};
          CUresult r = encode(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, buf, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
          if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
          const int smem_bytes = a.NS * a.K * a.S;
          for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0);
            feed<<<sms, 128, smem_bytes>>>(a, map);
            cudaEventRecord(e1);
            CK(cudaEventSynchronize(e1));
          }
          CK(cudaGetLastError());
          float ms;
          cudaEventElapsedTime(&ms, e0, e1);
          const double bytes = (double)sms * a.iters * a.K * a.S;
          const char *names[4] = {"linear, all lanes", "linear, lane 0", "tensor 2-D, all lanes", "tensor 2-D, lane 0"};
          printf("%-28s %6d %3d %3d %3d %9s %10.1f %9.1f\n", names[mode], a.S, a.K, a.NS, NW, dram ? "DRAM" : "L2", bytes / ms * 1e-6,
                 ms * 1e6 / ((double)a.iters * a.K));
        }
  return 0;
}

// dlbc_cuda.cu -- implementation of the C ABI declared in include/dlbc_cuda.h.
//
// Host-side orchestration only: context lifecycle, layout conversion on upload/download, the
// per-step launch schedule (single rank and z-slab multi rank with NCCL ghost-plane exchange
// overlapped with the interior kernels), outputs and diagnostics.  All arithmetic lives in the
// kernels (step_*.cuh, field_io.cu).  There is no CPU fallback anywhere in this file: every
// compute entry point needs a CUDA device and fails with an error message otherwise.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <string>
#include <vector>

#include "context.h"
#include "d2q9.h"
#include "field_io.h"
#include "nccl_dyn.h"

using namespace dlbc;

namespace {

thread_local std::string g_create_error;

struct KernelSet {
  launch_fn psi;
  force_fn force;
  launch_fn fused[2];   // [exact]
  launch_fn wave[2];    // [exact]
  launch_fn tma[2];     // [exact]
};

#define DLBC_SET(NC)                                                                                      \
  { launch_psi_c##NC, launch_force_c##NC, {launch_fused_fast_c##NC, launch_fused_exact_c##NC},            \
    {launch_wave_fast_c##NC, launch_wave_exact_c##NC},                                                    \
    {launch_tma_fast_c##NC, launch_tma_exact_c##NC} }
const KernelSet kKernels[DLBC_CUDA_MAX_COMPONENTS] = {DLBC_SET(1), DLBC_SET(2), DLBC_SET(3), DLBC_SET(4)};
#undef DLBC_SET

int set_error(dlbc_cuda_ctx *ctx, const std::string &msg, int code) {
  if (ctx) ctx->err = msg; else g_create_error = msg;
  return code ? code : 1;
}
int cuda_fail(dlbc_cuda_ctx *ctx, const char *what, cudaError_t e) {
  return set_error(ctx, std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")", 1000 + (int)e);
}
int nccl_fail(dlbc_cuda_ctx *ctx, const char *what, ncclResult_t r) {
  const char *s = nccl_api().GetErrorString ? nccl_api().GetErrorString(r) : "?";
  return set_error(ctx, std::string(what) + ": NCCL error " + std::to_string((int)r) + " (" + s + ")", 2000 + (int)r);
}

#define CK(expr)                                                   \
  do {                                                             \
    cudaError_t e_ = (expr);                                       \
    if (e_ != cudaSuccess) return cuda_fail(ctx, #expr, e_);       \
  } while (0)
#define NK(expr)                                                   \
  do {                                                             \
    ncclResult_t r_ = (expr);                                      \
    if (r_ != ncclSuccess) return nccl_fail(ctx, #expr, r_);       \
  } while (0)
#define TRY(expr)                \
  do {                           \
    int rc_ = (expr);            \
    if (rc_) return rc_;         \
  } while (0)
#define ARG(cond, msg)                                                 \
  do {                                                                 \
    if (!(cond)) return set_error(ctx, std::string("dlbc_cuda: ") + msg, 1); \
  } while (0)

// entry points that serve 2-D (D2Q9) contexts hand over to d2q9.cu before anything else
#define D2(call)                                                             \
  do {                                                                       \
    if (ctx && ctx->d2) {                                                    \
      cudaError_t e2_ = cudaSetDevice(ctx->device);                          \
      if (e2_ != cudaSuccess) return cuda_fail(ctx, "cudaSetDevice", e2_);   \
      return d2::call;                                                       \
    }                                                                        \
  } while (0)

bool is_pow2(double x) {
  int e;
  return x > 0.0 && std::frexp(x, &e) == 0.5;
}


int ensure_stage(dlbc_cuda_ctx *ctx, size_t bytes) {
  if (bytes <= ctx->stage_bytes) return 0;
  if (ctx->stage) {
    CK(cudaStreamSynchronize(ctx->compute));
    CK(cudaFree(ctx->stage));
    ctx->device_bytes -= (int64_t)ctx->stage_bytes;
    ctx->stage = nullptr;
    ctx->stage_bytes = 0;
  }
  CK(cudaMalloc(&ctx->stage, bytes));
  ctx->stage_bytes = bytes;
  ctx->device_bytes += (int64_t)bytes;
  return 0;
}

int release_stage(dlbc_cuda_ctx *ctx) {
  // large staging areas (outputs of a 512^3 lattice are GBs) are returned after use
  if (ctx->stage && ctx->stage_bytes > ((size_t)256 << 20)) {
    CK(cudaStreamSynchronize(ctx->compute));
    CK(cudaFree(ctx->stage));
    ctx->device_bytes -= (int64_t)ctx->stage_bytes;
    ctx->stage = nullptr;
    ctx->stage_bytes = 0;
  }
  return 0;
}

StepLaunch make_launch(dlbc_cuda_ctx *ctx, cudaStream_t s) {
  StepLaunch L;
  L.g = ctx->g;
  L.f.f_old = ctx->f[ctx->cur];
  L.f.f_new = ctx->f[1 - ctx->cur];
  L.f.psi = ctx->psi;
  L.f.rho_wall = ctx->rho_wall;
  L.f.mask = ctx->mask;
  L.f.nb = ctx->nbmask;
  L.p = ctx->p;
  L.eq = ctx->cfg.eqdist_form;
  L.psi = ctx->cfg.psi_form;
  L.sc = ctx->cfg.enable_shan_chen;
  L.mask = ctx->has_solid ? 1 : 0;
  L.stream = s;
  L.wave_sync = ctx->wave_sync[0];
  L.wave_tyb = ctx->opt_wave_tyb;
  L.wave_la = ctx->opt_wave_la;
  L.tma_pop3 = ctx->tma_maps ? &ctx->tma_pop3[ctx->cur] : nullptr;  // maps of f_old
  L.tma_pop1 = ctx->tma_maps ? &ctx->tma_pop1[ctx->cur] : nullptr;
  L.tma_psi = ctx->tma_maps ? &ctx->tma_psi : nullptr;
  return L;
}

// ---- NCCL ghost-plane exchange: replaces exchangeHalo (fields/parallel.d:40-143) -------------
// Rank r sends its plane nz of every array in the "up" list to rank r+1's ghost plane 0 and its
// plane 1 of every array in the "down" list to rank r-1's ghost plane nz+1 (periodic ring), and
// receives the mirror images.  Each plane of one array is one contiguous run of nx*ny elements in
// both device layouts: no packing.  Message order to one peer is [up..., down...] on both sides,
// so the ring also works when nranks == 2 and both neighbours are the same rank.
template <typename T>
struct PlaneXfer {
  std::vector<T *> send_up, recv_from_down;   // my plane nz   -> upper neighbour's ghost 0
  std::vector<T *> send_down, recv_from_up;   // my plane 1    -> lower neighbour's ghost nz+1
};

template <typename T>
int exchange_planes(dlbc_cuda_ctx *ctx, const PlaneXfer<T> &x, ncclDataType_t dt, cudaStream_t s) {
  NcclApi &N = nccl_api();
  const size_t plane = (size_t)ctx->g.plane;
  NK(N.GroupStart());
  for (T *p : x.send_up) NK(N.Send(p, plane, dt, ctx->up, ctx->nccl, s));
  for (T *p : x.send_down) NK(N.Send(p, plane, dt, ctx->down, ctx->nccl, s));
  for (T *p : x.recv_from_down) NK(N.Recv(p, plane, dt, ctx->down, ctx->nccl, s));
  for (T *p : x.recv_from_up) NK(N.Recv(p, plane, dt, ctx->up, ctx->nccl, s));
  NK(N.GroupEnd());
  return 0;
}

// fields laid out [narrays][zg][y][x] (psi, wall densities, mask): all arrays go both ways
template <typename T>
PlaneXfer<T> scalar_xfer(const dlbc_cuda_ctx *ctx, T *base, int narrays) {
  PlaneXfer<T> x;
  const Geom &g = ctx->g;
  for (int a = 0; a < narrays; a++) {
    T *p = base + (int64_t)a * g.sites_g;
    x.send_up.push_back(p + (int64_t)g.nz * g.plane);
    x.recv_from_down.push_back(p);
    x.send_down.push_back(p + g.plane);
    x.recv_from_up.push_back(p + (int64_t)(g.nz + 1) * g.plane);
  }
  return x;
}

// populations f[zg][c][q][y][x]: only the five directions per component that cross each face
// (+z: q in {5,9,10,11,17}; -z: q in {6,12,15,16,18}; connectivity.d:401-418).  The reference
// ships 2 full layers x 19 populations per face instead (fields/parallel.d:58-68).
PlaneXfer<double> population_xfer(const dlbc_cuda_ctx *ctx, double *f) {
  PlaneXfer<double> x;
  const Geom &g = ctx->g;
  for (int c = 0; c < ctx->cfg.ncomp; c++)
    for (int q = 0; q < Q; q++) {
      if (cz(q) == 1) {
        x.send_up.push_back(f + f_index(g, c, q, g.nz, 0));
        x.recv_from_down.push_back(f + f_index(g, c, q, 0, 0));
      } else if (cz(q) == -1) {
        x.send_down.push_back(f + f_index(g, c, q, 1, 0));
        x.recv_from_up.push_back(f + f_index(g, c, q, g.nz + 1, 0));
      }
    }
  return x;
}

// Make the ghost planes of f[cur] (and the static mask / wall densities) valid.  During time
// stepping the step kernels / the per-step exchange maintain them; this is the setup path after
// an upload (the reference's L.exchangeHalo() at main.d:142).
int ensure_ghosts(dlbc_cuda_ctx *ctx) {
  const Geom &g = ctx->g;
  if (!ctx->ghosts_valid) {
    double *f = ctx->f[ctx->cur];
    if (ctx->cfg.nranks == 1) {
      // a whole plane of all C*19 arrays is one contiguous block
      const size_t bytes = (size_t)g.fz * sizeof(double);
      CK(cudaMemcpyAsync(f, f + (int64_t)g.nz * g.fz, bytes, cudaMemcpyDeviceToDevice, ctx->compute));
      CK(cudaMemcpyAsync(f + (int64_t)(g.nz + 1) * g.fz, f + g.fz, bytes, cudaMemcpyDeviceToDevice, ctx->compute));
    } else {
      TRY(exchange_planes<double>(ctx, population_xfer(ctx, f), ncclDouble, ctx->compute));
    }
    ctx->ghosts_valid = true;
  }
  const bool need_wall = ctx->p.any_gwc && ctx->has_solid && ctx->cfg.enable_shan_chen;
  if (need_wall && !ctx->rho_wall_valid) {
    if (!ctx->rho_wall) {
      const size_t bytes = (size_t)ctx->cfg.ncomp * g.sites_g * sizeof(double);
      CK(cudaMalloc(&ctx->rho_wall, bytes));
      ctx->device_bytes += (int64_t)bytes;
    }
    for (int c = 0; c < ctx->cfg.ncomp; c++) {
      CK(launch_site_density(g, ctx->f[ctx->cur], c, ctx->rho_wall + (int64_t)c * g.sites_g, ctx->compute));
      ctx->launches++;
    }
    if (ctx->cfg.nranks == 1) {
      CK(launch_ghost_fill(g, ctx->rho_wall, ctx->cfg.ncomp, ctx->compute));
      ctx->launches++;
    } else {
      TRY(exchange_planes<double>(ctx, scalar_xfer(ctx, ctx->rho_wall, ctx->cfg.ncomp), ncclDouble, ctx->compute));
    }
    ctx->rho_wall_valid = true;
  }
  return 0;
}

// psi of the current populations' NEXT step input, all planes incl. ghosts (two-pass K1 + exchange)
int psi_all(dlbc_cuda_ctx *ctx, const StepLaunch &L) {
  const KernelSet &K = kKernels[ctx->cfg.ncomp - 1];
  CK(K.psi(L, 1, ctx->g.nz));
  ctx->launches++;
  if (ctx->cfg.nranks > 1) {
    TRY(exchange_planes<double>(ctx, scalar_xfer(ctx, ctx->psi, ctx->cfg.ncomp), ncclDouble, ctx->compute));
  }
  return 0;
}

struct KernelTimer {
  dlbc_cuda_ctx *ctx;
  bool on;
  std::vector<const char *> names;
  std::vector<cudaEvent_t> evs;
  cudaEvent_t get(size_t i) {
    while (ctx->ev_pool.size() <= i) {
      cudaEvent_t e;
      cudaEventCreate(&e);
      ctx->ev_pool.push_back(e);
    }
    return ctx->ev_pool[i];
  }
  void begin(const char *name) {
    if (!on) return;
    cudaEvent_t e = get(evs.size());
    cudaEventRecord(e, ctx->compute);
    evs.push_back(e);
    names.push_back(name);
  }
  void end() {
    if (!on) return;
    cudaEvent_t e = get(evs.size());
    cudaEventRecord(e, ctx->compute);
    evs.push_back(e);
  }
};

// Tensor maps of the TMA variant (step_tma.cuh): populations as the 5-D tensor f[zg][c][slot][y][x] with boxes
// [TMA_SEG + 4, 1, 3 or 1 slots, C, 1], psi as the 4-D tensor psi[c][zg][y][x] with box [TMA_SEG + 4, 1, 3, C].
// Out-of-range columns are zero-filled.  Needs an even nx (16-byte strides); C <= 2 (stage count, step_tma.cuh).
typedef CUresult (*tensor_map_encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                         const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                         CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
bool make_tma_maps(dlbc_cuda_ctx *ctx) {
  const Geom &g = ctx->g;
  const int C = ctx->cfg.ncomp;
  if (g.nx % 2 != 0 || C > 2) return false;
  static tensor_map_encode_fn encode = nullptr;
  if (!encode) {
    cudaDriverEntryPointQueryResult q;
    void *fn = nullptr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || !fn) return false;
    encode = (tensor_map_encode_fn)fn;
  }
  const cuuint32_t ones[5] = {1, 1, 1, 1, 1};
  for (int b = 0; b < 2; b++) {
    const cuuint64_t dim[5] = {(cuuint64_t)g.nx, (cuuint64_t)g.ny, (cuuint64_t)Q, (cuuint64_t)C, (cuuint64_t)(g.nz + 2)};
    const cuuint64_t str[4] = {(cuuint64_t)g.nx * 8, (cuuint64_t)g.plane * 8, (cuuint64_t)g.fc * 8, (cuuint64_t)g.fz * 8};
    for (int k = 0; k < 2; k++) {
      const cuuint32_t box[5] = {(cuuint32_t)(TMA_SEG + 4), 1, k == 0 ? 3u : 1u, (cuuint32_t)C, 1};
      CUtensorMap *m = k == 0 ? &ctx->tma_pop3[b] : &ctx->tma_pop1[b];
      if (encode(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 5, ctx->f[b], dim, str, box, ones, CU_TENSOR_MAP_INTERLEAVE_NONE,
                 CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return false;
    }
  }
  const cuuint64_t dim[4] = {(cuuint64_t)g.nx, (cuuint64_t)g.ny, (cuuint64_t)(g.nz + 2), (cuuint64_t)C};
  const cuuint64_t str[3] = {(cuuint64_t)g.nx * 8, (cuuint64_t)g.plane * 8, (cuuint64_t)g.sites_g * 8};
  const cuuint32_t box[4] = {(cuuint32_t)(TMA_SEG + 4), 1, 3, (cuuint32_t)C};
  return encode(&ctx->tma_psi, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 4, ctx->psi, dim, str, box, ones, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// The TMA variant has no mask path and needs the tensor maps (even nx, C <= 2).  AUTO falls back to the register
// path; an explicit request fails in step().
bool tma_ok(const dlbc_cuda_ctx *ctx) { return ctx->tma_maps && !ctx->has_solid; }
int effective_variant(const dlbc_cuda_ctx *ctx) {
  if (ctx->variant == DLBC_KERNEL_TMA && ctx->variant_auto && !tma_ok(ctx)) return DLBC_KERNEL_WAVE;
  return ctx->variant;
}

// One timestep: the statements of runTimeloop, main.d:179-195.
int step_once(dlbc_cuda_ctx *ctx, KernelTimer *kt) {
  const KernelSet &K = kKernels[ctx->cfg.ncomp - 1];
  const int exact = ctx->cfg.arithmetic == DLBC_ARITH_EXACT ? 1 : 0;
  const int nz = ctx->g.nz;
  const bool sc = ctx->cfg.enable_shan_chen != 0;
  const int variant = effective_variant(ctx);
  const bool single_pass = variant == DLBC_KERNEL_WAVE || variant == DLBC_KERNEL_TMA;
  StepLaunch L = make_launch(ctx, ctx->compute);
  launch_fn fused = variant == DLBC_KERNEL_TMA ? K.tma[exact] : variant == DLBC_KERNEL_WAVE ? K.wave[exact] : K.fused[exact];
  const char *fused_name = variant == DLBC_KERNEL_TMA ? "tma_kernel" : variant == DLBC_KERNEL_WAVE ? "wave_kernel" : "fused_kernel";

  if (ctx->cfg.nranks == 1) {
    if (sc && !single_pass) {
      if (kt) kt->begin("psi_kernel");
      CK(K.psi(L, 1, nz));
      ctx->launches++;
      if (kt) kt->end();
    }
    if (kt) kt->begin(fused_name);
    CK(fused(L, 1, nz));
    ctx->launches++;
    if (kt) kt->end();
  } else if (single_pass) {
    // z-slab ring, single-pass kernel.  The interior launch is persistent (it fills every SM until it
    // ends), so anything queued beside it starves: the boundary work therefore runs FIRST on the compute
    // stream -- psi of planes 1 and nz -> [comm: NCCL psi ghost exchange] -> planes 1 and nz -- and only the
    // exchange of the 5 crossing populations per side overlaps the interior planes 2..nz-1, which need
    // nothing from this step's exchanges (they recompute psi of planes 1 and nz from populations whose
    // ghost planes arrived at the end of the previous step).
    StepLaunch Lb = make_launch(ctx, ctx->compute), Lb2 = make_launch(ctx, ctx->compute);
    Lb.wave_sync = ctx->wave_sync[1];
    Lb2.wave_sync = ctx->wave_sync[2];
    CK(cudaStreamWaitEvent(ctx->compute, ctx->ev_ghost_f, 0));
    CK(K.psi(L, 1, 1)); ctx->launches++;
    if (nz > 1) { CK(K.psi(L, nz, 1)); ctx->launches++; }
    CK(cudaEventRecord(ctx->ev_bnd, ctx->compute));
    CK(cudaStreamWaitEvent(ctx->comm, ctx->ev_bnd, 0));
    TRY(exchange_planes<double>(ctx, scalar_xfer(ctx, ctx->psi, ctx->cfg.ncomp), ncclDouble, ctx->comm));
    CK(cudaEventRecord(ctx->ev_ghost_psi, ctx->comm));
    CK(cudaStreamWaitEvent(ctx->compute, ctx->ev_ghost_psi, 0));
    CK(fused(Lb, 1, 1)); ctx->launches++;
    if (nz > 1) { CK(fused(Lb2, nz, 1)); ctx->launches++; }
    CK(cudaEventRecord(ctx->ev_bnd, ctx->compute));
    CK(cudaStreamWaitEvent(ctx->comm, ctx->ev_bnd, 0));
    TRY(exchange_planes<double>(ctx, population_xfer(ctx, ctx->f[1 - ctx->cur]), ncclDouble, ctx->comm));
    CK(cudaEventRecord(ctx->ev_ghost_f, ctx->comm));
    if (nz > 2) { CK(fused(L, 2, nz - 2)); ctx->launches++; }
  } else {
    // z-slab ring.  compute stream: K1 boundary planes | K1 interior | K2 boundary | K2 interior;
    // comm stream: psi ghost exchange (overlaps K1 interior), f ghost exchange (overlaps K2
    // interior).  The reference exchanges 2 full layers of 19 populations per face and per
    // field, blocking, before every step (fields/parallel.d:70-103).
    const int lo = 1, hi = nz;
    // ghosts of f[cur] were written by the comm stream at the end of the previous step
    CK(cudaStreamWaitEvent(ctx->compute, ctx->ev_ghost_f, 0));
    if (sc) {
      CK(K.psi(L, lo, 1)); ctx->launches++;
      if (hi != lo) { CK(K.psi(L, hi, 1)); ctx->launches++; }
      CK(cudaEventRecord(ctx->ev_bnd, ctx->compute));
      CK(cudaStreamWaitEvent(ctx->comm, ctx->ev_bnd, 0));
      TRY(exchange_planes<double>(ctx, scalar_xfer(ctx, ctx->psi, ctx->cfg.ncomp), ncclDouble, ctx->comm));
      CK(cudaEventRecord(ctx->ev_ghost_psi, ctx->comm));
      if (nz > 2) { CK(K.psi(L, lo + 1, nz - 2)); ctx->launches++; }
      CK(cudaStreamWaitEvent(ctx->compute, ctx->ev_ghost_psi, 0));
    }
    CK(K.fused[exact](L, lo, 1)); ctx->launches++;
    if (hi != lo) { CK(K.fused[exact](L, hi, 1)); ctx->launches++; }
    CK(cudaEventRecord(ctx->ev_bnd, ctx->compute));
    CK(cudaStreamWaitEvent(ctx->comm, ctx->ev_bnd, 0));
    TRY(exchange_planes<double>(ctx, population_xfer(ctx, ctx->f[1 - ctx->cur]), ncclDouble, ctx->comm));
    CK(cudaEventRecord(ctx->ev_ghost_f, ctx->comm));
    if (nz > 2) { CK(K.fused[exact](L, lo + 1, nz - 2)); ctx->launches++; }
  }
  ctx->cur ^= 1;
  ctx->timestep++;
  ctx->force_valid = true;
  return 0;
}

int join_comm(dlbc_cuda_ctx *ctx) {
  if (ctx->cfg.nranks > 1) CK(cudaStreamWaitEvent(ctx->compute, ctx->ev_ghost_f, 0));
  return 0;
}

int check_ctx(dlbc_cuda_ctx *ctx) {
  if (!ctx) return set_error(nullptr, "dlbc_cuda: null context", 1);
  if (ctx->d2) return set_error(ctx, "dlbc_cuda: this entry point is not available on a 2-D (D2Q9) context", 1);
  cudaError_t e = cudaSetDevice(ctx->device);
  if (e != cudaSuccess) return cuda_fail(ctx, "cudaSetDevice", e);
  return 0;
}

// host <-> staging copies of xc consecutive x-layers of a [..][ny+2h][nz+2h][K] host array.
// `host_first` addresses the start (y/z halo included) of the first layer moved; the device side
// is dense and halo-free, [xc][ny][nz][K].
template <typename T>
int copy_layers(dlbc_cuda_ctx *ctx, T *host_first, T *dev_dense, int halo, int K, int64_t xc, bool to_device,
                cudaStream_t stream = nullptr) {
  const Geom &g = ctx->g;
  if (!stream) stream = ctx->compute;
  const int64_t NY = g.ny + 2 * halo, NZ = g.nz + 2 * halo;
  const size_t row = (size_t)g.nz * K * sizeof(T);
  const cudaMemcpyKind kind = to_device ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost;
  if (halo == 0) {
    if (to_device) CK(cudaMemcpyAsync((void *)dev_dense, (const void *)host_first, row * g.ny * xc, kind, stream));
    else CK(cudaMemcpyAsync((void *)host_first, (const void *)dev_dense, row * g.ny * xc, kind, stream));
    return 0;
  }
  for (int64_t i = 0; i < xc; i++) {
    T *h = host_first + (i * NY * NZ + (int64_t)halo * NZ + halo) * K;
    T *d = dev_dense + i * g.ny * g.nz * K;
    if (to_device) CK(cudaMemcpy2DAsync((void *)d, row, (const void *)h, (size_t)NZ * K * sizeof(T), row, g.ny, kind, stream));
    else CK(cudaMemcpy2DAsync((void *)h, (size_t)NZ * K * sizeof(T), (const void *)d, row, row, g.ny, kind, stream));
  }
  return 0;
}

int64_t chunk_layers(const dlbc_cuda_ctx *ctx, int K, size_t elem) {
  const int64_t layer = (int64_t)ctx->g.ny * ctx->g.nz * K * (int64_t)elem;
  int64_t xc = ((int64_t)192 << 20) / std::max<int64_t>(layer, 1);
  return std::max<int64_t>(1, std::min<int64_t>(xc, ctx->g.nx));
}

int download_field(dlbc_cuda_ctx *ctx, int kind, int c1, int c2, const double *vec_soa, double *out, int halo) {
  const Geom &g = ctx->g;
  const int K = (kind == 2 || kind == 3) ? 3 : 1;
  const size_t bytes = (size_t)g.nx * g.ny * g.nz * K * sizeof(double);
  TRY(ensure_stage(ctx, bytes));
  const double *src = kind == 3 ? vec_soa : ctx->f[ctx->cur];
  CK(launch_output(g, kind, src, ctx->mask, c1, c2, (double *)ctx->stage, ctx->compute));
  ctx->launches++;
  const int64_t NY = g.ny + 2 * halo, NZ = g.nz + 2 * halo;
  TRY(copy_layers<double>(ctx, out + (int64_t)halo * NY * NZ * K, (double *)ctx->stage, halo, K, g.nx, false));
  CK(cudaStreamSynchronize(ctx->compute));
  return release_stage(ctx);
}

// After the physical planes of ctx->mask have been written (upload or generation on the device): ghost planes,
// the "any solid site anywhere" decision, the neighbourhood bits, and the solid sites' values in both buffers.
int finish_mask(dlbc_cuda_ctx *ctx, int solid) {
  const Geom &g = ctx->g;
  if (ctx->cfg.nranks == 1) {
    CK(launch_ghost_fill_mask(g, ctx->mask, ctx->compute));
    ctx->launches++;
  } else {
    TRY(exchange_planes<uint8_t>(ctx, scalar_xfer(ctx, ctx->mask, 1), ncclUint8, ctx->compute));
    // "has_solid" also has to cover solids in the neighbours' boundary planes (our ghosts): count globally
    int *dflag = ctx->flags + 1;
    CK(cudaMemcpyAsync(dflag, &solid, sizeof(int), cudaMemcpyHostToDevice, ctx->compute));
    NK(nccl_api().AllReduce(dflag, dflag, 1, ncclInt32, ncclSum, ctx->nccl, ctx->compute));
    CK(cudaMemcpyAsync(&solid, dflag, sizeof(int), cudaMemcpyDeviceToHost, ctx->compute));
  }
  CK(cudaStreamSynchronize(ctx->compute));
  ctx->has_solid = solid > 0;
  if (ctx->has_solid) {
    if (!ctx->nbmask) {
      const size_t bytes = (size_t)g.sites_g * sizeof(uint32_t);
      CK(cudaMalloc(&ctx->nbmask, bytes));
      ctx->device_bytes += (int64_t)bytes;
    }
    CK(launch_neighbour_bits(g, ctx->mask, ctx->nbmask, ctx->compute));
    ctx->launches++;
    // a mask that arrives after the populations (or after some steps): make the static values of the solid
    // sites agree in both buffers (the ghost planes of f[cur] are refreshed first)
    TRY(ensure_ghosts(ctx));
    CK(launch_sync_solid(g, ctx->mask, ctx->f[ctx->cur], ctx->f[1 - ctx->cur], ctx->compute));
    ctx->launches++;
  }
  ctx->rho_wall_valid = false;
  ctx->force_valid = false;
  return 0;
}

// eqDistUnity (eqdist.d:310-315): eqDist of pop = (1, 0, ..., 0) at dv = 0 in the reference's operation order
void eq_dist_unity(int form, double eq[Q]) {
  if (form == DLBC_EQDIST_BDIST2) {
    const double aN1 = 1.0 / 36.0, aN0 = 1.0 / 3.0, cst = 1.0 - 0.0, const1 = aN1 * cst;
    for (int q = 0; q < Q; q++) eq[q] = (q == 0 ? aN0 * cst : (q < 7 ? 2.0 * (const1 + 0.0) : const1 + 0.0)) * 1.0;
  } else {
    for (int q = 0; q < Q; q++) eq[q] = (1.0 * weight(q)) * (1.0 + 0.0 + 0.0 - 0.0);
  }
}

// L.force[comp] of the last step into a fresh SoA array [3][sites_g] (the caller frees it): zeros before the
// first step and without Shan-Chen (resetForce, force.d:165-170)
int last_force(dlbc_cuda_ctx *ctx, int comp, double **vec_out) {
  const Geom &g = ctx->g;
  double *vec = nullptr;
  const size_t vbytes = (size_t)3 * g.sites_g * sizeof(double);
  CK(cudaMalloc(&vec, vbytes));
  int rc = 0;
  if (!ctx->force_valid || !ctx->cfg.enable_shan_chen) {
    cudaMemsetAsync(vec, 0, vbytes, ctx->compute);
  } else {
    // recompute psi of the last step's post-advection state from the previous buffer (its ghost
    // planes are still intact), then the force; the hot path never stores either
    ctx->cur ^= 1;  // f[cur] := the populations the last step read
    StepLaunch L = make_launch(ctx, ctx->compute);
    ctx->cur ^= 1;
    rc = psi_all(ctx, L);
    if (!rc) {
      cudaError_t e = kKernels[ctx->cfg.ncomp - 1].force(L, comp, vec, 1, g.nz);
      ctx->launches++;
      if (e != cudaSuccess) rc = cuda_fail(ctx, "force_kernel", e);
    }
  }
  if (rc) {
    cudaFree(vec);
    vec = nullptr;
  }
  *vec_out = vec;
  return rc;
}

}  // namespace

// =============================================================================================
extern "C" {

int dlbc_cuda_abi_version(void) { return DLBC_CUDA_ABI_VERSION; }

int dlbc_cuda_device_count(int *count) {
  dlbc_cuda_ctx *ctx = nullptr;
  ARG(count, "device_count: null pointer");
  cudaError_t e = cudaGetDeviceCount(count);
  if (e != cudaSuccess) {
    *count = 0;
    return cuda_fail(nullptr, "cudaGetDeviceCount", e);
  }
  return 0;
}

int dlbc_cuda_nccl_unique_id(uint8_t id[DLBC_CUDA_NCCL_ID_BYTES]) {
  dlbc_cuda_ctx *ctx = nullptr;
  static_assert(sizeof(ncclUniqueId) <= DLBC_CUDA_NCCL_ID_BYTES, "ncclUniqueId does not fit");
  ARG(id, "nccl_unique_id: null pointer");
  const char *miss = nccl_api().load();
  if (miss) return set_error(nullptr, miss, 1);
  ncclUniqueId uid;
  NK(nccl_api().GetUniqueId(&uid));
  std::memset(id, 0, DLBC_CUDA_NCCL_ID_BYTES);
  std::memcpy(id, &uid, sizeof(uid));
  return 0;
}

const char *dlbc_cuda_last_error(const dlbc_cuda_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int dlbc_cuda_host_alloc(void **ptr, size_t bytes) {
  dlbc_cuda_ctx *ctx = nullptr;
  ARG(ptr, "host_alloc: null pointer");
  CK(cudaHostAlloc(ptr, bytes, cudaHostAllocDefault));
  return 0;
}
int dlbc_cuda_host_free(void *ptr) {
  dlbc_cuda_ctx *ctx = nullptr;
  CK(cudaFreeHost(ptr));
  return 0;
}

int dlbc_cuda_destroy(dlbc_cuda_ctx *ctx) {
  if (!ctx) return 0;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  if (ctx->d2) d2::destroy(ctx->d2);
  if (ctx->nccl) nccl_api().CommDestroy(ctx->nccl);
  for (double *p : {ctx->f[0], ctx->f[1], ctx->psi, ctx->rho_wall, ctx->partial}) if (p) cudaFree(p);
  if (ctx->mask) cudaFree(ctx->mask);
  if (ctx->nbmask) cudaFree(ctx->nbmask);
  if (ctx->stage) cudaFree(ctx->stage);
  if (ctx->flags) cudaFree(ctx->flags);
  for (int *p : ctx->wave_sync) if (p) cudaFree(p);
  for (cudaEvent_t e : {ctx->ev_a, ctx->ev_b, ctx->ev_bnd, ctx->ev_ghost_psi, ctx->ev_ghost_f, ctx->ev_copy, ctx->ev_stage[0],
                        ctx->ev_stage[1]}) if (e) cudaEventDestroy(e);
  if (ctx->xfer) cudaStreamDestroy(ctx->xfer);
  for (cudaEvent_t e : ctx->ev_pool) cudaEventDestroy(e);
  for (cudaStream_t s : {ctx->compute, ctx->comm}) if (s) cudaStreamDestroy(s);
  delete ctx;
  return 0;
}

int dlbc_cuda_create(const dlbc_cuda_config *cfg, dlbc_cuda_ctx **out) {
  dlbc_cuda_ctx *ctx = nullptr;  // errors before the context exists go to the thread-local string
  ARG(cfg && out, "create: null pointer");
  *out = nullptr;
  ARG(cfg->abi_version == DLBC_CUDA_ABI_VERSION, "create: abi_version mismatch");
  ARG((cfg->d == 3 && cfg->q == 19) || (cfg->d == 2 && cfg->q == 9),
      "create: the lattices built are D3Q19 and D2Q9 (connectivity.d:21-32)");
  ARG(cfg->ncomp >= 1 && cfg->ncomp <= DLBC_CUDA_MAX_COMPONENTS, "create: ncomp out of range");
  if (cfg->d == 2) {
    // Lattice!(d2q9): its own small set of kernels and arrays (d2q9.cu), one rank; gn[2] is ignored (0 or 1)
    ARG(cfg->psi_form == DLBC_PSI_LINEAR || cfg->psi_form == DLBC_PSI_EXPONENTIAL, "create: bad psi_form");
    ARG(cfg->eqdist_form >= 0 && cfg->eqdist_form <= 2, "create: bad eqdist_form");
    for (int c = 0; c < cfg->ncomp; c++) ARG(cfg->tau[c] > 0.0 && std::isfinite(cfg->tau[c]), "create: tau must be positive");
    int ndev2 = 0;
    cudaError_t e2 = cudaGetDeviceCount(&ndev2);
    if (e2 != cudaSuccess || ndev2 == 0)
      return set_error(nullptr, std::string("dlbc_cuda_create: no CUDA device (") + cudaGetErrorString(e2) +
                                    "); this library has no CPU fallback", 1000 + (int)e2);
    ARG(cfg->device >= 0 && cfg->device < ndev2, "create: device ordinal out of range");
    d2::Lattice2 *l2 = nullptr;
    const int rc = d2::create(*cfg, &l2, g_create_error);
    if (rc) return rc;
    ctx = new dlbc_cuda_ctx();
    ctx->cfg = *cfg;
    ctx->device = cfg->device;
    ctx->d2 = l2;
    *out = ctx;
    return 0;
  }
  ARG(cfg->nranks >= 1 && cfg->rank >= 0 && cfg->rank < cfg->nranks, "create: bad rank/nranks");
  for (int d = 0; d < 3; d++) ARG(cfg->gn[d] >= 1 && cfg->gn[d] <= 65535, "create: gn out of range (1..65535)");
  // lattice.d:178,233-241: the global size must be divisible by the process grid
  ARG(cfg->gn[2] % cfg->nranks == 0, "create: gn[2] is not divisible by nranks (lattice.d:233-241)");
  ARG(cfg->psi_form == DLBC_PSI_LINEAR || cfg->psi_form == DLBC_PSI_EXPONENTIAL, "create: bad psi_form");
  ARG(cfg->eqdist_form >= 0 && cfg->eqdist_form <= 2, "create: bad eqdist_form");
  ARG(cfg->arithmetic == DLBC_ARITH_FAST || cfg->arithmetic == DLBC_ARITH_EXACT, "create: bad arithmetic mode");
  ARG(cfg->kernel_variant >= DLBC_KERNEL_AUTO && cfg->kernel_variant <= DLBC_KERNEL_TMA, "create: bad kernel_variant");
  for (int c = 0; c < cfg->ncomp; c++) ARG(cfg->tau[c] > 0.0 && std::isfinite(cfg->tau[c]), "create: tau must be positive");

  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return set_error(nullptr, std::string("dlbc_cuda_create: no CUDA device (") + cudaGetErrorString(e) +
                                  "); this library has no CPU fallback", 1000 + (int)e);
  ARG(cfg->device >= 0 && cfg->device < ndev, "create: device ordinal out of range");

  ctx = new dlbc_cuda_ctx();
  ctx->cfg = *cfg;
  ctx->device = cfg->device;
#define CKC(expr)                                             \
  do {                                                        \
    cudaError_t e_ = (expr);                                  \
    if (e_ != cudaSuccess) {                                  \
      cuda_fail(ctx, #expr, e_);                              \
      g_create_error = ctx->err;                              \
      dlbc_cuda_destroy(ctx);                                 \
      return 1000 + (int)e_;                                  \
    }                                                         \
  } while (0)
  CKC(cudaSetDevice(ctx->device));

  Geom &g = ctx->g;
  g.nx = (int)cfg->gn[0];
  g.ny = (int)cfg->gn[1];
  g.nz = (int)(cfg->gn[2] / cfg->nranks);
  g.periodic_z_local = cfg->nranks == 1;
  // per-thread addressing is 32-bit offsets within three consecutive plane blocks (lattice.cuh)
  if ((int64_t)g.nx * g.ny * cfg->ncomp * Q * 3 >= ((int64_t)1 << 31)) {
    g_create_error = "dlbc_cuda: create: nx*ny*ncomp*57 must be below 2^31 (32-bit in-plane addressing)";
    dlbc_cuda_destroy(ctx);
    return 1;
  }
  g.plane = g.nx * g.ny;
  g.fc = Q * g.plane;
  g.fz = cfg->ncomp * Q * g.plane;
  g.ncomp = cfg->ncomp;
  g.sites_g = (int64_t)(g.nz + 2) * g.plane;

  StepParams &p = ctx->p;
  std::memset(&p, 0, sizeof(p));
  p.ncomp = cfg->ncomp;
  for (int c = 0; c < cfg->ncomp; c++) {
    p.tau[c] = cfg->tau[c];
    p.omega[c] = 1.0 / cfg->tau[c];
    p.one_minus_omega[c] = 1.0 - p.omega[c];
    p.inv_tau[c] = p.omega[c];
    p.tau_pow2[c] = is_pow2(cfg->tau[c]);
    p.gwc[c] = cfg->gwc[c];
    if (cfg->gwc[c] != 0.0) p.any_gwc = 1;
    for (int c2 = 0; c2 < cfg->ncomp; c2++) p.gcc[c * MAXC + c2] = cfg->gcc[c * DLBC_CUDA_MAX_COMPONENTS + c2];
  }
  for (int d = 0; d < 3; d++) p.gacc[d] = cfg->global_acc[d];

  // AUTO: the L2-ordered single launch fed by TMA (fastest measured, profiles/README.md), on z-slabs too unless
  // the slab is thin: its two extra boundary launches cost ~0.4 ms per step at 512^2 planes, which the
  // two-kernel schedule wins back below ~128 planes
  ctx->variant_auto = cfg->kernel_variant == DLBC_KERNEL_AUTO;
  ctx->variant = !ctx->variant_auto ? cfg->kernel_variant
                 : (cfg->nranks == 1 || ctx->g.nz >= 160 ? DLBC_KERNEL_WAVE : DLBC_KERNEL_TWO_PASS);
  if (!cfg->enable_shan_chen) ctx->variant = DLBC_KERNEL_TWO_PASS;  // nothing to order: one fused kernel per step

  if (const char *e_l2 = std::getenv("DLBC_L2_PERSIST_MB")) {
    // experiment switch: size of the L2 set-aside for persisting (evict_last) lines; unset = the driver's default
    int max_persist = 0;
    CKC(cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, ctx->device));
    size_t want = (size_t)std::atol(e_l2) << 20;
    if (want > (size_t)max_persist) want = (size_t)max_persist;
    CKC(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want));
    size_t got = 0;
    cudaDeviceGetLimit(&got, cudaLimitPersistingL2CacheSize);
    std::fprintf(stderr, "dlbc_cuda: persisting L2 set-aside %zu MB (max %d MB)\n", got >> 20, max_persist >> 20);
  }
  CKC(cudaStreamCreateWithFlags(&ctx->compute, cudaStreamNonBlocking));
  {
    // the boundary chain of a multi-rank step must overtake the interior kernel's queued blocks
    int prio_lo = 0, prio_hi = 0;
    CKC(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    CKC(cudaStreamCreateWithPriority(&ctx->comm, cudaStreamNonBlocking, prio_hi));
  }
  CKC(cudaStreamCreateWithFlags(&ctx->xfer, cudaStreamNonBlocking));
  CKC(cudaEventCreateWithFlags(&ctx->ev_copy, cudaEventDisableTiming));
  CKC(cudaEventCreateWithFlags(&ctx->ev_stage[0], cudaEventDisableTiming));
  CKC(cudaEventCreateWithFlags(&ctx->ev_stage[1], cudaEventDisableTiming));
  CKC(cudaEventCreate(&ctx->ev_a));
  CKC(cudaEventCreate(&ctx->ev_b));
  CKC(cudaEventCreateWithFlags(&ctx->ev_bnd, cudaEventDisableTiming));
  CKC(cudaEventCreateWithFlags(&ctx->ev_ghost_psi, cudaEventDisableTiming));
  CKC(cudaEventCreateWithFlags(&ctx->ev_ghost_f, cudaEventDisableTiming));

  const size_t fbytes = (size_t)cfg->ncomp * Q * g.sites_g * sizeof(double);
  const size_t sbytes = (size_t)cfg->ncomp * g.sites_g * sizeof(double);
  CKC(cudaMalloc(&ctx->f[0], fbytes));
  CKC(cudaMalloc(&ctx->f[1], fbytes));
  CKC(cudaMalloc(&ctx->psi, sbytes));
  CKC(cudaMalloc(&ctx->mask, (size_t)g.sites_g));
  CKC(cudaMalloc(&ctx->partial, (size_t)(4 * REDUCE_BLOCKS + 4) * sizeof(double)));
  CKC(cudaMalloc(&ctx->flags, 2 * sizeof(int)));
  for (int i = 0; i < (cfg->nranks > 1 ? 3 : 1); i++) {
    CKC(cudaMalloc(&ctx->wave_sync[i], wave_sync_alloc_ints(g) * sizeof(int)));
    ctx->device_bytes += (int64_t)(wave_sync_alloc_ints(g) * sizeof(int));
  }
  ctx->device_bytes += (int64_t)(2 * fbytes + sbytes + g.sites_g + (4 * REDUCE_BLOCKS + 4) * sizeof(double) + 8);
  ctx->tma_maps = make_tma_maps(ctx);
  CKC(cudaMemsetAsync(ctx->f[0], 0, fbytes, ctx->compute));
  CKC(cudaMemsetAsync(ctx->f[1], 0, fbytes, ctx->compute));
  CKC(cudaMemsetAsync(ctx->psi, 0, sbytes, ctx->compute));
  CKC(cudaMemsetAsync(ctx->mask, 0, (size_t)g.sites_g, ctx->compute));
  CKC(cudaEventRecord(ctx->ev_ghost_f, ctx->compute));
  CKC(cudaStreamSynchronize(ctx->compute));

  if (cfg->nranks > 1) {
    const char *miss = nccl_api().load();
    if (miss) {
      g_create_error = miss;
      dlbc_cuda_destroy(ctx);
      return 1;
    }
    ncclUniqueId uid;
    std::memcpy(&uid, cfg->nccl_id, sizeof(uid));
    ncclResult_t r = nccl_api().CommInitRank(&ctx->nccl, cfg->nranks, uid, cfg->rank);
    if (r != ncclSuccess) {
      nccl_fail(ctx, "ncclCommInitRank", r);
      g_create_error = ctx->err;
      ctx->nccl = nullptr;
      dlbc_cuda_destroy(ctx);
      return 2000 + (int)r;
    }
    ctx->up = (cfg->rank + 1) % cfg->nranks;
    ctx->down = (cfg->rank + cfg->nranks - 1) % cfg->nranks;
  }
#undef CKC
  *out = ctx;
  return 0;
}

// ---- state transfer -------------------------------------------------------------------------
int dlbc_cuda_upload_populations_xrange(dlbc_cuda_ctx *ctx, int comp, const double *aos, int halo, int64_t x_begin,
                                        int64_t x_count) {
  TRY(check_ctx(ctx));
  ARG(aos, "upload_populations: null pointer");
  ARG(comp >= 0 && comp < ctx->cfg.ncomp, "upload_populations: component out of range");
  ARG(halo >= 0, "upload_populations: negative halo");
  ARG(x_begin >= 0 && x_count >= 0 && x_begin + x_count <= ctx->g.nx, "upload_populations: x range out of bounds");
  const Geom &g = ctx->g;
  TRY(join_comm(ctx));
  const int64_t xcmax = chunk_layers(ctx, Q, sizeof(double));
  const int64_t half_elems = xcmax * g.ny * g.nz * Q;
  TRY(ensure_stage(ctx, (size_t)2 * half_elems * sizeof(double)));
  // `aos` points at layer x_begin of the host array (the first of the layers moved).
  // Double-buffered: the host -> staging copy of a chunk (xfer stream) overlaps the layout conversion of
  // the previous chunk (compute stream), across calls too.  The call returns when the COPY is done, i.e.
  // when the host buffer may be reused; the conversion kernels stay queued on the compute stream, in
  // front of whatever the host enqueues next.
  const int64_t NY = g.ny + 2 * halo, NZ = g.nz + 2 * halo;
  for (int64_t x0 = x_begin; x0 < x_begin + x_count; x0 += xcmax) {
    const int64_t xc = std::min(xcmax, x_begin + x_count - x0);
    const int h = ctx->stage_half;
    ctx->stage_half ^= 1;
    const double *half = (const double *)ctx->stage + (int64_t)h * half_elems;
    CK(cudaStreamWaitEvent(ctx->xfer, ctx->ev_stage[h], 0));  // the kernel that last read this half is done
    TRY(copy_layers<const double>(ctx, aos + (x0 - x_begin) * NY * NZ * Q, half, halo, Q, xc, true, ctx->xfer));
    CK(cudaEventRecord(ctx->ev_copy, ctx->xfer));
    CK(cudaStreamWaitEvent(ctx->compute, ctx->ev_copy, 0));
    CK(launch_aos_to_soa(g, half, (int)xc, (int)x0, comp, ctx->f[ctx->cur], ctx->f[1 - ctx->cur], ctx->compute));
    ctx->launches++;
    CK(cudaEventRecord(ctx->ev_stage[h], ctx->compute));
    static const char *e_ov = std::getenv("DLBC_UPLOAD_OVERLAP");
    if (e_ov && e_ov[0] == '0') CK(cudaStreamSynchronize(ctx->compute));  // measurement: no overlap
    else CK(cudaEventSynchronize(ctx->ev_copy));
  }
  ctx->ghosts_valid = false;
  ctx->rho_wall_valid = false;
  ctx->force_valid = false;
  return 0;
}

int dlbc_cuda_upload_populations(dlbc_cuda_ctx *ctx, int comp, const double *aos, int halo) {
  D2(upload_populations(ctx->d2, comp, aos, halo, ctx->err));
  TRY(check_ctx(ctx));
  ARG(aos, "upload_populations: null pointer");
  ARG(halo >= 0, "upload_populations: negative halo");
  // the full-array form receives Field.arr._data.ptr, i.e. the address of layer -halo
  const int64_t NY = ctx->g.ny + 2 * halo, NZ = ctx->g.nz + 2 * halo;
  return dlbc_cuda_upload_populations_xrange(ctx, comp, aos + (int64_t)halo * NY * NZ * Q, halo, 0, ctx->g.nx);
}

int dlbc_cuda_download_populations_xrange(dlbc_cuda_ctx *ctx, int comp, double *aos, int halo, int64_t x_begin,
                                          int64_t x_count) {
  TRY(check_ctx(ctx));
  ARG(aos, "download_populations: null pointer");
  ARG(comp >= 0 && comp < ctx->cfg.ncomp, "download_populations: component out of range");
  ARG(halo >= 0, "download_populations: negative halo");
  ARG(x_begin >= 0 && x_count >= 0 && x_begin + x_count <= ctx->g.nx, "download_populations: x range out of bounds");
  const Geom &g = ctx->g;
  TRY(join_comm(ctx));
  const int64_t xcmax = chunk_layers(ctx, Q, sizeof(double));
  TRY(ensure_stage(ctx, (size_t)xcmax * g.ny * g.nz * Q * sizeof(double)));
  const int64_t NY = g.ny + 2 * halo, NZ = g.nz + 2 * halo;
  for (int64_t x0 = x_begin; x0 < x_begin + x_count; x0 += xcmax) {
    const int64_t xc = std::min(xcmax, x_begin + x_count - x0);
    CK(launch_soa_to_aos(g, ctx->f[ctx->cur], (int)xc, (int)x0, comp, (double *)ctx->stage, ctx->compute));
    ctx->launches++;
    TRY(copy_layers<double>(ctx, aos + (x0 - x_begin) * NY * NZ * Q, (double *)ctx->stage, halo, Q, xc, false));
    CK(cudaStreamSynchronize(ctx->compute));
  }
  return 0;
}

int dlbc_cuda_download_populations(dlbc_cuda_ctx *ctx, int comp, double *aos, int halo) {
  D2(download_populations(ctx->d2, comp, aos, halo, ctx->err));
  TRY(check_ctx(ctx));
  ARG(aos, "download_populations: null pointer");
  ARG(halo >= 0, "download_populations: negative halo");
  const int64_t NY = ctx->g.ny + 2 * halo, NZ = ctx->g.nz + 2 * halo;
  return dlbc_cuda_download_populations_xrange(ctx, comp, aos + (int64_t)halo * NY * NZ * Q, halo, 0, ctx->g.nx);
}

int dlbc_cuda_upload_mask(dlbc_cuda_ctx *ctx, const int32_t *mask, int halo) {
  D2(upload_mask(ctx->d2, mask, halo, ctx->err));
  TRY(check_ctx(ctx));
  ARG(mask, "upload_mask: null pointer");
  ARG(halo >= 0, "upload_mask: negative halo");
  const Geom &g = ctx->g;
  TRY(join_comm(ctx));
  TRY(ensure_stage(ctx, (size_t)g.nx * g.ny * g.nz * sizeof(int32_t)));
  const int64_t NY = g.ny + 2 * halo, NZ = g.nz + 2 * halo;
  TRY(copy_layers<const int32_t>(ctx, mask + (int64_t)halo * NY * NZ, (const int32_t *)ctx->stage, halo, 1, g.nx, true));
  CK(cudaMemsetAsync(ctx->flags, 0, 2 * sizeof(int), ctx->compute));
  CK(launch_mask_in(g, (const int32_t *)ctx->stage, ctx->mask, ctx->flags, ctx->compute));
  ctx->launches++;
  int flags[2] = {0, 0};
  CK(cudaMemcpyAsync(flags, ctx->flags, sizeof(flags), cudaMemcpyDeviceToHost, ctx->compute));
  CK(cudaStreamSynchronize(ctx->compute));
  ARG(flags[0] == 0, "upload_mask: mask values must be 0 (None) or 1 (Solid) (mask.d:84-93)");
  TRY(finish_mask(ctx, flags[1]));
  return release_stage(ctx);
}

int dlbc_cuda_download_mask(dlbc_cuda_ctx *ctx, int32_t *mask, int halo) {
  D2(download_mask(ctx->d2, mask, halo, ctx->err));
  TRY(check_ctx(ctx));
  ARG(mask, "download_mask: null pointer");
  ARG(halo >= 0, "download_mask: negative halo");
  const Geom &g = ctx->g;
  TRY(ensure_stage(ctx, (size_t)g.nx * g.ny * g.nz * sizeof(int32_t)));
  CK(launch_mask_out(g, ctx->mask, (int32_t *)ctx->stage, ctx->compute));
  ctx->launches++;
  const int64_t NY = g.ny + 2 * halo, NZ = g.nz + 2 * halo;
  TRY(copy_layers<int32_t>(ctx, mask + (int64_t)halo * NY * NZ, (int32_t *)ctx->stage, halo, 1, g.nx, false));
  CK(cudaStreamSynchronize(ctx->compute));
  return release_stage(ctx);
}

// ---- the hot path ---------------------------------------------------------------------------
int dlbc_cuda_step(dlbc_cuda_ctx *ctx, int nsteps) {
  if (ctx && ctx->d2) {
    cudaError_t e2 = cudaSetDevice(ctx->device);
    if (e2 != cudaSuccess) return cuda_fail(ctx, "cudaSetDevice", e2);
    const int rc = d2::step(ctx->d2, nsteps, ctx->err);
    if (rc == 0) ctx->timestep += nsteps;
    return rc;
  }
  TRY(check_ctx(ctx));
  ARG(nsteps >= 0, "step: negative step count");
  if (nsteps == 0) return 0;
  ARG(effective_variant(ctx) != DLBC_KERNEL_TMA || tma_ok(ctx),
      "step: the TMA kernel needs an even nx, at most 2 components and no solid sites; use AUTO or WAVE");
  TRY(ensure_ghosts(ctx));
  if (ctx->cfg.nranks > 1 && ctx->timestep == 0) CK(cudaEventRecord(ctx->ev_ghost_f, ctx->compute));
  for (int i = 0; i < nsteps; i++) TRY(step_once(ctx, nullptr));
  return 0;
}

int dlbc_cuda_sync(dlbc_cuda_ctx *ctx) {
  D2(sync(ctx->d2, ctx->err));
  TRY(check_ctx(ctx));
  CK(cudaStreamSynchronize(ctx->compute));
  CK(cudaStreamSynchronize(ctx->comm));
  CK(cudaGetLastError());
  return 0;
}

int64_t dlbc_cuda_timestep(const dlbc_cuda_ctx *ctx) { return ctx ? ctx->timestep : -1; }

// ---- outputs --------------------------------------------------------------------------------
int dlbc_cuda_download_density(dlbc_cuda_ctx *ctx, int comp, double *out, int halo) {
  D2(download_field(ctx->d2, d2::F2_DENSITY, comp, comp, out, halo, ctx->err));
  TRY(check_ctx(ctx));
  ARG(out && halo >= 0, "download_density: bad argument");
  ARG(comp >= 0 && comp < ctx->cfg.ncomp, "download_density: component out of range");
  TRY(join_comm(ctx));
  return download_field(ctx, 0, comp, comp, nullptr, out, halo);
}
int dlbc_cuda_download_colour(dlbc_cuda_ctx *ctx, int comp1, int comp2, double *out, int halo) {
  D2(download_field(ctx->d2, d2::F2_COLOUR, comp1, comp2, out, halo, ctx->err));
  TRY(check_ctx(ctx));
  ARG(out && halo >= 0, "download_colour: bad argument");
  ARG(comp1 >= 0 && comp1 < ctx->cfg.ncomp && comp2 >= 0 && comp2 < ctx->cfg.ncomp, "download_colour: component out of range");
  TRY(join_comm(ctx));
  return download_field(ctx, 1, comp1, comp2, nullptr, out, halo);
}
int dlbc_cuda_download_velocity(dlbc_cuda_ctx *ctx, int comp, double *out, int halo) {
  D2(download_field(ctx->d2, d2::F2_VELOCITY, comp, comp, out, halo, ctx->err));
  TRY(check_ctx(ctx));
  ARG(out && halo >= 0, "download_velocity: bad argument");
  ARG(comp >= 0 && comp < ctx->cfg.ncomp, "download_velocity: component out of range");
  TRY(join_comm(ctx));
  return download_field(ctx, 2, comp, comp, nullptr, out, halo);
}

int dlbc_cuda_download_force(dlbc_cuda_ctx *ctx, int comp, double *out, int halo) {
  D2(download_field(ctx->d2, d2::F2_FORCE, comp, comp, out, halo, ctx->err));
  TRY(check_ctx(ctx));
  ARG(out && halo >= 0, "download_force: bad argument");
  ARG(comp >= 0 && comp < ctx->cfg.ncomp, "download_force: component out of range");
  TRY(join_comm(ctx));
  double *vec = nullptr;
  TRY(last_force(ctx, comp, &vec));
  int rc = download_field(ctx, 3, comp, comp, vec, out, halo);
  cudaStreamSynchronize(ctx->compute);
  cudaFree(vec);
  return rc;
}

// pressure (density.d:306-326) of the current state, every site
int dlbc_cuda_download_pressure(dlbc_cuda_ctx *ctx, double *out, int halo) {
  TRY(check_ctx(ctx));
  ARG(out && halo >= 0, "download_pressure: bad argument");
  const Geom &g = ctx->g;
  TRY(join_comm(ctx));
  TRY(ensure_stage(ctx, (size_t)g.nx * g.ny * g.nz * sizeof(double)));
  CK(launch_pressure(g, ctx->f[ctx->cur], ctx->p, ctx->cfg.psi_form, (double *)ctx->stage, ctx->compute));
  ctx->launches++;
  const int64_t NY = g.ny + 2 * halo, NZ = g.nz + 2 * halo;
  TRY(copy_layers<double>(ctx, out + (int64_t)halo * NY * NZ, (double *)ctx->stage, halo, 1, g.nx, false));
  CK(cudaStreamSynchronize(ctx->compute));
  return release_stage(ctx);
}

// corrected velocity (velocity.d:177-185): raw velocity + tau * F / rho with the force of the last step
int dlbc_cuda_download_corrected_velocity(dlbc_cuda_ctx *ctx, int comp, double *out, int halo) {
  TRY(check_ctx(ctx));
  ARG(out && halo >= 0, "download_corrected_velocity: bad argument");
  ARG(comp >= 0 && comp < ctx->cfg.ncomp, "download_corrected_velocity: component out of range");
  const Geom &g = ctx->g;
  TRY(join_comm(ctx));
  double *vec = nullptr;
  TRY(last_force(ctx, comp, &vec));
  int rc = ensure_stage(ctx, (size_t)g.nx * g.ny * g.nz * 3 * sizeof(double));
  if (!rc) {
    cudaError_t e = launch_corrected_velocity(g, ctx->f[ctx->cur], ctx->mask, comp, ctx->cfg.tau[comp], vec,
                                              (double *)ctx->stage, ctx->compute);
    ctx->launches++;
    if (e != cudaSuccess) rc = cuda_fail(ctx, "corrected_velocity_kernel", e);
  }
  const int64_t NY = g.ny + 2 * halo, NZ = g.nz + 2 * halo;
  if (!rc) rc = copy_layers<double>(ctx, out + (int64_t)halo * NY * NZ * 3, (double *)ctx->stage, halo, 3, g.nx, false);
  cudaStreamSynchronize(ctx->compute);
  cudaFree(vec);
  if (rc) return rc;
  return release_stage(ctx);
}

// volumeAveragedDensity (laplace.d:236-272): mean density of every component over the fluid sites of the cube
// offset +- half_width (global coordinates, not wrapped: the reference's documented behaviour)
int dlbc_cuda_volume_averaged_density(dlbc_cuda_ctx *ctx, const int64_t offset[3], int half_width, double *out) {
  TRY(check_ctx(ctx));
  ARG(offset && out && half_width >= 0, "volume_averaged_density: bad argument");
  const Geom &g = ctx->g;
  const int C = ctx->cfg.ncomp;
  TRY(join_comm(ctx));
  int64_t lo[3], hi[3];
  for (int d = 0; d < 3; d++) {
    lo[d] = offset[d] - half_width;
    hi[d] = offset[d] + half_width;
  }
  double *res = ctx->partial;  // [2 C] of the reduction scratch
  CK(launch_cube_density(g, ctx->f[ctx->cur], ctx->mask, C, (int64_t)ctx->cfg.rank * g.nz, lo, hi, res, ctx->compute));
  ctx->launches++;
  if (ctx->cfg.nranks > 1) NK(nccl_api().AllReduce(res, res, 2 * C, ncclDouble, ncclSum, ctx->nccl, ctx->compute));
  double h[2 * DLBC_CUDA_MAX_COMPONENTS];
  CK(cudaMemcpyAsync(h, res, 2 * C * sizeof(double), cudaMemcpyDeviceToHost, ctx->compute));
  CK(cudaStreamSynchronize(ctx->compute));
  for (int c = 0; c < C; c++) out[c] = h[2 * c] / h[2 * c + 1];  // (0 / 0 = NaN for an empty cube, as the reference)
  return 0;
}

// ---- initial conditions and geometry generated on the device ----------------------------
int dlbc_cuda_init_mask(dlbc_cuda_ctx *ctx, int kind, int axis, int wall_offset, int thickness, const int32_t has_wall[6]) {
  TRY(check_ctx(ctx));
  ARG(kind >= DLBC_MASK_NONE && kind <= DLBC_MASK_WALLS, "init_mask: bad kind");
  ARG(axis >= 0 && axis < 3, "init_mask: axis out of range (mask.d:161-163)");
  ARG(kind != DLBC_MASK_WALLS || has_wall, "init_mask: Walls needs dimensionHasWall");
  const Geom &g = ctx->g;
  TRY(join_comm(ctx));
  int hw[6] = {0, 0, 0, 0, 0, 0};
  if (has_wall)
    for (int i = 0; i < 6; i++) hw[i] = has_wall[i] != 0;
  CK(cudaMemsetAsync(ctx->flags, 0, 2 * sizeof(int), ctx->compute));
  CK(launch_init_mask(g, kind, axis, wall_offset, thickness, hw, ctx->cfg.gn, (int64_t)ctx->cfg.rank * g.nz, ctx->mask,
                      ctx->flags + 1, ctx->compute));
  ctx->launches++;
  int solid = 0;
  CK(cudaMemcpyAsync(&solid, ctx->flags + 1, sizeof(int), cudaMemcpyDeviceToHost, ctx->compute));
  CK(cudaStreamSynchronize(ctx->compute));
  return finish_mask(ctx, solid);
}

int dlbc_cuda_init_populations(dlbc_cuda_ctx *ctx, int comp, int kind, double density, double perturb, uint32_t seed,
                               int64_t skip) {
  TRY(check_ctx(ctx));
  ARG(comp >= 0 && comp < ctx->cfg.ncomp, "init_populations: component out of range");
  ARG(kind >= DLBC_INIT_EQDIST && kind <= DLBC_INIT_EQDIST_PERTURB_FRAC, "init_populations: bad kind");
  ARG(skip >= 0, "init_populations: negative skip");
  const Geom &g = ctx->g;
  TRY(join_comm(ctx));
  double eq[Q];
  eq_dist_unity(ctx->cfg.eqdist_form, eq);
  uint32_t *raw = nullptr;
  double a = 0.0, b = 0.0;
  if (kind != DLBC_INIT_EQDIST) {
    // one draw per site of the rank's full array incl. its halo of 2, in memory order (byElementForward)
    const int64_t n = (int64_t)(g.nx + 4) * (g.ny + 4) * (g.nz + 4);
    TRY(ensure_stage(ctx, (size_t)n * sizeof(uint32_t)));
    raw = (uint32_t *)ctx->stage;
    CK(launch_mt19937(seed, skip, n, raw, ctx->compute));
    ctx->launches++;
    if (kind == DLBC_INIT_EQDIST_RANDOM) {
      a = 0.0;
      b = 2.0;
    } else {
      a = -perturb;
      b = perturb;
    }
  }
  CK(launch_init_populations(g, kind, density, a, std::nextafter(b, -INFINITY), eq, raw, comp, ctx->f[ctx->cur],
                             ctx->f[1 - ctx->cur], ctx->compute));
  ctx->launches++;
  CK(cudaStreamSynchronize(ctx->compute));
  ctx->ghosts_valid = false;
  ctx->rho_wall_valid = false;
  ctx->force_valid = false;
  return release_stage(ctx);
}

int dlbc_cuda_init_spheres(dlbc_cuda_ctx *ctx, int comp, int ncentres, const double *centres, int skip_axis, double radius,
                           double thickness, double density1, double density2) {
  TRY(check_ctx(ctx));
  ARG(comp >= 0 && comp < ctx->cfg.ncomp, "init_spheres: component out of range");
  ARG(centres && (ncentres == 1 || ncentres == 2), "init_spheres: one or two centres");
  ARG(skip_axis >= -1 && skip_axis < 3, "init_spheres: bad axis");
  TRY(join_comm(ctx));
  double eq[Q];
  eq_dist_unity(ctx->cfg.eqdist_form, eq);
  CK(launch_init_profile(ctx->g, 0, ncentres, centres, skip_axis, 0, radius, thickness, density1, density2, nullptr,
                         (int64_t)ctx->cfg.rank * ctx->g.nz, eq, comp, ctx->f[ctx->cur], ctx->f[1 - ctx->cur], ctx->compute));
  ctx->launches++;
  ctx->ghosts_valid = false;
  ctx->rho_wall_valid = false;
  ctx->force_valid = false;
  return 0;
}

int dlbc_cuda_init_axis_profile(dlbc_cuda_ctx *ctx, int comp, int axis, const double *values) {
  TRY(check_ctx(ctx));
  ARG(comp >= 0 && comp < ctx->cfg.ncomp, "init_axis_profile: component out of range");
  ARG(axis >= 0 && axis < 3 && values, "init_axis_profile: bad argument");
  TRY(join_comm(ctx));
  double eq[Q];
  eq_dist_unity(ctx->cfg.eqdist_form, eq);
  const size_t bytes = (size_t)ctx->cfg.gn[axis] * sizeof(double);
  TRY(ensure_stage(ctx, bytes));
  CK(cudaMemcpyAsync(ctx->stage, values, bytes, cudaMemcpyHostToDevice, ctx->compute));
  CK(launch_init_profile(ctx->g, 1, 0, nullptr, -1, axis, 0.0, 0.0, 0.0, 0.0, (const double *)ctx->stage,
                         (int64_t)ctx->cfg.rank * ctx->g.nz, eq, comp, ctx->f[ctx->cur], ctx->f[1 - ctx->cur], ctx->compute));
  ctx->launches++;
  CK(cudaStreamSynchronize(ctx->compute));
  ctx->ghosts_valid = false;
  ctx->rho_wall_valid = false;
  ctx->force_valid = false;
  return 0;
}

int dlbc_cuda_init_wall(dlbc_cuda_ctx *ctx) {
  TRY(check_ctx(ctx));
  TRY(join_comm(ctx));
  if (ctx->has_solid) {
    CK(launch_init_wall(ctx->g, ctx->mask, ctx->f[0], ctx->f[1], ctx->compute));
    ctx->launches++;
  }
  ctx->rho_wall_valid = false;
  ctx->force_valid = false;
  return 0;
}

static int reduce_component(dlbc_cuda_ctx *ctx, int comp, double out4[4]) {
  const Geom &g = ctx->g;
  TRY(join_comm(ctx));
  CK(launch_reduce(g, ctx->f[ctx->cur], comp, ctx->mask, ctx->partial, ctx->compute));
  ctx->launches += 2;
  double *res = ctx->partial + 4 * REDUCE_BLOCKS;
  if (ctx->cfg.nranks > 1) NK(nccl_api().AllReduce(res, res, 4, ncclDouble, ncclSum, ctx->nccl, ctx->compute));
  CK(cudaMemcpyAsync(out4, res, 4 * sizeof(double), cudaMemcpyDeviceToHost, ctx->compute));
  CK(cudaStreamSynchronize(ctx->compute));
  return 0;
}

int dlbc_cuda_total_mass(dlbc_cuda_ctx *ctx, int comp, double *out) {
  if (ctx && ctx->d2) {
    double r2[4];
    if (!out) return set_error(ctx, "dlbc_cuda: total_mass: null pointer", 1);
    cudaSetDevice(ctx->device);
    const int rc = d2::reduce(ctx->d2, comp, r2, ctx->err);
    if (rc == 0) *out = r2[0];
    return rc;
  }
  TRY(check_ctx(ctx));
  ARG(out, "total_mass: null pointer");
  ARG(comp >= 0 && comp < ctx->cfg.ncomp, "total_mass: component out of range");
  double r[4];
  TRY(reduce_component(ctx, comp, r));
  *out = r[0];
  return 0;
}
int dlbc_cuda_total_momentum(dlbc_cuda_ctx *ctx, int comp, double out[3]) {
  if (ctx && ctx->d2) {
    double r2[4];
    if (!out) return set_error(ctx, "dlbc_cuda: total_momentum: null pointer", 1);
    cudaSetDevice(ctx->device);
    const int rc = d2::reduce(ctx->d2, comp, r2, ctx->err);
    if (rc == 0) { out[0] = r2[1]; out[1] = r2[2]; out[2] = 0.0; }
    return rc;
  }
  TRY(check_ctx(ctx));
  ARG(out, "total_momentum: null pointer");
  ARG(comp >= 0 && comp < ctx->cfg.ncomp, "total_momentum: component out of range");
  double r[4];
  TRY(reduce_component(ctx, comp, r));
  out[0] = r[1]; out[1] = r[2]; out[2] = r[3];
  return 0;
}

// ---- introspection / measurement ------------------------------------------------------------
int dlbc_cuda_tables(int64_t *velocities, int64_t *bounce_out, double *weights, double *css) {
  dlbc_cuda_ctx *ctx = nullptr;
  ARG(velocities && bounce_out && weights && css, "tables: null pointer");
  for (int q = 0; q < Q; q++) {
    velocities[3 * q + 0] = cx(q);
    velocities[3 * q + 1] = cy(q);
    velocities[3 * q + 2] = cz(q);
    bounce_out[q] = bounce(q);
    weights[q] = weight(q);
  }
  *css = 1.0 / 3.0;
  return 0;
}

int dlbc_cuda_site_function(int device, int fn, int form, int n, const double *in, const double *in2, double *out) {
  dlbc_cuda_ctx *ctx = nullptr;
  ARG(fn >= DLBC_SITE_DENSITY && fn <= DLBC_SITE_EQDIST && n >= 0 && in && out, "site_function: bad argument");
  ARG(fn != DLBC_SITE_EQDIST || in2, "site_function: the equilibrium needs the velocity shifts dv[n][3]");
  ARG(form >= 0 && form <= 2 && (fn != DLBC_SITE_PSI || form <= 1), "site_function: bad form");
  if (n == 0) return 0;
  CK(cudaSetDevice(device));
  const size_t nin = (size_t)n * (fn == DLBC_SITE_PSI ? 1 : Q), nout = (size_t)n * (fn == DLBC_SITE_VELOCITY ? 3 : fn == DLBC_SITE_EQDIST ? Q : 1);
  const size_t nin2 = !in2 ? 0 : (size_t)n * (fn == DLBC_SITE_EQDIST ? 3 : 1);
  double *d = nullptr;
  CK(cudaMalloc(&d, (nin + nin2 + nout) * sizeof(double)));
  cudaError_t e = cudaMemcpy(d, in, nin * sizeof(double), cudaMemcpyHostToDevice);
  if (e == cudaSuccess && nin2) e = cudaMemcpy(d + nin, in2, nin2 * sizeof(double), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = launch_site_function(fn, form, n, d, nin2 ? d + nin : nullptr, d + nin + nin2, nullptr);
  if (e == cudaSuccess) e = cudaMemcpy(out, d + nin + nin2, nout * sizeof(double), cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (e != cudaSuccess) return cuda_fail(nullptr, "site_function", e);
  return 0;
}

int dlbc_cuda_launch_count(const dlbc_cuda_ctx *ctx, int64_t *count) {
  if (!ctx || !count) return 1;
  *count = ctx->d2 ? d2::launches(ctx->d2) : ctx->launches;
  return 0;
}

int dlbc_cuda_device_bytes(const dlbc_cuda_ctx *ctx, int64_t *bytes) {
  if (!ctx || !bytes) return 1;
  *bytes = ctx->d2 ? d2::device_bytes(ctx->d2) : ctx->device_bytes;
  return 0;
}

int dlbc_cuda_kernel_variant(const dlbc_cuda_ctx *ctx, int *variant) {
  if (!ctx || !variant) return 1;
  *variant = ctx->d2 ? DLBC_KERNEL_TWO_PASS : effective_variant(ctx);
  return 0;
}

int dlbc_cuda_set_option(dlbc_cuda_ctx *ctx, const char *name, int64_t value) {
  TRY(check_ctx(ctx));
  ARG(name, "set_option: null name");
  ARG(value >= 0 && value <= 65535, "set_option: value out of range");
  if (!std::strcmp(name, "wave_tyb")) ctx->opt_wave_tyb = (int)value;
  else if (!std::strcmp(name, "wave_la")) {
    ARG(value == 0 || value >= 2, "set_option: wave_la must be >= 2 (a fused item reads psi of the next plane)");
    ctx->opt_wave_la = (int)value;
  } else return set_error(ctx, std::string("dlbc_cuda: set_option: unknown option ") + name, 1);
  return 0;
}

int dlbc_cuda_timed_steps(dlbc_cuda_ctx *ctx, int nsteps, double *elapsed_ms) {
  TRY(check_ctx(ctx));
  ARG(nsteps >= 0 && elapsed_ms, "timed_steps: bad argument");
  ARG(effective_variant(ctx) != DLBC_KERNEL_TMA || tma_ok(ctx), "timed_steps: the TMA kernel needs an even nx, at most 2 components and no solid sites");
  TRY(ensure_ghosts(ctx));
  if (ctx->cfg.nranks > 1 && ctx->timestep == 0) CK(cudaEventRecord(ctx->ev_ghost_f, ctx->compute));
  TRY(join_comm(ctx));
  CK(cudaStreamSynchronize(ctx->compute));
  CK(cudaEventRecord(ctx->ev_a, ctx->compute));
  for (int i = 0; i < nsteps; i++) TRY(step_once(ctx, nullptr));
  TRY(join_comm(ctx));
  CK(cudaEventRecord(ctx->ev_b, ctx->compute));
  CK(cudaEventSynchronize(ctx->ev_b));
  float ms = 0.f;
  CK(cudaEventElapsedTime(&ms, ctx->ev_a, ctx->ev_b));
  *elapsed_ms = (double)ms;
  return 0;
}

int dlbc_cuda_kernel_times(dlbc_cuda_ctx *ctx, int cap, const char **names, double *ms, int *n) {
  TRY(check_ctx(ctx));
  ARG(names && ms && n && cap >= 0, "kernel_times: bad argument");
  ARG(ctx->cfg.nranks == 1, "kernel_times: single-rank contexts only");
  TRY(ensure_ghosts(ctx));
  KernelTimer kt{ctx, true, {}, {}};
  TRY(step_once(ctx, &kt));
  CK(cudaStreamSynchronize(ctx->compute));
  *n = 0;
  for (size_t i = 0; i < kt.names.size() && (int)i < cap; i++) {
    float t = 0.f;
    CK(cudaEventElapsedTime(&t, kt.evs[2 * i], kt.evs[2 * i + 1]));
    names[i] = kt.names[i];
    ms[i] = (double)t;
    (*n)++;
  }
  return 0;
}

}  // extern "C"

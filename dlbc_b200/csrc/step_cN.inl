// step_cN.inl -- instantiates the step kernels for DLBC_NC components in arithmetic mode
// DLBC_EXACT (1: translation unit built with -fmad=false; 0: FMA contraction allowed).
// Included by step_c<N>_{exact,fast}.cu.  The psi and force kernels (adds, multiplies in
// reference order, exp) live in the EXACT unit only.
#include "kernels.h"
#include "step_two_pass.cuh"
#include "step_wave.cuh"
#include "step_tma.cuh"
#include <cstdlib>

namespace dlbc {

#define DLBC_CAT2(a, b) a##b
#define DLBC_CAT(a, b) DLBC_CAT2(a, b)

static inline dim3 block_for(const Geom &g) { return dim3(g.nx >= 128 ? 128 : ((g.nx + 31) / 32) * 32); }
static inline dim3 grid_for(const Geom &g, dim3 b, int z_count) { return dim3((g.nx + b.x - 1) / b.x, g.ny, z_count); }

#if DLBC_EXACT
cudaError_t DLBC_CAT(launch_psi_c, DLBC_NC)(const StepLaunch &L, int z_begin, int z_count) {
  constexpr int C = DLBC_NC;
  if (z_count <= 0) return cudaSuccess;
  const dim3 b = block_for(L.g), gr = grid_for(L.g, b, z_count);
#define PSI_CASE(PSI, MASK) psi_kernel<C, PSI, MASK><<<gr, b, 0, L.stream>>>(L.g, L.f, z_begin)
  if (L.psi == 0) { if (L.mask) PSI_CASE(0, true); else PSI_CASE(0, false); }
  else            { if (L.mask) PSI_CASE(1, true); else PSI_CASE(1, false); }
#undef PSI_CASE
  return cudaGetLastError();
}

cudaError_t DLBC_CAT(launch_force_c, DLBC_NC)(const StepLaunch &L, int comp, double *out, int z_begin, int z_count) {
  constexpr int C = DLBC_NC;
  if (z_count <= 0) return cudaSuccess;
  const dim3 b = block_for(L.g), gr = grid_for(L.g, b, z_count);
  if (L.mask) force_kernel<C, true><<<gr, b, 0, L.stream>>>(L.g, L.f, L.p, comp, out, z_begin);
  else        force_kernel<C, false><<<gr, b, 0, L.stream>>>(L.g, L.f, L.p, comp, out, z_begin);
  return cudaGetLastError();
}
#define DLBC_MODE_NAME(base) DLBC_CAT(DLBC_CAT(base, _exact_c), DLBC_NC)
constexpr bool kExact = true;
#else
#define DLBC_MODE_NAME(base) DLBC_CAT(DLBC_CAT(base, _fast_c), DLBC_NC)
constexpr bool kExact = false;
#endif

template <int C, int EQ, bool SC>
static void fused_dispatch(const StepLaunch &L, dim3 gr, dim3 b, int z_begin) {
  if (L.mask) fused_kernel<C, EQ, SC, true, kExact><<<gr, b, 0, L.stream>>>(L.g, L.f, L.p, z_begin);
  else        fused_kernel<C, EQ, SC, false, kExact><<<gr, b, 0, L.stream>>>(L.g, L.f, L.p, z_begin);
}

cudaError_t DLBC_MODE_NAME(launch_fused)(const StepLaunch &L, int z_begin, int z_count) {
  constexpr int C = DLBC_NC;
  if (z_count <= 0) return cudaSuccess;
  const dim3 b = block_for(L.g), gr = grid_for(L.g, b, z_count);
  switch (L.eq * 2 + (L.sc ? 1 : 0)) {
    case 0: fused_dispatch<C, 0, false>(L, gr, b, z_begin); break;
    case 1: fused_dispatch<C, 0, true>(L, gr, b, z_begin); break;
    case 2: fused_dispatch<C, 1, false>(L, gr, b, z_begin); break;
    case 3: fused_dispatch<C, 1, true>(L, gr, b, z_begin); break;
    case 4: fused_dispatch<C, 2, false>(L, gr, b, z_begin); break;
    case 5: fused_dispatch<C, 2, true>(L, gr, b, z_begin); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

// ---- L2-ordered single launch: schedule shared by the register path (step_wave.cuh) and the TMA path
// (step_tma.cuh); `seg` = sites per item ----
static cudaError_t wave_plan(const StepLaunch &L, int z_begin, int z_count, int seg, int alt, Wave &w) {
  const Geom &g = L.g;
  if (g.periodic_z_local && !(z_begin == 1 && z_count == g.nz)) return cudaErrorInvalidValue;
  w.z_begin = z_begin;
  w.z_end = z_begin + z_count;
  w.pre = g.periodic_z_local ? 1 : 0;
  w.pl = g.periodic_z_local ? 1 : (z_begin - 1 < 1 ? 1 : z_begin - 1);
  w.ph = g.periodic_z_local ? g.nz - 1 : (w.z_end > g.nz ? g.nz : w.z_end);
  // band-plane of ~16k sites (5 MB of populations at C = 2): measured best on 512^2 planes (profiles/README.md)
  w.tyb = L.wave_tyb > 0 ? L.wave_tyb : (int)(16384 / g.nx);
  if (w.tyb < 8 && L.wave_tyb <= 0) w.tyb = 8;
  if (w.tyb > g.ny) w.tyb = g.ny;
  w.seg = seg;
  w.xb = (g.nx + w.seg - 1) / w.seg;
  w.nsub = 1;
  w.alt = alt;
  // lag of F behind P: at least 2 slots (F(z) reads psi(z+1)); more when a band-plane is smaller than
  // the ~64k sites the GPU keeps in flight, so that F items rarely have to wait (512^2 planes, TYB = 32:
  // LA 4 measured best, profiles/README.md)
  int la = L.wave_la > 0 ? L.wave_la : (int)((65536 + (int64_t)w.tyb * g.nx - 1) / ((int64_t)w.tyb * g.nx));
  if (la < 2) la = 2;
  if (L.wave_la <= 0 && la > 8) la = 8;
  w.la = la;
  w.rp = (w.tyb + 2) * w.xb;
  w.rf = w.tyb * w.xb;
  // F(zf) runs in slot zf - pl + 1 + la (step_wave.cuh); the last one is zf = z_end - 1
  const int np = w.ph - w.pl + 1, nf = (w.z_end - 1 - w.pl + 1) + w.la;
  w.slots = 1 + (np > nf ? np : nf);
  w.nbands = (g.ny + w.tyb - 1) / w.tyb;
  w.sync = L.wave_sync;
  if (!w.sync) return cudaErrorInvalidValue;
  const int64_t items = (int64_t)w.nbands * w.slots * (w.rp + w.rf);
  if (items >= ((int64_t)1 << 31)) return cudaErrorInvalidConfiguration;
  return cudaMemsetAsync(w.sync, 0, wave_sync_ints(g, w.tyb) * sizeof(int), L.stream);
}

static int device_slot() {
  int dev = 0;
  cudaGetDevice(&dev);
  return dev >= 0 && dev < 64 ? dev : 0;
}

template <int C, int EQ, int PSI, bool MASK>
static cudaError_t wave_launch(const StepLaunch &L, int z_begin, int z_count) {
  Wave w;
  cudaError_t e = wave_plan(L, z_begin, z_count, WAVE_SEG, 0, w);
  if (e != cudaSuccess) return e;
  // persistent grid: as many blocks as can be resident (fewer when the launch is small); per device
  static int resident[64] = {};
  const int dev = device_slot();
  if (!resident[dev]) {
    int sms = 0, per_sm = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, wave_kernel<C, EQ, PSI, MASK, kExact>, WAVE_SEG, 0);
    resident[dev] = sms * (per_sm > 0 ? per_sm : 1);
  }
  const int64_t items = (int64_t)w.nbands * w.slots * (w.rp + w.rf);
  const unsigned grid = (unsigned)(items < resident[dev] ? items : resident[dev]);
  wave_kernel<C, EQ, PSI, MASK, kExact><<<grid, WAVE_SEG, 0, L.stream>>>(L.g, L.f, L.p, w);
  return cudaGetLastError();
}

template <int C, int EQ>
static cudaError_t wave_eq(const StepLaunch &L, int z_begin, int z_count) {
  if (L.psi == 0) return L.mask ? wave_launch<C, EQ, 0, true>(L, z_begin, z_count) : wave_launch<C, EQ, 0, false>(L, z_begin, z_count);
  return L.mask ? wave_launch<C, EQ, 1, true>(L, z_begin, z_count) : wave_launch<C, EQ, 1, false>(L, z_begin, z_count);
}

cudaError_t DLBC_MODE_NAME(launch_wave)(const StepLaunch &L, int z_begin, int z_count) {
  if (z_count <= 0) return cudaSuccess;
  if (!L.sc) return cudaErrorNotSupported;
  switch (L.eq) {
    case 0: return wave_eq<DLBC_NC, 0>(L, z_begin, z_count);
    case 1: return wave_eq<DLBC_NC, 1>(L, z_begin, z_count);
    case 2: return wave_eq<DLBC_NC, 2>(L, z_begin, z_count);
    default: return cudaErrorInvalidValue;
  }
}

// ---- the same schedule fed by tensor-map TMA copies from producer warps (step_tma.cuh) ----
template <int C, int EQ, int PSI>
static cudaError_t tma_launch(const StepLaunch &L, int z_begin, int z_count) {
  if constexpr (C > 2) {
    return cudaErrorNotSupported;  // C >= 3: the stages of a 57-double site do not fit (and neither do its registers)
  } else {
    // no mask path; 16-byte box starts and strides need an even nx; the host layer builds the tensor maps
    if (L.mask || L.g.nx % 2 != 0 || !L.tma_pop3 || !L.tma_pop1 || !L.tma_psi) return cudaErrorNotSupported;
    Wave w;
    cudaError_t e = wave_plan(L, z_begin, z_count, TMA_SEG, 1, w);
    if (e != cudaSuccess) return e;
    auto kern = tma_kernel<C, EQ, PSI, kExact>;
    constexpr int smem = tma_smem_bytes<C>();
    static int sms[64] = {};
    const int dev = device_slot();
    if (!sms[dev]) {
      e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
      if (e != cudaSuccess) return e;
      cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev);
    }
    // one persistent CTA per SM; fewer when the launch holds fewer items than the CTAs could take at once
    const int64_t items = (int64_t)w.nbands * w.slots * (w.rp + w.rf);
    const int64_t want = (items + 2 * TMA_LANES - 1) / (2 * TMA_LANES);
    const unsigned grid = (unsigned)(want < sms[dev] ? (want < 1 ? 1 : want) : sms[dev]);
    kern<<<grid, TMA_THREADS, smem, L.stream>>>(L.g, L.f, L.p, w, *L.tma_pop3, *L.tma_pop1, *L.tma_psi);
    return cudaGetLastError();
  }
}

template <int C, int EQ>
static cudaError_t tma_eq(const StepLaunch &L, int z_begin, int z_count) {
  return L.psi == 0 ? tma_launch<C, EQ, 0>(L, z_begin, z_count) : tma_launch<C, EQ, 1>(L, z_begin, z_count);
}

cudaError_t DLBC_MODE_NAME(launch_tma)(const StepLaunch &L, int z_begin, int z_count) {
  if (z_count <= 0) return cudaSuccess;
  if (!L.sc) return cudaErrorNotSupported;
  switch (L.eq) {
    case 0: return tma_eq<DLBC_NC, 0>(L, z_begin, z_count);
    case 1: return tma_eq<DLBC_NC, 1>(L, z_begin, z_count);
    case 2: return tma_eq<DLBC_NC, 2>(L, z_begin, z_count);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace dlbc

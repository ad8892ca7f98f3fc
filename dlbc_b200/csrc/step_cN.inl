// step_cN.inl -- instantiates the step kernels for DLBC_NC components in arithmetic mode
// DLBC_EXACT (1: translation unit built with -fmad=false; 0: FMA contraction allowed).
// Included by step_c<N>_{exact,fast}.cu.  The psi and force kernels (adds, multiplies in
// reference order, exp) live in the EXACT unit only.
#include "kernels.h"
#include "step_two_pass.cuh"
#include "step_zmarch.cuh"
#include "step_zpipe.cuh"
#include "step_wave.cuh"
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

// ---- z-marching single-pass variant (Shan-Chen only; C <= 2: 19 C doubles stay in registers) ----
#if DLBC_NC <= 2
template <int C, int EQ, int PSI, bool MASK, int TX, int TY, int MINB>
static cudaError_t zmarch_launch(const StepLaunch &L, int z_begin, int z_count, int chunk) {
  auto kern = zmarch_kernel<C, EQ, PSI, MASK, kExact, TX, TY, MINB>;
  constexpr size_t smem = (size_t)(C * Q * TX * TY + 4 * C * (TX + 2) * (TY + 2)) * sizeof(double);
  {
    // per device (function attributes are not shared between the devices of one process)
    static bool configured[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64 || !configured[dev]) {
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
      if (dev >= 0 && dev < 64) configured[dev] = true;
    }
  }
  ZMarch zm{z_begin, z_begin + z_count, chunk};
  dim3 grid((L.g.nx + TX - 1) / TX, (L.g.ny + TY - 1) / TY, (z_count + chunk - 1) / chunk);
  kern<<<grid, dim3(TX, TY), smem, L.stream>>>(L.g, L.f, L.p, zm);
  return cudaGetLastError();
}

template <int C, int EQ, int PSI, bool MASK>
static cudaError_t zmarch_tile(const StepLaunch &L, int z_begin, int z_count) {
  // tuning knobs (measurement only; defaults are the measured best)
  static const char *tile = std::getenv("DLBC_ZM_TILE");
  static const char *ch = std::getenv("DLBC_ZM_CHUNK");
  int chunk = ch ? std::atoi(ch) : 64;
  if (chunk < 1) chunk = 64;
  if (tile && tile[0] == '6' && tile[3] == '4') return zmarch_launch<C, EQ, PSI, MASK, 64, 4, 2>(L, z_begin, z_count, chunk);
  if (tile && tile[0] == '3') return zmarch_launch<C, EQ, PSI, MASK, 32, 8, 2>(L, z_begin, z_count, chunk);
  return zmarch_launch<C, EQ, PSI, MASK, 64, 8, 1>(L, z_begin, z_count, chunk);
}

template <int C, int EQ>
static cudaError_t zmarch_eq(const StepLaunch &L, int z_begin, int z_count) {
  if (L.psi == 0) return L.mask ? zmarch_tile<C, EQ, 0, true>(L, z_begin, z_count) : zmarch_tile<C, EQ, 0, false>(L, z_begin, z_count);
  return L.mask ? zmarch_tile<C, EQ, 1, true>(L, z_begin, z_count) : zmarch_tile<C, EQ, 1, false>(L, z_begin, z_count);
}
#endif

cudaError_t DLBC_MODE_NAME(launch_zmarch)(const StepLaunch &L, int z_begin, int z_count) {
  if (z_count <= 0) return cudaSuccess;
#if DLBC_NC <= 2
  if (!L.sc) return cudaErrorNotSupported;
  switch (L.eq) {
    case 0: return zmarch_eq<DLBC_NC, 0>(L, z_begin, z_count);
    case 1: return zmarch_eq<DLBC_NC, 1>(L, z_begin, z_count);
    case 2: return zmarch_eq<DLBC_NC, 2>(L, z_begin, z_count);
    default: return cudaErrorInvalidValue;
  }
#else
  (void)z_begin;
  return cudaErrorNotSupported;
#endif
}

// ---- asynchronously pipelined z-march (step_zpipe.cuh) ----
#if DLBC_NC <= 2
template <int C, int EQ, int PSI, bool MASK, int TX, int TY>
static cudaError_t zpipe_launch(const StepLaunch &L, int z_begin, int z_count, int chunk) {
  auto kern = zpipe_kernel<C, EQ, PSI, MASK, kExact, TX, TY>;
  constexpr int NT = TX * TY, NRIM = 2 * (TX + 2) + 2 * TY, RP = (TX + 2) * (TY + 2);
  constexpr size_t smem = (size_t)(2 * C * Q * NT + C * Q * NRIM + 4 * C * RP) * sizeof(double);
  static_assert(smem <= 227 * 1024, "zpipe tile does not fit in shared memory");
  {
    // per device (function attributes are not shared between the devices of one process)
    static bool configured[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64 || !configured[dev]) {
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
      if (dev >= 0 && dev < 64) configured[dev] = true;
    }
  }
  ZMarch zm{z_begin, z_begin + z_count, chunk};
  dim3 grid((L.g.nx + TX - 1) / TX, (L.g.ny + TY - 1) / TY, (z_count + chunk - 1) / chunk);
  kern<<<grid, dim3(TX, TY), smem, L.stream>>>(L.g, L.f, L.p, zm);
  return cudaGetLastError();
}

template <int C, int EQ, int PSI, bool MASK>
static cudaError_t zpipe_tile(const StepLaunch &L, int z_begin, int z_count) {
  static const char *tile = std::getenv("DLBC_ZP_TILE");
  static const char *ch = std::getenv("DLBC_ZP_CHUNK");
  int chunk = ch ? std::atoi(ch) : 128;
  if (chunk < 1) chunk = 128;
  if (tile && tile[0] == '6') return zpipe_launch<C, EQ, PSI, MASK, 64, 4>(L, z_begin, z_count, chunk);
  return zpipe_launch<C, EQ, PSI, MASK, 32, 8>(L, z_begin, z_count, chunk);
}

template <int C, int EQ>
static cudaError_t zpipe_eq(const StepLaunch &L, int z_begin, int z_count) {
  if (L.psi == 0) return L.mask ? zpipe_tile<C, EQ, 0, true>(L, z_begin, z_count) : zpipe_tile<C, EQ, 0, false>(L, z_begin, z_count);
  return L.mask ? zpipe_tile<C, EQ, 1, true>(L, z_begin, z_count) : zpipe_tile<C, EQ, 1, false>(L, z_begin, z_count);
}
#endif

cudaError_t DLBC_MODE_NAME(launch_zpipe)(const StepLaunch &L, int z_begin, int z_count) {
  if (z_count <= 0) return cudaSuccess;
#if DLBC_NC <= 2
  if (!L.sc) return cudaErrorNotSupported;
  switch (L.eq) {
    case 0: return zpipe_eq<DLBC_NC, 0>(L, z_begin, z_count);
    case 1: return zpipe_eq<DLBC_NC, 1>(L, z_begin, z_count);
    case 2: return zpipe_eq<DLBC_NC, 2>(L, z_begin, z_count);
    default: return cudaErrorInvalidValue;
  }
#else
  (void)z_begin;
  return cudaErrorNotSupported;
#endif
}

// ---- L2-ordered single launch (step_wave.cuh) ----
template <int C, int EQ, int PSI, bool MASK, bool SPLIT>
static cudaError_t wave_launch(const StepLaunch &L, int z_begin, int z_count) {
  static const char *e_tyb = std::getenv("DLBC_WAVE_TYB");
  static const char *e_la = std::getenv("DLBC_WAVE_LA");
  static const char *e_pf = std::getenv("DLBC_WAVE_PF");      // prefetch distance in tickets (0 = off = default)
  static const char *e_pfk = std::getenv("DLBC_WAVE_PFKEEP"); // 1 (default): prefetch with L2::evict_last
  const Geom &g = L.g;
  if (g.periodic_z_local && !(z_begin == 1 && z_count == g.nz)) return cudaErrorInvalidValue;
  Wave w;
  w.z_begin = z_begin;
  w.z_end = z_begin + z_count;
  w.pre = g.periodic_z_local ? 1 : 0;
  w.pl = g.periodic_z_local ? 1 : (z_begin - 1 < 1 ? 1 : z_begin - 1);
  w.ph = g.periodic_z_local ? g.nz - 1 : (w.z_end > g.nz ? g.nz : w.z_end);
  // band-plane of ~16k sites (5 MB of populations at C = 2): measured best on 512^2 planes (profiles/README.md)
  w.tyb = e_tyb ? std::atoi(e_tyb) : (int)(16384 / g.nx);
  if (w.tyb < 8) w.tyb = 8;
  if (w.tyb > g.ny) w.tyb = g.ny;
  w.xb = (g.nx + WAVE_SEG - 1) / WAVE_SEG;
  // lag of F behind P: at least 2 slots (F(z) reads psi(z+1)); more when a band-plane is smaller than
  // the ~80k sites the GPU keeps in flight, so that F blocks rarely have to wait.  With per-row completion
  // counters (rowsync) an F item only waits for the 9 row segments it reads, and one slot less is best
  // (512^2 planes, TYB = 32: LA 4 with rowsync 6.96 GLUPS, LA 5 without 6.81; profiles/README.md).
  static const char *e_rs0 = std::getenv("DLBC_WAVE_ROWSYNC");
  const int64_t lag_sites = (e_rs0 ? std::atoi(e_rs0) : 1) ? 65536 : 81920;
  int la = e_la ? std::atoi(e_la) : (int)((lag_sites + (int64_t)w.tyb * g.nx - 1) / ((int64_t)w.tyb * g.nx));
  if (la < 2) la = 2;
  if (!e_la && la > 8) la = 8;
  w.la = la;
  w.rp = (w.tyb + 2) * w.xb;
  w.rf = w.tyb * w.xb;
  // F(zf) runs in slot zf - pl + 1 + la (step_wave.cuh); the last one is zf = z_end - 1
  const int np = w.ph - w.pl + 1, nf = (w.z_end - 1 - w.pl + 1) + w.la;
  w.slots = 1 + (np > nf ? np : nf);
  w.nbands = (g.ny + w.tyb - 1) / w.tyb;
  w.sync = L.wave_sync;
  if (!w.sync) return cudaErrorInvalidValue;
  const int64_t blocks = (int64_t)w.nbands * w.slots * (w.rp + w.rf);
  if (blocks >= ((int64_t)1 << 31)) return cudaErrorInvalidConfiguration;
  cudaError_t e = cudaMemsetAsync(w.sync, 0, wave_sync_ints(g, w.tyb) * sizeof(int), L.stream);
  if (e != cudaSuccess) return e;
  // persistent grid: as many blocks as can be resident (fewer when the launch is small)
  static int resident = 0;
  if (!resident) {
    int dev = 0, sms = 0, per_sm = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, wave_kernel<C, EQ, PSI, MASK, kExact, SPLIT>, WAVE_SEG * (SPLIT ? C : 1), 0);
    resident = sms * (per_sm > 0 ? per_sm : 1);
  }
  const unsigned grid = (unsigned)(blocks < resident ? blocks : resident);
  w.pf = e_pf ? std::atoi(e_pf) : 0;   // measured slower when on (profiles/r01_wave_experiments.md): L2 capacity, not latency, is the limit
  w.pf_keep = e_pfk ? std::atoi(e_pfk) : 1;
  static const char *e_batch = std::getenv("DLBC_WAVE_BATCH");
  static const char *e_rs = std::getenv("DLBC_WAVE_ROWSYNC");
  w.rowsync = e_rs ? std::atoi(e_rs) : 1;
  w.batch = e_batch ? std::atoi(e_batch) : 1;  // with LA = 5: +2-4 % (LA = 4: -8 %), profiles/README.md
  static const char *e_free = std::getenv("DLBC_WAVE_FREE");  // 1: barrier-free form (wave_free_kernel)
  if (!SPLIT && (e_free ? std::atoi(e_free) != 0 : false)) {
    wave_free_kernel<C, EQ, PSI, MASK, kExact><<<grid, WAVE_SEG, 0, L.stream>>>(g, L.f, L.p, w);
    return cudaGetLastError();
  }
  wave_kernel<C, EQ, PSI, MASK, kExact, SPLIT><<<grid, dim3(WAVE_SEG, SPLIT ? C : 1), 0, L.stream>>>(g, L.f, L.p, w);
  return cudaGetLastError();
}

template <int C, int EQ, bool SPLIT>
static cudaError_t wave_psi(const StepLaunch &L, int z_begin, int z_count) {
  if (L.psi == 0) return L.mask ? wave_launch<C, EQ, 0, true, SPLIT>(L, z_begin, z_count) : wave_launch<C, EQ, 0, false, SPLIT>(L, z_begin, z_count);
  return L.mask ? wave_launch<C, EQ, 1, true, SPLIT>(L, z_begin, z_count) : wave_launch<C, EQ, 1, false, SPLIT>(L, z_begin, z_count);
}

template <int C, int EQ>
static cudaError_t wave_eq(const StepLaunch &L, int z_begin, int z_count) {
  // DLBC_WAVE_SPLIT=1 selects the one-thread-per-(site, component) form (step_wave.cuh); measured
  // slower than one thread per site (profiles/README.md), so it is off by default
  static const char *e_split = std::getenv("DLBC_WAVE_SPLIT");
  const bool split = C > 1 && (e_split ? std::atoi(e_split) != 0 : false);
  if constexpr (C > 1) { if (split) return wave_psi<C, EQ, true>(L, z_begin, z_count); }
  return wave_psi<C, EQ, false>(L, z_begin, z_count);
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

}  // namespace dlbc

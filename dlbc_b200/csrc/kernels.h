// kernels.h -- host-callable launchers of the step kernels (implemented in step_c*.cu).
#pragma once
#include <cuda.h>  // CUtensorMap (type only: the driver entry point is resolved at run time)
#include <cuda_runtime.h>
#include "lattice.cuh"

namespace dlbc {

struct StepLaunch {
  Geom g;
  Fields f;
  StepParams p;
  int eq, psi, sc, mask;
  cudaStream_t stream;
  int *wave_sync;  // ticket + completion counters of the wave / TMA variants (wave_sync_ints() ints)
  int wave_tyb;    // band height in rows, 0 = default (dlbc_cuda_set_option "wave_tyb")
  int wave_la;     // lag of the fused items behind the psi items in slots, 0 = default ("wave_la")
  // TMA variant: tensor maps of f_old (3-slot and 1-slot boxes) and of psi, built by the host layer
  // (dlbc_cuda.cu make_tma_maps); nullptr when the lattice has no TMA path
  const CUtensorMap *tma_pop3, *tma_pop1, *tma_psi;
};

constexpr int TMA_SEG = 64;  // sites per item of the TMA variant == x extent of its boxes minus the 4 pad columns

// size of StepLaunch::wave_sync: 1 ticket + one completion counter per (band, ghost-inclusive plane, row of
// the band incl. its two halo rows); allocated for the worst case (band height 1: 3 rows per band)
inline size_t wave_sync_ints(const Geom &g, int tyb) {
  const int nb = (g.ny + tyb - 1) / tyb;
  return 1 + (size_t)nb * (g.nz + 2) * (tyb + 2);
}
// upper bound of wave_sync_ints over all band heights (ceil(ny / t) * (t + 2) <= 3 ny + 16)
inline size_t wave_sync_alloc_ints(const Geom &g) { return 1 + (size_t)(3 * g.ny + 16) * (g.nz + 2); }

// z_begin / z_count address ghost-inclusive planes: the physical planes are 1..nz.
// Each returns cudaGetLastError() of its launch.  One translation unit per (component count,
// arithmetic mode) so the heavily unrolled kernels compile in parallel and so that the EXACT
// units can be built with -fmad=false (no FMA contraction: every fp64 operation rounded
// separately, as the reference's SSE2 code does) while the FAST units keep FMA contraction.
typedef cudaError_t (*launch_fn)(const StepLaunch &, int z_begin, int z_count);
// Shan-Chen force of component `comp` (L.force[comp], force.d:238-309) into out[3][sites_g].
typedef cudaError_t (*force_fn)(const StepLaunch &, int comp, double *out, int z_begin, int z_count);

#define DLBC_DECLARE_STEP(NC)                                                        \
  cudaError_t launch_psi_c##NC(const StepLaunch &, int, int);                        \
  cudaError_t launch_force_c##NC(const StepLaunch &, int, double *, int, int);       \
  cudaError_t launch_fused_exact_c##NC(const StepLaunch &, int, int);                \
  cudaError_t launch_fused_fast_c##NC(const StepLaunch &, int, int);                 \
  cudaError_t launch_wave_exact_c##NC(const StepLaunch &, int, int);                 \
  cudaError_t launch_wave_fast_c##NC(const StepLaunch &, int, int);                  \
  cudaError_t launch_tma_exact_c##NC(const StepLaunch &, int, int);                  \
  cudaError_t launch_tma_fast_c##NC(const StepLaunch &, int, int);
DLBC_DECLARE_STEP(1)
DLBC_DECLARE_STEP(2)
DLBC_DECLARE_STEP(3)
DLBC_DECLARE_STEP(4)
#undef DLBC_DECLARE_STEP

}  // namespace dlbc

// kernels.h -- host-callable launchers of the step kernels (implemented in step_c*.cu).
#pragma once
#include <cuda_runtime.h>
#include "lattice.cuh"

namespace dlbc {

struct StepLaunch {
  Geom g;
  Fields f;
  StepParams p;
  int eq, psi, sc, mask, exact;
  cudaStream_t stream;
};

// z_begin / z_count address ghost-inclusive planes: the physical planes are 1..nz.
// Each returns cudaGetLastError() of its launch.  One function per component count so the
// heavily unrolled kernels compile in parallel translation units.
typedef cudaError_t (*launch_fn)(const StepLaunch &, int z_begin, int z_count);
cudaError_t launch_psi_c1(const StepLaunch &, int, int);
cudaError_t launch_psi_c2(const StepLaunch &, int, int);
cudaError_t launch_psi_c3(const StepLaunch &, int, int);
cudaError_t launch_psi_c4(const StepLaunch &, int, int);
cudaError_t launch_fused_c1(const StepLaunch &, int, int);
cudaError_t launch_fused_c2(const StepLaunch &, int, int);
cudaError_t launch_fused_c3(const StepLaunch &, int, int);
cudaError_t launch_fused_c4(const StepLaunch &, int, int);

}  // namespace dlbc

// context.h -- the library's per-context state (host side).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/dlbc_cuda.h"
#include "kernels.h"

struct ncclComm;
namespace dlbc { namespace d2 { struct Lattice2; } }

struct dlbc_cuda_ctx {
  dlbc_cuda_config cfg{};
  dlbc::d2::Lattice2 *d2 = nullptr;   // a 2-D (D2Q9) context: everything else below is unused, see d2q9.cu
  dlbc::Geom g{};
  dlbc::StepParams p{};
  int device = 0;
  int variant = DLBC_KERNEL_TWO_PASS;  // requested / AUTO-resolved variant (effective_variant() applies the TMA size rule)
  bool variant_auto = false;
  int opt_wave_tyb = 0, opt_wave_la = 0;  // dlbc_cuda_set_option

  cudaStream_t compute = nullptr;  // step kernels, layout conversion, host <-> device copies
  cudaStream_t comm = nullptr;     // per-step ghost-plane exchange (NCCL), overlapped with compute
  cudaStream_t xfer = nullptr;     // host -> staging copies of uploads, overlapped with the layout-conversion kernels

  double *f[2] = {nullptr, nullptr};  // populations, A/B: [C][19][sites_g]
  int cur = 0;                        // f[cur] holds the current state
  double *psi = nullptr;              // [C][sites_g]
  double *rho_wall = nullptr;         // [C][sites_g], lazily built when gwc != 0 and solids exist
  uint8_t *mask = nullptr;            // [sites_g]
  uint32_t *nbmask = nullptr;         // [sites_g] neighbourhood solid flags (Fields::nb), allocated when solids exist
  bool has_solid = false;
  bool rho_wall_valid = false;
  bool ghosts_valid = false;          // f[cur] ghost planes hold the neighbours' boundary planes
  bool force_valid = false;           // f[1-cur] / psi hold the inputs of the last step

  // staging for uploads/downloads/outputs
  void *stage = nullptr;              // device, stage_bytes
  size_t stage_bytes = 0;
  int stage_half = 0;                 // uploads double-buffer the staging area: half used by the next chunk
  cudaEvent_t ev_copy = nullptr;      // xfer: chunk copied into its half (the host buffer is free again)
  cudaEvent_t ev_stage[2] = {nullptr, nullptr};  // compute: the conversion kernel has consumed half i
  double *partial = nullptr;          // device, reduction partials
  int *flags = nullptr;               // device, [2]: bad mask value seen, solid count
  // TMA variant: tensor maps of both population buffers (boxes of 3 slots / 1 slot) and of psi (step_tma.cuh)
  CUtensorMap tma_pop3[2], tma_pop1[2], tma_psi;
  bool tma_maps = false;
  int *wave_sync[3] = {nullptr, nullptr, nullptr};  // wave variant: ticket + counters (interior, lower, upper boundary launch)

  // multi-rank
  ncclComm *nccl = nullptr;
  int up = 0, down = 0;               // neighbour ranks in +z / -z

  // events for measurement
  cudaEvent_t ev_a = nullptr, ev_b = nullptr;
  // cross-stream ordering of the multi-rank step (no timing)
  cudaEvent_t ev_bnd = nullptr;        // compute: boundary planes written
  cudaEvent_t ev_ghost_psi = nullptr;  // comm: psi ghost planes received
  cudaEvent_t ev_ghost_f = nullptr;    // comm: population ghost planes of the new buffer received
  std::vector<cudaEvent_t> ev_pool;

  int64_t timestep = 0;
  int64_t launches = 0;
  int64_t device_bytes = 0;
  std::string err;
};

// step_tma.cuh -- kernel variant DLBC_KERNEL_TMA: the wave schedule of step_wave.cuh (psi items P and
// fused items F of one launch, ticket-ordered band by band so that the second gather of every population
// is an L2 hit) with the data movement taken off the compute warps.
//
// One persistent CTA per SM, warp-specialised:
//   * warps 0-3 are PRODUCERS.  Thread 0 takes tickets; a ticket is one lattice row of one kind (P or F) and is
//     cut into items of TMA_SEG sites, which the four warps stage round-robin into a ring of shared-memory
//     stages with tensor-map TMA copies (cp.async.bulk.tensor, SASS UTMALDG, completion counted on an
//     mbarrier).  The issue rate of TMA copies is ~45 ns per copy and warp whatever their size
//     (profiles/r02/ubench_tma_*.txt), so an item must be FEW copies: the device layout stores directions that
//     stream from the same neighbouring row in consecutive planes (qslot, lb_site.cuh) and the box of one copy
//     spans 3 planes x all components:
//       - populations: 9 boxes (5 x [SEG+4, 1, 3 slots, C] + 4 x [SEG+4, 1, 1 slot, C]) of the 5-D tensor
//         f[zg][c][slot][y][x]: box g holds the row y - c_y, plane z - c_z its directions are pulled from;
//       - F items: 3 boxes [SEG+4, 1, 3 planes, C] of psi[c][zg][y][x] (rows y-1, y, y+1; planes z-1..z+1).
//     A box starts at x0 - 2 (fp64 box starts must be 16-byte aligned) and is SEG + 4 wide, so the +-1 shifts in
//     x are index offsets for the consumer.  Out-of-range columns (x < 0, x >= nx) are zero-filled by the TMA:
//     the one thread per edge item that needs a wrapped neighbour loads it itself.
//     F tickets first wait until the nine psi rows they read are complete (P items publish per (band, plane,
//     row) with red.release.gpu; relaxed polls, one acquire fence, fence.proxy.async before the copies).
//   * warps 4.. are CONSUMERS in groups of TMA_SEG threads.  A group claims the next staged item, moves its
//     operands shared -> registers with immediate-offset LDS, hands the stage back, and runs the same per-site
//     functions as every other variant (lb_site.cuh: identical bits), storing psi (P) or the post-collision
//     populations (F) straight from registers.
//
// Deadlock freedom: as in step_wave.cuh -- a ticket only ever waits for smaller tickets; slots of the ring are
// numbered in ticket order within the CTA, each slot has exactly one producer warp, and consumers wait for
// nothing but copies that have been issued.
//
// Scope: no solid sites (AUTO takes the register path when a mask has solids), nx even, C <= 2 (the ring must
// hold at least as many stages as there are consumer groups).
#pragma once
#include <cuda.h>
#include "step_wave.cuh"

namespace dlbc {

constexpr int TMA_ROW = TMA_SEG + 4;         // doubles per staged row: element i <-> lattice x = x0 - 2 + i
constexpr int TMA_ROWB = TMA_ROW * 8;
constexpr int TMA_GW = TMA_SEG / 32;         // warps per consumer group
constexpr int TMA_PWARPS = 4;                // producer warps (warpgroup 0)
constexpr int TMA_PCWARPS = 4;               // consumer warps of the P items (warpgroup 1)
constexpr int TMA_FCWARPS = 8;               // consumer warps of the F items (warpgroups 2, 3)
constexpr int TMA_THREADS = 32 * (TMA_PWARPS + TMA_PCWARPS + TMA_FCWARPS);
// register split (setmaxnreg, per warpgroup): the kernel is compiled for 65536 / 512 = 128 registers per thread;
// the producers shrink to 40, the P consumers (19 populations at a time, a sum, an exp) to 56, and the F consumers
// (19 C populations + force + moments per thread; no spills: L1 is all but given to the stages) grow to 208:
// 128 * 40 + 128 * 56 + 256 * 208 = 65536
constexpr int TMA_REGS_PRODUCER = 40, TMA_REGS_PCONS = 56, TMA_REGS_FCONS = 208;
static_assert(TMA_THREADS == 512 && TMA_PWARPS == 4 && TMA_PCWARPS == 4, "register split assumes 4 warpgroups of 128");
constexpr int TMA_SMEM_BUDGET = 227 * 1024 - 2048;  // dynamic shared memory available next to the static control block

// ---- stage layout -----------------------------------------------------------------------------------------
// population boxes 0..4 (3 slots each), 5..8 (1 slot each), then the three psi boxes (rows y-1, y, y+1); every
// box starts on a 128-byte boundary (TMA destination alignment).  Box g <-> (c_y, c_z) of its directions.
__host__ __device__ constexpr int tma_up128(int b) { return (b + 127) / 128 * 128; }
template <int C> __host__ __device__ constexpr int tma_box3() { return tma_up128(TMA_ROWB * 3 * C); }
template <int C> __host__ __device__ constexpr int tma_box1() { return tma_up128(TMA_ROWB * C); }
template <int C> __host__ __device__ constexpr int tma_pop_bytes() { return 5 * tma_box3<C>() + 4 * tma_box1<C>(); }
// two rings: P stages hold the 9 population boxes, F stages those + the 3 psi boxes
template <int C> __host__ __device__ constexpr int tma_pstage_bytes() { return tma_pop_bytes<C>(); }
template <int C> __host__ __device__ constexpr int tma_fstage_bytes() { return tma_pop_bytes<C>() + 3 * tma_box3<C>(); }
constexpr int TMA_NF = 4;  // F stages: one per producer warp (a stage is always filled by the same warp)
template <int C>
__host__ __device__ constexpr int tma_np() {  // P stages: what is left, in multiples of ... at most 8
  return (TMA_SMEM_BUDGET - TMA_NF * tma_fstage_bytes<C>()) / tma_pstage_bytes<C>() > 8
             ? 8 : (TMA_SMEM_BUDGET - TMA_NF * tma_fstage_bytes<C>()) / tma_pstage_bytes<C>();
}
template <int C> __host__ __device__ constexpr int tma_smem_bytes() { return TMA_NF * tma_fstage_bytes<C>() + tma_np<C>() * tma_pstage_bytes<C>(); }
__host__ __device__ constexpr int tma_box_of(int q) {  // (c_y, c_z) -> box
  return cz(q) == 0 ? (cy(q) == 0 ? 0 : (cy(q) == 1 ? 1 : 2))
                    : (cy(q) == 0 ? (cz(q) == 1 ? 3 : 4) : (cz(q) == 1 ? (cy(q) == 1 ? 5 : 6) : (cy(q) == 1 ? 7 : 8)));
}
__host__ __device__ constexpr int tma_box_slot0(int b) { return b < 5 ? 3 * b : 10 + b; }
// element (double) index of direction q of component c inside a stage, for the site at staged column 0
template <int C>
__host__ __device__ constexpr int tma_pop_elem(int c, int q) {
  const int b = tma_box_of(q);
  const int k = qslot(q) - tma_box_slot0(b);
  return b < 5 ? (b * tma_box3<C>()) / 8 + (c * 3 + k) * TMA_ROW
               : (5 * tma_box3<C>() + (b - 5) * tma_box1<C>()) / 8 + c * TMA_ROW;
}
// psi of component c at the neighbour in direction (dy, dz), staged column 0
template <int C>
__host__ __device__ constexpr int tma_psi_elem(int c, int dy, int dz) {
  return (tma_pop_bytes<C>() + (dy + 1) * tma_box3<C>()) / 8 + (c * 3 + dz + 1) * TMA_ROW;
}

template <int N>
struct TmaRing {
  unsigned long long full[N];   // a producer warp arrives with expect_tx, the copies complete_tx
  unsigned long long empty[N];  // the consumer group's warps arrive once their operands are in registers
  int desc[N][8];               // kind (1 P, 2 F, 3 end), zg, y, x0, sites, counter index, slot number
  int claim;                    // next slot a consumer group takes
};
template <int C>
struct TmaCtl {
  TmaRing<tma_np<C>()> p;
  TmaRing<TMA_NF> f;
  int ticket[2];  // producer warps: the ticket being staged / the next one
};

// ---- PTX wrappers ----------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n"
      "selp.u32 %0, 1, 0, P1;\n"
      "}\n" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
// wait for the completion of the phase with this parity (watchdog: see step_wave.cuh)
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  Watchdog wd;
  while (!mbar_try_wait(bar, parity)) wd.tick();
}
// tensor-map copies global -> shared (SASS UTMALDG), completion in bytes on an mbarrier, L2 eviction hint
__device__ __forceinline__ void tma_load_5d(uint32_t dst, const CUtensorMap *map, int c0, int c1, int c2, int c3, int c4,
                                            uint32_t bar, uint64_t pol) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4, %5, %6, %7}], [%2], %8;\n" ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4), "l"(pol)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap *map, int c0, int c1, int c2, int c3, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];\n" ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void named_bar(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;\n" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ int wrap_y(int y, int ny) { return y < 0 ? y + ny : (y >= ny ? y - ny : y); }

// c_y / c_z of population box b for a run-time b: 2-bit codes packed at compile time
__host__ __device__ constexpr int tma_box_cy_c(int b) { return b == 1 || b == 5 || b == 7 ? 1 : (b == 2 || b == 6 || b == 8 ? -1 : 0); }
__host__ __device__ constexpr int tma_box_cz_c(int b) { return b == 3 || b == 5 || b == 6 ? 1 : (b == 4 || b == 7 || b == 8 ? -1 : 0); }
__host__ __device__ constexpr bool tma_boxes_consistent() {
  for (int q = 0; q < Q; q++) {
    const int b = tma_box_of(q);
    if (tma_box_cy_c(b) != cy(q) || tma_box_cz_c(b) != cz(q)) return false;
    if (qslot(q) < tma_box_slot0(b) || qslot(q) >= tma_box_slot0(b) + (b < 5 ? 3 : 1)) return false;
  }
  return true;
}
static_assert(tma_boxes_consistent(), "box <-> (c_y, c_z) <-> storage slots");
__host__ __device__ constexpr unsigned tma_pack_cy() { unsigned m = 0; for (int b = 0; b < 9; b++) m |= (unsigned)(tma_box_cy_c(b) + 1) << (2 * b); return m; }
__host__ __device__ constexpr unsigned tma_pack_cz() { unsigned m = 0; for (int b = 0; b < 9; b++) m |= (unsigned)(tma_box_cz_c(b) + 1) << (2 * b); return m; }
__device__ __forceinline__ int tma_box_cy(int b) { constexpr unsigned m = tma_pack_cy(); return (int)((m >> (2 * b)) & 3u) - 1; }
__device__ __forceinline__ int tma_box_cz(int b) { constexpr unsigned m = tma_pack_cz(); return (int)((m >> (2 * b)) & 3u) - 1; }

// ---- consumer side ------------------------------------------------------------------------------------
// lb/advection.d:61-89 from the staged rows: f[q] = old[p - c_q][q] sits at staged column t + 2 - c_x(q) of its
// row (tma_pop_elem).  The wrapped neighbours of the first / last site of a lattice row are zero-filled by the
// TMA: left (x = 0) pulls the c_x = +1 directions from x = nx - 1, right (x = nx - 1) the c_x = -1 directions
// from x = 0, straight from global memory (f_old is read-only during the step).
__device__ __forceinline__ void tma_pull_edge(const Geom &g, const double *f_old, int zg, int y, int c, bool left, bool right,
                                              double (&f)[Q]) {
  const int xs = left ? g.nx - 1 : 0;
#pragma unroll
  for (int q = 0; q < Q; q++) {
    if (cx(q) == 0) continue;
    if (cx(q) == 1 ? left : right) {
      const double *a = f_old + (int64_t)(zg - cz(q)) * g.fz + (int64_t)c * g.fc + (int64_t)qslot(q) * g.plane +
                        (int64_t)wrap_y(y - cy(q), g.ny) * g.nx + xs;
      f[q] = ld_global_nc(a);
    }
  }
}

// A consumer WARP takes the next half-item (32 sites) of its ring -- slot n = h / 2, half h % 2 -- and waits for
// the slot's copies; no barrier between the warps of a stage.  Slots are claimed dynamically, so slot n - N (the
// stage's previous use) may still be waiting for ITS copies: the parity wait then passes one phase early.  The
// slot number in the descriptor tells: it can only be n once slot n - N has been consumed and released.
// `before_block()` runs once if the warp is about to wait (the P consumers flush their pending releases there:
// whatever it waits for may, through an F ticket polled by the producers, depend on them).
template <int N, class BeforeBlock>
__device__ __forceinline__ int tma_claim(TmaRing<N> &ring, int lane, int &half, const volatile int *&d, BeforeBlock before_block) {
  int h = 0;
  if (lane == 0) h = atomicAdd(&ring.claim, 1);
  h = __shfl_sync(0xffffffffu, h, 0);
  const int n = h / TMA_GW;
  half = h - n * TMA_GW;
  const int st = n % N;
  d = ring.desc[st];
  const uint32_t full = smem_u32(&ring.full[st]);
  const uint32_t par = (n / N) & 1;
  if (!mbar_try_wait(full, par)) before_block();
  for (;;) {
    mbar_wait(full, par);
    if (d[6] == n) break;
    mbar_wait(full, par ^ 1);  // the previous use lands first
  }
  return st;
}

template <int C, int EQ, int PSI, bool EXACT>
__global__ void __launch_bounds__(TMA_THREADS, 1)
tma_kernel(Geom g, Fields fl, StepParams P, Wave w, const __grid_constant__ CUtensorMap map_pop3,
           const __grid_constant__ CUtensorMap map_pop1, const __grid_constant__ CUtensorMap map_psi) {
  extern __shared__ __align__(128) unsigned char smem[];  // [TMA_NF F stages][NP P stages]
  __shared__ TmaCtl<C> ctl;
  constexpr int NP = tma_np<C>(), NF = TMA_NF;
  constexpr int PB = tma_pstage_bytes<C>(), FB = tma_fstage_bytes<C>();
  static_assert(NP >= 2, "no room for two P stages");
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < NP; i++) {
      mbar_init(smem_u32(&ctl.p.full[i]), 1);
      mbar_init(smem_u32(&ctl.p.empty[i]), TMA_GW);
      ctl.p.desc[i][6] = -1;
    }
    for (int i = 0; i < NF; i++) {
      mbar_init(smem_u32(&ctl.f.full[i]), 1);
      mbar_init(smem_u32(&ctl.f.empty[i]), TMA_GW);
      ctl.f.desc[i][6] = -1;
    }
    ctl.p.claim = 0;
    ctl.f.claim = 0;
    ctl.ticket[0] = atomicAdd(w.sync, 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  }
  __syncthreads();

  const int total = w.nbands * w.slots * (w.rp + w.rf);  // tickets (w.xb == 1: one row each)
  const int nsub = w.nsub;                               // items per ticket

  if (warp < TMA_PWARPS) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;\n" ::"n"(TMA_REGS_PRODUCER));
    // ================================ producers ================================
    const uint64_t pol_keep = make_policy<POL_KEEP>();      // P: first touch of the populations, keep for the F items
    const uint64_t pol_stream = make_policy<POL_STREAM>();  // F: last touch
    int nbase_p = 0, nbase_f = 0;  // slots of each ring used by this CTA's earlier tickets
    for (int it = 0;; it ^= 1) {
      const int t = ctl.ticket[it];
      if (t >= total) break;
      int nxt = 0;
      if (threadIdx.x == 0) nxt = atomicAdd(w.sync, 1);  // in flight while this ticket is staged
      const WaveItem wi = wave_decode(g, w, t);
      if (wi.kind != 0) {
        const bool isf = wi.kind == 2;
        if (isf) {
          if (lane < 9) {
            const int *c = wave_wait_counter(g, w, wi, lane);
            if (c) wave_poll_until(c, nsub * TMA_GW, 64);  // one count per half-item
            wave_acquire();
          }
          __syncwarp();
          asm volatile("fence.proxy.async.global;\n" ::: "memory");  // generic-proxy psi stores -> async-proxy reads
        }
        const uint64_t pol = isf ? pol_stream : pol_keep;
        const int nbase = isf ? nbase_f : nbase_p;
        const int N = isf ? NF : NP;
        for (int i = 0; i < nsub; i++) {
          const int n = nbase + i;
          const int st = n % N;
          // a stage is always filled by the same warp, so its uses are staged in order: the parity wait on the
          // empty barrier below tells "use k - 1 released" from "not yet" only if use k - 2 is known to be over
          if (st % TMA_PWARPS != warp) continue;
          const uint32_t full = smem_u32(isf ? &ctl.f.full[st] : &ctl.p.full[st]);
          const uint32_t base = smem_u32(smem) + (isf ? st * FB : NF * FB + st * PB);
          const int x0 = i * TMA_SEG;
          mbar_wait(smem_u32(isf ? &ctl.f.empty[st] : &ctl.p.empty[st]), ((n / N) & 1) ^ 1);
          if (lane == 0) {
            int *d = isf ? ctl.f.desc[st] : ctl.p.desc[st];
            d[0] = wi.kind; d[1] = wi.zg; d[2] = wi.y; d[3] = x0; d[4] = min(TMA_SEG, g.nx - x0);
            d[5] = wi.cnt + wi.zg * (w.tyb + 2) + wi.row;
            d[6] = n;
            mbar_arrive_tx(full, (uint32_t)(TMA_ROWB * C * (isf ? Q + 9 : Q)));
          }
          __syncwarp();
          if (lane < 9) {
            const int b = lane;
            const int ys = wrap_y(wi.y - tma_box_cy(b), g.ny);
            const int zs = wi.zg - tma_box_cz(b);
            if (b < 5) tma_load_5d(base + b * tma_box3<C>(), &map_pop3, x0 - 2, ys, 3 * b, 0, zs, full, pol);
            else tma_load_5d(base + 5 * tma_box3<C>() + (b - 5) * tma_box1<C>(), &map_pop1, x0 - 2, ys, 10 + b, 0, zs, full, pol);
          } else if (lane < 12 && isf) {
            const int dy = lane - 10;
            tma_load_4d(base + tma_pop_bytes<C>() + (dy + 1) * tma_box3<C>(), &map_psi, x0 - 2, wrap_y(wi.y + dy, g.ny),
                        wi.zg - 1, 0, full);
          }
        }
        if (isf) nbase_f += nsub; else nbase_p += nsub;
      }
      if (threadIdx.x == 0) ctl.ticket[it ^ 1] = nxt;
      named_bar(1, 32 * TMA_PWARPS);
    }
    // one end half-item per consumer warp
    if (threadIdx.x == 0) {
      for (int e = 0; e < TMA_PCWARPS / TMA_GW; e++) {
        const int n = nbase_p + e, st = n % NP;
        mbar_wait(smem_u32(&ctl.p.empty[st]), ((n / NP) & 1) ^ 1);
        ctl.p.desc[st][0] = 3;
        ctl.p.desc[st][6] = n;
        mbar_arrive(smem_u32(&ctl.p.full[st]));
      }
      for (int e = 0; e < TMA_FCWARPS / TMA_GW; e++) {
        const int n = nbase_f + e, st = n % NF;
        mbar_wait(smem_u32(&ctl.f.empty[st]), ((n / NF) & 1) ^ 1);
        ctl.f.desc[st][0] = 3;
        ctl.f.desc[st][6] = n;
        mbar_arrive(smem_u32(&ctl.f.full[st]));
      }
    }
  } else if (warp < TMA_PWARPS + TMA_PCWARPS) {
    // ================================ P consumers ================================
    // psi of the post-advection densities (density.d:89-96, force.d:360-367), one component at a time.  Each warp
    // works on its own half-items.  Publishing a row segment costs a gpu-scope fence (~1 us until the psi stores
    // are acknowledged): up to four are batched behind one fence, unless the warp would have to wait.
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;\n" ::"n"(TMA_REGS_PCONS));
    int npend = 0, mypend = 0;  // lane i < npend holds the counter index of the i-th unpublished half-item
    auto flush = [&]() {
      __syncwarp();  // the psi stores of all lanes precede (warp barrier, cumulativity) the fences below
      if (lane < npend) {
        asm volatile("fence.proxy.async.global;\n" ::: "memory");
        asm volatile("fence.acq_rel.gpu;\n" ::: "memory");
        asm volatile("red.relaxed.gpu.global.add.s32 [%0], %1;\n" ::"l"(w.sync + mypend), "r"(1) : "memory");
      }
      npend = 0;
    };
    for (;;) {
      const volatile int *d;
      int half;
      const int st = tma_claim(ctl.p, lane, half, d, flush);
      if (d[0] == 3) break;
      const int zg = d[1], y = d[2], x0 = d[3], cnt = d[4], cidx = d[5];
      const int t = half * 32 + lane;  // site within the item
      const double *sp = reinterpret_cast<const double *>(smem + NF * FB + st * PB) + t;
      const bool act = t < cnt;
      const int sxy = y * g.nx + x0 + t;
      // items at the ends of a row: the first / last site's wrapped x neighbour is not in the box
      const bool left = act && x0 + t == 0, right = act && x0 + t == g.nx - 1;
      double v[C];
#pragma unroll
      for (int c = 0; c < C; c++) {
        double f[Q];
        if (act) {
#pragma unroll
          for (int q = 0; q < Q; q++) f[q] = sp[tma_pop_elem<C>(c, q) + 2 - cx(q)];
          if (left || right) tma_pull_edge(g, fl.f_old, zg, y, c, left, right, f);
        }
        v[c] = act ? psi_of<PSI>(density19(f)) : 0.0;
        if (c == C - 1) {
          __syncwarp();
          if (lane == 0) mbar_arrive(smem_u32(&ctl.p.empty[st]));  // the last operands are in registers: hand the stage back
        }
      }
      if (act) {
#pragma unroll
        for (int c = 0; c < C; c++) {
          psi_plane(fl.psi, g, zg, c)[sxy] = v[c];
          if (g.periodic_z_local) {
            if (zg == g.nz) psi_plane(fl.psi, g, 0, c)[sxy] = v[c];      // -> ghost plane 0
            if (zg == 1) psi_plane(fl.psi, g, g.nz + 1, c)[sxy] = v[c];  // -> ghost plane nz+1
          }
        }
      }
      if (lane == npend) mypend = cidx;
      if (++npend == 4) flush();
    }
    flush();
  } else {
    // ================================ F consumers ================================
    // pull + Shan-Chen force + collide (advection.d, force.d:238-309, collision.d:90-119)
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;\n" ::"n"(TMA_REGS_FCONS));
    for (;;) {
      const volatile int *d;
      int half;
      const int st = tma_claim(ctl.f, lane, half, d, []() {});
      if (d[0] == 3) break;
      const int zg = d[1], y = d[2], x0 = d[3], cnt = d[4];
      const int t = half * 32 + lane;
      const double *sp = reinterpret_cast<const double *>(smem + st * FB) + t;
      const bool act = t < cnt;
      const int sxy = y * g.nx + x0 + t;
      const bool edge_item = x0 == 0 || x0 + cnt == g.nx;
      const bool left = act && x0 + t == 0, right = act && x0 + t == g.nx - 1;
      double f[C][Q], F[C][3], psi1[C];
#pragma unroll
      for (int c = 0; c < C; c++) { F[c][0] = 0.0; F[c][1] = 0.0; F[c][2] = 0.0; psi1[c] = 0.0; }
      if (act) {
        // the force first: its 18 C psi operands are dead again before the 19 C populations are loaded
#pragma unroll
        for (int c = 0; c < C; c++) psi1[c] = sp[tma_psi_elem<C>(c, 0, 0) + 2];
        if (!edge_item) {
          sc_force<C, EXACT>(P, psi1, F, [&](int f2, int q) -> double {
            return sp[tma_psi_elem<C>(f2, cy(q), cz(q)) + 2 + cx(q)];
          });
        } else {
          // wrapped x neighbours (zero-filled by the TMA) come straight from L2: coherent loads, the producer's
          // acquire precedes this thread's wait on the full barrier
          const int xs = left ? g.nx - 1 : 0;
          sc_force<C, EXACT>(P, psi1, F, [&](int f2, int q) -> double {
            if (cx(q) != 0 && (cx(q) == -1 ? left : right))
              return ld_global_cg(fl.psi + (int64_t)f2 * g.sites_g + (int64_t)(zg + cz(q)) * g.plane +
                                  (int64_t)wrap_y(y + cy(q), g.ny) * g.nx + xs);
            return sp[tma_psi_elem<C>(f2, cy(q), cz(q)) + 2 + cx(q)];
          });
        }
#pragma unroll
        for (int c = 0; c < C; c++)
#pragma unroll
          for (int q = 0; q < Q; q++) f[c][q] = sp[tma_pop_elem<C>(c, q) + 2 - cx(q)];
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&ctl.f.empty[st]));
      if (act) {
        if (left || right) {
#pragma unroll
          for (int c = 0; c < C; c++) tma_pull_edge(g, fl.f_old, zg, y, c, left, right, f[c]);
        }
        double rho[C], m[C][3];
#pragma unroll
        for (int c = 0; c < C; c++) {
          rho[c] = density19(f[c]);
          momentum19(f[c], m[c]);
        }
        double u[3] = {0.0, 0.0, 0.0};
        if (EQ == 2) common_velocity<C, EXACT>(P, m, rho, u);
#pragma unroll
        for (int c = 0; c < C; c++) {
          collide_site<EQ, EXACT>(f[c], rho[c], m[c], F[c], u, P, c);
          store19<POL_STREAM>(g, fl.f_new, zg, c, sxy, f[c]);
        }
      }
    }
  }
}

}  // namespace dlbc

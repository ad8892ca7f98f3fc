// step_tma.cuh -- kernel variant DLBC_KERNEL_TMA: the wave schedule of step_wave.cuh (psi items P and
// fused items F of one launch, ticket-ordered band by band so that the second gather of every population
// is an L2 hit) with the data movement taken off the compute warps.
//
// One persistent CTA per SM = four independent LANES, each a little pipeline of its own:
//   * a PRODUCER warp takes tickets two at a time (a ticket is an item: a row segment of TMA_SEG sites of one
//     kind; in a slot the items alternate P, F, P, F, ... so a pair is normally one of each) and stages them
//     into the lane's P stage and F stage with tensor-map TMA copies
//     (cp.async.bulk.tensor, SASS UTMALDG, completion counted on an mbarrier).  The issue rate of TMA copies is
//     ~45 ns per copy and warp whatever their size (profiles/r02/ubench_tma_*.txt), so an item must be FEW
//     copies: the device layout stores directions that stream from the same neighbouring row in consecutive
//     planes (qslot, lb_site.cuh) and the box of one copy spans 3 planes x all components:
//       - populations: 9 boxes (5 x [SEG+4, 1, 3 slots, C] + 4 x [SEG+4, 1, 1 slot, C]) of the 5-D tensor
//         f[zg][c][slot][y][x]: box b holds the row y - c_y, plane z - c_z its directions are pulled from;
//       - F items: 3 boxes [SEG+4, 1, 3 planes, C] of psi[c][zg][y][x] (rows y-1, y, y+1; planes z-1..z+1).
//     A box starts at x0 - 2 (fp64 box starts must be 16-byte aligned) and is SEG + 4 wide, so the +-1 shifts in
//     x are index offsets for the consumer.  Out-of-range columns (x < 0, x >= nx) are zero-filled by the TMA:
//     the one thread per edge item that needs a wrapped neighbour loads it itself.
//     Before an F ticket is staged the producer waits until the nine psi rows it reads are complete (P items
//     publish per (band, plane, row); relaxed polls, one acquire fence, fence.proxy.async before the copies).
//     Only this lane waits: the other three -- and the other 147 CTAs -- carry on, as the independent blocks of
//     the register-path kernel do.
//   * a P-CONSUMER warp (56 registers) turns the lane's P stage into psi: 2 sites per thread, one component at a
//     time, densities first, stage handed back, then exp and the stores.  Publishing costs a gpu-scope fence
//     (~1 us): up to four items are published behind one fence unless the warp would have to wait.
//   * two F-CONSUMER warps (208 registers: no spills) take half an F item each: operands shared -> registers with
//     immediate-offset LDS, stage handed back, then the same per-site functions as every other variant
//     (lb_site.cuh: identical bits) and the stores of the post-collision populations.
// Every stage has exactly one producer and one consumer (pair), used strictly in turn: plain phase counters, no
// claims, no barriers between warps.
//
// Deadlock freedom: as in step_wave.cuh -- a ticket only ever waits for smaller tickets, each of which is held
// by a lane that is itself staging or waiting for still smaller ones; consumers wait for nothing but copies that
// have been issued.
//
// Scope: no solid sites (AUTO takes the register path when a mask has solids), nx even, C <= 2 (stage sizes).
#pragma once
#include <cuda.h>
#include "step_wave.cuh"

namespace dlbc {

constexpr int TMA_ROW = TMA_SEG + 4;         // doubles per staged row: element i <-> lattice x = x0 - 2 + i
constexpr int TMA_ROWB = TMA_ROW * 8;
constexpr int TMA_LANES = 4;                 // independent pipelines per CTA
constexpr int TMA_THREADS = 32 * TMA_LANES * 4;  // per lane: producer, P consumer, two F consumers
static_assert(TMA_SEG == 64, "an F item is two warps wide, a P item two sites per thread");
// register split (setmaxnreg, per warpgroup): the kernel is compiled for 65536 / 512 = 128 registers per thread;
// warps 0-3 (producers) shrink to 48, warps 4-7 (P consumers: 19 populations at a time, a sum, an exp) to 80, and
// warps 8-15 (F consumers: 19 C populations + force + moments per thread, no spills -- L1 is all but given to the
// stages) grow to 192: 128 * 48 + 128 * 80 + 256 * 192 = 65536
constexpr int TMA_REGS_PRODUCER = 48, TMA_REGS_PCONS = 80, TMA_REGS_FCONS = 192;
constexpr int TMA_SMEM_BUDGET = 227 * 1024 - 2048;  // dynamic shared memory available next to the static control block

// ---- stage layout -----------------------------------------------------------------------------------------
// population boxes 0..4 (3 slots each), 5..8 (1 slot each), then the three psi boxes (rows y-1, y, y+1); every
// box starts on a 128-byte boundary (TMA destination alignment).  Box g <-> (c_y, c_z) of its directions.
__host__ __device__ constexpr int tma_up128(int b) { return (b + 127) / 128 * 128; }
template <int C> __host__ __device__ constexpr int tma_box3() { return tma_up128(TMA_ROWB * 3 * C); }
template <int C> __host__ __device__ constexpr int tma_box1() { return tma_up128(TMA_ROWB * C); }
template <int C> __host__ __device__ constexpr int tma_pop_bytes() { return 5 * tma_box3<C>() + 4 * tma_box1<C>(); }
// per lane one P stage (the 9 population boxes) and one F stage (those + the 3 psi boxes)
template <int C> __host__ __device__ constexpr int tma_pstage_bytes() { return tma_pop_bytes<C>(); }
template <int C> __host__ __device__ constexpr int tma_fstage_bytes() { return tma_pop_bytes<C>() + 3 * tma_box3<C>(); }
template <int C> __host__ __device__ constexpr int tma_smem_bytes() { return TMA_LANES * (tma_fstage_bytes<C>() + tma_pstage_bytes<C>()); }
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

struct TmaCtl {
  unsigned long long full_p[TMA_LANES], empty_p[TMA_LANES];  // full: the producer arrives with expect_tx, the copies
  unsigned long long full_f[TMA_LANES], empty_f[TMA_LANES];  // complete_tx; empty: the consumer warp(s) arrive
  int desc_p[TMA_LANES][8], desc_f[TMA_LANES][8];            // kind (1 P, 2 F, 3 end), zg, y, x0, sites, counter index
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
  for (int spins = 0; !mbar_try_wait(bar, parity); spins++) {
    __nanosleep(40);  // a spinning warp competes for issue slots with the working warps of its sub-partition
    if ((spins & 63) == 63) wd.tick();
  }
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

template <int C, int EQ, int PSI, bool EXACT>
__global__ void __launch_bounds__(TMA_THREADS, 1)
tma_kernel(Geom g, Fields fl, StepParams P, Wave w, const __grid_constant__ CUtensorMap map_pop3,
           const __grid_constant__ CUtensorMap map_pop1, const __grid_constant__ CUtensorMap map_psi) {
  extern __shared__ __align__(128) unsigned char smem[];  // [4 F stages][4 P stages]
  __shared__ TmaCtl ctl;
  constexpr int PB = tma_pstage_bytes<C>(), FB = tma_fstage_bytes<C>();
  static_assert(tma_smem_bytes<C>() <= TMA_SMEM_BUDGET, "stages do not fit");
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < TMA_LANES; i++) {
      mbar_init(smem_u32(&ctl.full_p[i]), 1);
      mbar_init(smem_u32(&ctl.empty_p[i]), 1);
      mbar_init(smem_u32(&ctl.full_f[i]), 1);
      mbar_init(smem_u32(&ctl.empty_f[i]), 2);
    }
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  }
  __syncthreads();

  const int total = w.nbands * w.slots * (w.rp + w.rf);  // tickets == items of TMA_SEG sites; even

  if (warp < TMA_LANES) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;\n" ::"n"(TMA_REGS_PRODUCER));
    // ================================ producer of lane `warp` ================================
    const int ln = warp;
    const uint64_t pol_keep = make_policy<POL_KEEP>();      // P: first touch of the populations, keep for the F items
    const uint64_t pol_stream = make_policy<POL_STREAM>();  // F: last touch
    const uint32_t full_p = smem_u32(&ctl.full_p[ln]), empty_p = smem_u32(&ctl.empty_p[ln]);
    const uint32_t full_f = smem_u32(&ctl.full_f[ln]), empty_f = smem_u32(&ctl.empty_f[ln]);
    const uint32_t base_f = smem_u32(smem) + ln * FB, base_p = smem_u32(smem) + TMA_LANES * FB + ln * PB;
    int kp = 0, kf = 0;  // uses of the lane's P / F stage so far
    // stage the item of ticket `wi` (nothing to do for kind 0)
    auto stage = [&](const WaveItem &wi) {
      if (wi.kind == 0) return;
      const bool isf = wi.kind == 2;
      const uint32_t full = isf ? full_f : full_p, base = isf ? base_f : base_p;
      const int x0 = wi.x0;
      mbar_wait(isf ? empty_f : empty_p, ((isf ? kf : kp) & 1) ^ 1);
      if (lane == 0) {
        int *d = isf ? ctl.desc_f[ln] : ctl.desc_p[ln];
        d[0] = wi.kind; d[1] = wi.zg; d[2] = wi.y; d[3] = x0; d[4] = min(TMA_SEG, g.nx - x0);
        d[5] = wi.cnt + wi.zg * (w.tyb + 2) + wi.row;
        mbar_arrive_tx(full, (uint32_t)(TMA_ROWB * C * (isf ? Q + 9 : Q)));
      }
      __syncwarp();
      if (lane < 9) {
        const int b = lane;
        const int ys = wrap_y(wi.y - tma_box_cy(b), g.ny);
        const int zs = wi.zg - tma_box_cz(b);
        const uint64_t pol = isf ? pol_stream : pol_keep;
        if (b < 5) tma_load_5d(base + b * tma_box3<C>(), &map_pop3, x0 - 2, ys, 3 * b, 0, zs, full, pol);
        else tma_load_5d(base + 5 * tma_box3<C>() + (b - 5) * tma_box1<C>(), &map_pop1, x0 - 2, ys, 10 + b, 0, zs, full, pol);
      } else if (lane < 12 && isf) {
        const int dy = lane - 10;
        tma_load_4d(base + tma_pop_bytes<C>() + (dy + 1) * tma_box3<C>(), &map_psi, x0 - 2, wrap_y(wi.y + dy, g.ny),
                    wi.zg - 1, 0, full);
      }
      if (isf) kf++; else kp++;
    };
    // an F ticket reads psi rows written by P items of this launch: wait for them (acquire), then let the
    // async proxy see them
    auto await_psi = [&](const WaveItem &wi) {
      if (wi.kind != 2) return;
      if (lane < 9) {
        const int *c = wave_wait_counter(g, w, wi, lane);
        if (c) wave_poll_until(c, w.xb, 64);
      }
      __syncwarp();    // the polling lanes leave their loops one by one: ONE converged fence, not one per straggler
      wave_acquire();  // (the warp barrier makes every lane's polls precede every lane's fence)
      asm volatile("fence.proxy.async.global;\n" ::: "memory");
    };
    int t = 0;
    if (lane == 0) t = atomicAdd(w.sync, 2);
    t = __shfl_sync(0xffffffffu, t, 0);
    while (t < total) {
      int nxt = 0;
      if (lane == 0) nxt = atomicAdd(w.sync, 2);  // in flight while this pair is staged
      WaveItem a = wave_decode(g, w, t), b = wave_decode(g, w, t + 1);
      if (a.kind == 2) { const WaveItem c = a; a = b; b = c; }
      stage(a);  // (a P item first: it never waits)
      await_psi(b);
      stage(b);
      t = __shfl_sync(0xffffffffu, nxt, 0);
    }
    // end markers
    if (lane == 0) {
      mbar_wait(empty_p, (kp & 1) ^ 1);
      ctl.desc_p[ln][0] = 3;
      mbar_arrive(full_p);
      mbar_wait(empty_f, (kf & 1) ^ 1);
      ctl.desc_f[ln][0] = 3;
      mbar_arrive(full_f);
    }
  } else if (warp < 2 * TMA_LANES) {
    // ================================ P consumer of lane `warp - 4` ================================
    // psi of the post-advection densities (density.d:89-96, force.d:360-367)
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;\n" ::"n"(TMA_REGS_PCONS));
    const int ln = warp - TMA_LANES;
    const uint32_t full = smem_u32(&ctl.full_p[ln]), empty = smem_u32(&ctl.empty_p[ln]);
    const volatile int *d = ctl.desc_p[ln];
    const double *stage = reinterpret_cast<const double *>(smem + TMA_LANES * FB + ln * PB);
    int npend = 0, mypend = 0;  // lane i < npend holds the counter index of the i-th unpublished item
    auto flush = [&]() {
      __syncwarp();  // the psi stores of all lanes precede (warp barrier, cumulativity) the fences below
      if (lane < npend) {
        asm volatile("fence.proxy.async.global;\n" ::: "memory");
        asm volatile("fence.acq_rel.gpu;\n" ::: "memory");
        asm volatile("red.relaxed.gpu.global.add.s32 [%0], %1;\n" ::"l"(w.sync + mypend), "r"(1) : "memory");
      }
      npend = 0;
    };
    for (int k = 0;; k++) {
      if (!mbar_try_wait(full, k & 1)) {
        flush();  // whatever this warp is about to wait for may, through an F ticket, depend on its own rows
        mbar_wait(full, k & 1);
      }
      if (d[0] == 3) break;
      const int zg = d[1], y = d[2], x0 = d[3], cnt = d[4], cidx = d[5];
#pragma unroll 1
      for (int h = 0; h < 2; h++) {
        const int t = h * 32 + lane;  // two sites per thread, one after the other (56 registers)
        const bool act = t < cnt;
        // the ends of a lattice row: the first / last site's wrapped x neighbour is not in the box
        const bool left = act && x0 + t == 0, right = act && x0 + t == g.nx - 1;
        double rho[C];
#pragma unroll
        for (int c = 0; c < C; c++) {
          double f[Q];
#pragma unroll
          for (int q = 0; q < Q; q++) f[q] = act ? stage[tma_pop_elem<C>(c, q) + t + 2 - cx(q)] : 0.0;
          if (left || right) tma_pull_edge(g, fl.f_old, zg, y, c, left, right, f);
          rho[c] = density19(f);
        }
        if (h == 1) {
          __syncwarp();
          if (lane == 0) mbar_arrive(empty);  // the last densities are in registers: hand the stage back
        }
        if (act) {
          const int sxy = y * g.nx + x0 + t;
#pragma unroll
          for (int c = 0; c < C; c++) {
            const double v = psi_of<PSI>(rho[c]);
            psi_plane(fl.psi, g, zg, c)[sxy] = v;
            if (g.periodic_z_local) {
              if (zg == g.nz) psi_plane(fl.psi, g, 0, c)[sxy] = v;      // -> ghost plane 0
              if (zg == 1) psi_plane(fl.psi, g, g.nz + 1, c)[sxy] = v;  // -> ghost plane nz+1
            }
          }
        }
      }
      if (lane == npend) mypend = cidx;
      if (++npend == 4) flush();
    }
    flush();
  } else {
    // ================================ F consumers of lane `(warp - 8) % 4` ================================
    // pull + Shan-Chen force + collide (advection.d, force.d:238-309, collision.d:90-119), half an item per warp
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;\n" ::"n"(TMA_REGS_FCONS));
    // the two warps of a lane sit on the same SM sub-partition (warp index mod 4) and wake at the same barrier:
    // they run the same long straight-line code side by side and share its instruction-cache lines
    const int ln = (warp - 2 * TMA_LANES) & 3;
    const int t = ((warp - 2 * TMA_LANES) >> 2) * 32 + lane;  // site within the item
    const uint32_t full = smem_u32(&ctl.full_f[ln]), empty = smem_u32(&ctl.empty_f[ln]);
    const volatile int *d = ctl.desc_f[ln];
    const double *sp = reinterpret_cast<const double *>(smem + ln * FB) + t;
    for (int k = 0;; k++) {
      mbar_wait(full, k & 1);
      if (d[0] == 3) break;
      const int zg = d[1], y = d[2], x0 = d[3], cnt = d[4];
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
      if (lane == 0) mbar_arrive(empty);  // operands are in registers: hand the stage back
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

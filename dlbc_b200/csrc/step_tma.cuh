// step_tma.cuh -- kernel variant DLBC_KERNEL_TMA: the wave schedule of step_wave.cuh (psi items P and
// fused items F of one launch, ticket-ordered band by band so that the second gather of every population
// is an L2 hit) with the data movement taken off the compute warps.
//
// One persistent CTA per SM, warp-specialised:
//   * warp 0 is the PRODUCER.  It takes tickets, decodes them into items (a row segment of TMA_SEG sites),
//     waits -- for F items -- until the nine psi row segments the item reads are complete, and then stages
//     everything the item reads into a ring of shared-memory stages with TMA bulk copies
//     (cp.async.bulk.shared.global, one per lattice row segment, completion counted on an mbarrier):
//       - the C*19 population rows of the pull gather: row (c, q) of plane z - c_z(q), lattice row
//         y - c_y(q), shifted by c_x(q) -- each a contiguous run of the plane-major SoA layout (lattice.cuh);
//       - F items: the 9 psi rows (planes z-1..z+1, rows y-1..y+1) of every component;
//       - MASK: the neighbourhood solid bits of the segment.
//     Loads in flight occupy shared memory, not registers, and do not depend on what the compute warps are
//     doing: ~5 stages of 30 KB per SM stay in flight (C = 2), against 76 registers x 512 threads before.
//   * warps 1.. are CONSUMERS in groups of TMA_SEG threads.  A group claims the next staged item, reads its
//     operands from shared memory with immediate-offset LDS (no per-access address arithmetic), hands the
//     stage back, and runs the same per-site functions as every other variant (lb_site.cuh: identical
//     bits), storing psi (P) or the post-collision populations (F) straight from registers.
//
// Ordering: as in step_wave.cuh -- a P item publishes completion per (band, plane, row) with
// red.release.gpu after a group barrier; the producer polls the nine counters of an F item (relaxed loads,
// then one fence.acq_rel.gpu) BEFORE issuing the item's psi copies, followed by fence.proxy.async: the
// copies are async-proxy reads of data written through the generic proxy.  The producer only waits for
// items with smaller tickets, and consumers never wait for anything but copies already issued, so the
// launch cannot deadlock whatever the number of resident CTAs.
//
// Needs nx even (16-byte copy granularity; nx % 4 == 0 with a mask): other sizes take the register path.
#pragma once
#include "step_wave.cuh"

namespace dlbc {

#ifndef DLBC_TMA_SEG
#define DLBC_TMA_SEG 64
#endif
#ifndef DLBC_TMA_GROUPS
#define DLBC_TMA_GROUPS 4
#endif
constexpr int TMA_SEG = DLBC_TMA_SEG;        // sites per item == threads per consumer group
constexpr int TMA_ROW = TMA_SEG + 4;         // doubles per staged row: element i <-> lattice x = x0 - 2 + i
constexpr int TMA_ROWB = TMA_ROW * 8;
constexpr int TMA_GROUPS = DLBC_TMA_GROUPS;  // consumer groups per CTA
constexpr int TMA_GW = TMA_SEG / 32;         // warps per group
constexpr int TMA_THREADS = 128 + TMA_GROUPS * TMA_SEG;  // warpgroup 0: the producer warp (+ 3 warps that only give up their registers)
static_assert((TMA_GROUPS * TMA_SEG) % 128 == 0, "consumers must fill whole warpgroups (setmaxnreg is per warpgroup)");
// register split (setmaxnreg): the kernel is compiled for 65536 / TMA_THREADS registers per thread; warpgroup 0
// shrinks to TMA_REGS_PRODUCER and the consumer warpgroups grow to TMA_REGS_CONSUMER (19 C populations +
// force + moments per thread must stay in registers: a spill would go through an L1 that the stages left ~10 KB of)
constexpr int TMA_REGS_PRODUCER = 56;
constexpr int TMA_REGS_CONSUMER = ((65536 - 128 * TMA_REGS_PRODUCER) / (TMA_GROUPS * TMA_SEG)) / 8 * 8 > 240
                                      ? 240 : ((65536 - 128 * TMA_REGS_PRODUCER) / (TMA_GROUPS * TMA_SEG)) / 8 * 8;
constexpr int TMA_MAX_STAGES = 16;
constexpr int TMA_SMEM_BUDGET = 227 * 1024 - 2048;  // dynamic shared memory available next to the static control block

// stage layout: [C*19 population rows][C*9 psi rows][TMA_SEG neighbourhood words]
template <int C>
__host__ __device__ constexpr int tma_stage_bytes() { return (C * (Q + 9) * TMA_ROWB + TMA_SEG * 4 + 127) / 128 * 128; }
template <int C>
__host__ __device__ constexpr int tma_stages() {
  return TMA_SMEM_BUDGET / tma_stage_bytes<C>() < TMA_MAX_STAGES ? TMA_SMEM_BUDGET / tma_stage_bytes<C>() : TMA_MAX_STAGES;
}

struct TmaCtl {
  unsigned long long full[TMA_MAX_STAGES];   // producer arrives with expect_tx, the bulk copies complete_tx
  unsigned long long empty[TMA_MAX_STAGES];  // the consumer group's warps arrive once their operands are in registers
  int desc[TMA_MAX_STAGES][8];               // kind (1 P, 2 F, 3 end), zg, y, x0, n, counter index
  int claim;                                 // next item a consumer group takes
  int bcast[TMA_GROUPS][2];
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
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}\n" ::"r"(bar), "r"(parity) : "memory");
}
// global -> shared bulk copy (SASS UBLKCP), completion in bytes on an mbarrier
template <bool HINT>
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar, uint64_t pol) {
  if (HINT)
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;\n"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar), "l"(pol) : "memory");
  else
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void group_bar(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;\n" ::"r"(id), "r"(nthreads) : "memory");
}

// Stage one row segment: staged element i <-> lattice x = x0 - 2 + i (periodic in x); elements 2 .. n+1 always,
// the two below (lo) / above (hi) only when the row is read shifted by +1 / -1.  One copy when the run is
// contiguous in memory, one more 16-byte copy for a pad that wraps around the lattice.  n, x0, nx even.
template <bool HINT>
__device__ __forceinline__ void tma_row(uint32_t dst, const double *src_row, int nx, int x0, int n, bool lo, bool hi,
                                        uint32_t bar, uint64_t pol) {
  int xs = x0, cnt = n;
  uint32_t d = dst + 16;
  if (lo) {
    if (x0 >= 2) { xs -= 2; cnt += 2; d -= 16; }
    else bulk_g2s<HINT>(dst, src_row + (nx - 2), 16, bar, pol);
  }
  if (hi) {
    if (x0 + n + 2 <= nx) cnt += 2;
    else bulk_g2s<HINT>(dst + 16 + n * 8, src_row, 16, bar, pol);
  }
  bulk_g2s<HINT>(d, src_row + xs, cnt * 8, bar, pol);
}

// c_x / c_y / c_z of a run-time q (the constexpr tables of lb_site.cuh are for unrolled loops): 2-bit codes
__host__ __device__ constexpr uint64_t pack_codes(int (*fn)(int)) {
  uint64_t m = 0;
  for (int q = 0; q < Q; q++) m |= (uint64_t)(fn(q) + 1) << (2 * q);
  return m;
}
__device__ __forceinline__ int cx_rt(int q) { constexpr uint64_t m = pack_codes(cx); return (int)((m >> (2 * q)) & 3) - 1; }
__device__ __forceinline__ int cy_rt(int q) { constexpr uint64_t m = pack_codes(cy); return (int)((m >> (2 * q)) & 3) - 1; }
__device__ __forceinline__ int cz_rt(int q) { constexpr uint64_t m = pack_codes(cz); return (int)((m >> (2 * q)) & 3) - 1; }

__device__ __forceinline__ int wrap_y(int y, int ny) { return y < 0 ? y + ny : (y >= ny ? y - ny : y); }

// ---- consumer side: operands from a stage -----------------------------------------------------------
// sp: the stage's population rows + this thread's site index, i.e. element [row][t] of [C*19][TMA_ROW]
// lb/advection.d:61-89 from the staged rows: f[q] = old[p - c_q][q] sits at staged x index t + 2 - c_x(q);
// a bounced link (bit bounce(q) of the neighbourhood word) takes old[p][bounce q] from global memory instead
// (that value is streamed by nobody: its downstream site is solid).
template <bool MASK>
__device__ __forceinline__ void tma_pull19(const double *sp, int c, uint32_t bits, const double *ps_global, const Geom &g,
                                           double (&f)[Q]) {
#pragma unroll
  for (int q = 0; q < Q; q++) {
    if (MASK && q > 0 && ((bits >> bounce(q)) & 1u)) f[q] = ld_global_nc(ptr_at(ps_global, bounce(q) * g.plane));
    else f[q] = sp[(c * Q + q) * TMA_ROW + 2 - cx(q)];
  }
}

template <int C, int EQ, int PSI, bool MASK, bool EXACT>
__global__ void __launch_bounds__(TMA_THREADS, 1) tma_kernel(Geom g, Fields fl, StepParams P, Wave w) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ TmaCtl ctl;
  constexpr int NS = tma_stages<C>();
  constexpr int SB = tma_stage_bytes<C>();
  constexpr int PSI_OFF = C * Q * TMA_ROWB;
  constexpr int BITS_OFF = C * (Q + 9) * TMA_ROWB;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < NS; i++) {
      mbar_init(smem_u32(&ctl.full[i]), 1);
      mbar_init(smem_u32(&ctl.empty[i]), TMA_GW);
    }
    ctl.claim = 0;
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  }
  __syncthreads();

  const int total = w.nbands * w.slots * (w.rp + w.rf);

  if (warp < 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;\n" ::"n"(TMA_REGS_PRODUCER));
    if (warp != 0) return;
    // ================================ producer ================================
    const uint64_t pol_keep = make_policy<POL_KEEP>();      // P: first touch of the populations, keep for the F items
    const uint64_t pol_stream = make_policy<POL_STREAM>();  // F: last touch
    int t = 0;
    if (lane == 0) t = atomicAdd(w.sync, 1);
    t = __shfl_sync(0xffffffffu, t, 0);
    int n = 0;  // items pushed so far
    while (t < total) {
      int nxt = 0;
      if (lane == 0) nxt = atomicAdd(w.sync, 1);  // in flight while this item is staged
      const WaveItem wi = wave_decode(g, w, t);
      if (wi.kind != 0) {
        const int st = n % NS;
        const uint32_t full = smem_u32(&ctl.full[st]);
        const uint32_t base = smem_u32(smem) + st * SB;
        const int cnt = min(TMA_SEG, g.nx - wi.x0);  // sites of this (possibly ragged) segment
        mbar_wait(smem_u32(&ctl.empty[st]), ((n / NS) & 1) ^ 1);
        if (wi.kind == 2) {
          if (lane < 9) {
            const int *c = wave_wait_counter(g, w, wi, lane);
            if (c) while (ld_poll(c) < w.xb) __nanosleep(64);
            wave_acquire();
          }
          __syncwarp();
          asm volatile("fence.proxy.async.global;\n" ::: "memory");  // generic-proxy psi stores -> async-proxy reads
        }
        if (lane == 0) {
          int *d = ctl.desc[st];
          d[0] = wi.kind; d[1] = wi.zg; d[2] = wi.y; d[3] = wi.x0; d[4] = cnt; d[5] = wi.cnt + wi.zg * (w.tyb + 2) + wi.row;
          uint32_t tx = (uint32_t)C * (Q * cnt * 8 + 160);
          if (wi.kind == 2) tx += (uint32_t)C * (9 * cnt * 8 + 160);
          if (MASK) tx += (uint32_t)cnt * 4;
          mbar_arrive_tx(full, tx);
        }
        __syncwarp();
        const uint64_t pol = wi.kind == 1 ? pol_keep : pol_stream;
        for (int r = lane; r < C * Q; r += 32) {
          const int c = r / Q, q = r - c * Q;
          const int ys = wrap_y(wi.y - cy_rt(q), g.ny);
          const double *src = fl.f_old + (int64_t)(wi.zg - cz_rt(q)) * g.fz + (int64_t)c * g.fc + (int64_t)q * g.plane +
                              (int64_t)ys * g.nx;
          const int sx = cx_rt(q);
          tma_row<true>(base + r * TMA_ROWB, src, g.nx, wi.x0, cnt, sx == 1, sx == -1, full, pol);
        }
        if (wi.kind == 2) {
          for (int r = lane; r < C * 9; r += 32) {
            const int c = r / 9, k = r - c * 9;
            const int dz = k / 3 - 1, dy = k - (k / 3) * 3 - 1;
            const int ys = wrap_y(wi.y + dy, g.ny);
            const double *src = fl.psi + (int64_t)c * g.sites_g + (int64_t)(wi.zg + dz) * g.plane + (int64_t)ys * g.nx;
            const bool pads = dy == 0 || dz == 0;  // corner rows are only read at dx = 0
            tma_row<false>(base + PSI_OFF + r * TMA_ROWB, src, g.nx, wi.x0, cnt, pads, pads, full, 0);
          }
        }
        if (MASK && lane == 31)
          bulk_g2s<false>(base + BITS_OFF, fl.nb + (int64_t)wi.zg * g.plane + (int64_t)wi.y * g.nx + wi.x0, cnt * 4, full, 0);
        n++;
      }
      t = __shfl_sync(0xffffffffu, nxt, 0);
    }
    // one end marker per consumer group
    if (lane == 0) {
      for (int e = 0; e < TMA_GROUPS; e++, n++) {
        const int st = n % NS;
        mbar_wait(smem_u32(&ctl.empty[st]), ((n / NS) & 1) ^ 1);
        ctl.desc[st][0] = 3;
        mbar_arrive(smem_u32(&ctl.full[st]));
      }
    }
  } else {
    // ================================ consumers ================================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;\n" ::"n"(TMA_REGS_CONSUMER));
    const int grp = (threadIdx.x - 128) / TMA_SEG;
    const int t = (threadIdx.x - 128) - grp * TMA_SEG;  // site within the item
    for (int iter = 0;; iter ^= 1) {
      if (t == 0) ctl.bcast[grp][iter] = atomicAdd(&ctl.claim, 1);
      group_bar(1 + grp, TMA_SEG);
      const int n = ctl.bcast[grp][iter];
      const int st = n % NS;
      mbar_wait(smem_u32(&ctl.full[st]), (n / NS) & 1);
      const int *d = ctl.desc[st];
      const int kind = d[0];
      if (kind == 3) break;
      const int zg = d[1], y = d[2], x0 = d[3], cnt = d[4], cidx = d[5];
      const unsigned char *stage = smem + st * SB;
      const double *sp = reinterpret_cast<const double *>(stage) + t;
      const uint32_t empty = smem_u32(&ctl.empty[st]);
      const bool in = t < cnt;
      const int sxy = y * g.nx + x0 + t;
      uint32_t bits = 0u;
      if (MASK && in) bits = reinterpret_cast<const uint32_t *>(stage + BITS_OFF)[t];
      const bool act = in && !(MASK && (bits & 1u));  // solid sites: nothing to do (advection.d:83-85, force.d:273)
      const double *psg = f_plane(fl.f_old, g, zg, 0) + sxy;

      if (kind == 1) {
        // ---- P item: psi of the post-advection densities (density.d:89-96, force.d:360-367)
        double f[C][Q];
        if (act) {
#pragma unroll
          for (int c = 0; c < C; c++) tma_pull19<MASK>(sp, c, bits, psg + c * g.fc, g, f[c]);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty);  // operands are in registers: hand the stage back
        if (act) {
#pragma unroll
          for (int c = 0; c < C; c++) {
            const double v = psi_of<PSI>(density19(f[c]));
            psi_plane(fl.psi, g, zg, c)[sxy] = v;
            if (g.periodic_z_local) {
              if (zg == g.nz) psi_plane(fl.psi, g, 0, c)[sxy] = v;      // -> ghost plane 0
              if (zg == 1) psi_plane(fl.psi, g, g.nz + 1, c)[sxy] = v;  // -> ghost plane nz+1
            }
          }
        }
        group_bar(1 + grp, TMA_SEG);  // all psi stores of the group precede (CTA scope) the release below
        if (t == 0) {
          asm volatile("fence.proxy.async.global;\n" ::: "memory");
          red_release_add(w.sync + cidx, 1);
        }
      } else {
        // ---- F item: pull + Shan-Chen force + collide (advection.d, force.d:238-309, collision.d:90-119)
        double f[C][Q], F[C][3], psi1[C];
#pragma unroll
        for (int c = 0; c < C; c++) { F[c][0] = 0.0; F[c][1] = 0.0; F[c][2] = 0.0; psi1[c] = 0.0; }
        if (act) {
          // the force first: its 18 C psi operands are dead again before the 19 C populations are loaded
          const double *sq = reinterpret_cast<const double *>(stage + PSI_OFF) + t + 2;
#pragma unroll
          for (int c = 0; c < C; c++) psi1[c] = sq[(c * 9 + 4) * TMA_ROW];
          sc_force<C, EXACT>(P, psi1, F, [&](int f2, int q) -> double {
            if (MASK && ((bits >> q) & 1u)) return psi1[f2];
            return sq[(f2 * 9 + (cz(q) + 1) * 3 + (cy(q) + 1)) * TMA_ROW + cx(q)];
          });
#pragma unroll
          for (int c = 0; c < C; c++) tma_pull19<MASK>(sp, c, bits, psg + c * g.fc, g, f[c]);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty);
        if (act) {
          if (MASK && P.any_gwc) wall_force<C, true>(g, fl, P, zg, make_nbr(g, x0 + t, y), bits, psi1, F);
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
}

}  // namespace dlbc

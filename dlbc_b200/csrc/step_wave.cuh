// step_wave.cuh -- kernel variant DLBC_KERNEL_WAVE: the two-pass timestep (psi sites + fused
// pull/force/collide sites, step_two_pass.cuh) as ONE launch whose blocks are ordered so that
// the second read of every population is an L2 hit instead of an HBM read.
//
// Why: psi at the 18 neighbours needs the post-advection density there, so the populations are
// gathered twice per step.  As two kernels the second gather comes from HBM again
// (472 C + 1 = 945 B per site update at C = 2).  The 126 MB L2 can hold the populations of a few
// tens of thousands of sites, so if the psi work of a piece of the lattice runs just before the
// collide work of the same piece, the second gather hits L2 and HBM traffic drops to
// ~(304 + 16) C + 1 bytes.  A whole 512 x 512 plane is 80 MB of populations -- too big -- so the
// lattice is cut into y-bands of TYB rows; the launch walks band by band, and within a band plane
// by plane along z, interleaving
//     P(p):  psi of plane p for rows yb-1 .. yb+TYB (the band plus a one-row halo, recomputed by
//            the neighbouring band too: 2/TYB extra psi work, deterministic, identical bits)
//     F(z):  fused pull + Shan-Chen force + collide of plane z, rows yb .. yb+TYB-1
// with F(z) scheduled LA planes behind P(z + LA) (band-plane = TYB * nx sites, ~5-10 MB).
//
// Ordering without a grid barrier: every block takes a ticket (atomicAdd) when it starts and
// derives its work item from the ticket, so "earlier ticket" == "already running or finished".
// P blocks publish completion per (band, plane) with a release reduction; an F block polls
// until the three psi planes it reads are complete.  It only ever waits for earlier tickets,
// hence no deadlock whatever the number of resident blocks.
//
// No shared memory and no block barrier: the same latency hiding as the stand-alone fused kernel
// (87 % of the measured copy bandwidth).
// Bit-identical to the other variants: the same per-site functions.
#pragma once
#include "lattice.cuh"
#include "step_two_pass.cuh"

namespace dlbc {

struct Wave {
  int z_begin, z_end;  // F planes [z_begin, z_end) (ghost-inclusive indices, 1..nz)
  int pl, ph;          // P planes of the main sequence [pl, ph]
  int pre;             // periodic slab: slot 0 computes P(nz) (which also fills ghost plane 0)
  int tyb;             // band height (rows)
  int la;              // F lags P by `la` slots
  int xb;              // blocks per row
  int rp, rf;          // P / F blocks per slot: (tyb + 2) * xb, tyb * xb
  int slots;           // slots per band
  int nbands;
  int pf;              // L2 prefetch distance in tickets (0 = off): see wave_prefetch
  int pf_keep;         // prefetch with L2::evict_last (else default priority)
  int batch;           // P items gather all components before summing (psi_site BATCH)
  int *sync;           // [0] ticket, [1 + band * (nz + 2) + plane] completed P blocks
};

// Polling load of a completion counter.  relaxed.gpu (LDG.STRONG.GPU) rather than acquire.gpu: an
// acquire makes ptxas emit CCTL.IVALL, which wipes the whole SM's L1 at every poll.  No stale L1
// line of a psi plane can exist -- L1 is invalidated at launch, and a psi line is only ever loaded
// by an F block after its counter is complete -- so ordering is all that is needed, and the
// branch on the polled value plus the __syncthreads that follows provide it.
__device__ __forceinline__ int ld_poll(const int *p) {
  int v;
  asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ void red_release_add(int *p, int v) {
  asm volatile("red.release.gpu.global.add.s32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}

constexpr int WAVE_SEG = 128;  // sites per item == threads per block
#ifndef WAVE_PREFETCH
#define WAVE_PREFETCH true
#endif

// c_y / c_z of a run-time q (the constexpr tables of lb_site.cuh are for unrolled loops): 2-bit codes
__host__ __device__ constexpr uint64_t pack_codes(int (*fn)(int)) {
  uint64_t m = 0;
  for (int q = 0; q < Q; q++) m |= (uint64_t)(fn(q) + 1) << (2 * q);
  return m;
}
__device__ __forceinline__ int cy_rt(int q) { constexpr uint64_t m = pack_codes(cy); return (int)((m >> (2 * q)) & 3) - 1; }
__device__ __forceinline__ int cz_rt(int q) { constexpr uint64_t m = pack_codes(cz); return (int)((m >> (2 * q)) & 3) - 1; }

// DRAM -> L2 prefetch of everything the P item with ticket tp will gather: C*19 row segments of
// WAVE_SEG doubles, one prefetch per 128-byte line, spread over the block's threads.  A P item is
// pure streaming (38 loads, 38 adds, one exp per site) and runs at 16 warps/SM because it shares
// the launch with the 128-register F items: without the prefetch its warps sat on two full DRAM
// latencies per site (35 % of the kernel's warp time, profiles/r01_wave_before.md).  Issued one
// round of tickets ahead, the prefetch costs no registers and no scoreboard, and the P item's own
// loads then hit L2.  The x +- 1 shift of the gather stays inside the same lines except for one
// element at either end, which belongs to the neighbouring item's segment.
template <int C>
__device__ __forceinline__ void wave_prefetch(const Geom &g, const Fields &fl, const Wave &w, int tp,
                                              int per_band, int per_slot) {
  const int band = tp / per_band;
  const int r = tp - band * per_band;
  const int k = r / per_slot;
  const int j = r - k * per_slot;
  if (j >= w.rp) return;
  int p = -1;
  if (k == 0) { if (w.pre) p = g.nz; }
  else if (w.pl + (k - 1) <= w.ph) p = w.pl + (k - 1);
  if (p < 0) return;
  const int yb = band * w.tyb;
  const int tb = min(w.tyb, g.ny - yb);
  const int row = j / w.xb, xblk = j - row * w.xb;
  if (row > tb + 1) return;
  int y = yb - 1 + row;
  y = y < 0 ? y + g.ny : (y >= g.ny ? y - g.ny : y);
  const int x0 = xblk * WAVE_SEG;
  const int lpr = (min(WAVE_SEG, g.nx - x0) + 15) >> 4;  // 128-byte lines per row segment
  const int nlines = C * Q * lpr;
  for (int l = threadIdx.x; l < nlines; l += WAVE_SEG) {
    const int cq = l / lpr, s = l - cq * lpr;
    const int c = cq / Q, q = cq - c * Q;
    int ys = y - cy_rt(q);
    ys = ys < 0 ? ys + g.ny : (ys >= g.ny ? ys - g.ny : ys);
    const double *a = fl.f_old + (int64_t)(p - cz_rt(q)) * g.fz + (int64_t)c * g.fc + (int64_t)q * g.plane +
                      ys * g.nx + x0 + s * 16;
    if (w.pf_keep) asm volatile("prefetch.global.L2::evict_last [%0];\n" ::"l"(a));
    else asm volatile("prefetch.global.L2 [%0];\n" ::"l"(a));
  }
}

// Persistent 128-thread blocks (4 per SM): each block loops over tickets; an item is a 128-site row
// segment.  The atomicAdd for the NEXT ticket is issued before the current item's work and
// consumed after it, so its L2 round trip is hidden.  A block working on ticket t only ever waits
// for items with smaller tickets, each of which is being worked on by a resident block or is the
// prefetched ticket of a block that is itself working on a still smaller one: the chain ends at
// a running item, so there is no deadlock.
// (Measured alternatives, profiles/README.md: single-warp blocks with 128-site items triple the
// number of sites in flight and overflow L2; single-warp blocks with 32-site items need a twice
// longer lag before the F items stop waiting.  Both were slower.)

template <int C, int EQ, int PSI, bool MASK, bool EXACT>
__global__ void __launch_bounds__(WAVE_SEG, (MASK ? 3 : 4)) wave_kernel(Geom g, Fields fl, StepParams P, Wave w) {
  __shared__ int s_ticket[2];
  const int per_slot = w.rp + w.rf;
  const int per_band = w.slots * per_slot;
  const int total = w.nbands * per_band;
  if (threadIdx.x == 0) s_ticket[0] = atomicAdd(w.sync, 1);
  __syncthreads();
  int t = s_ticket[0];
  for (int it = 1; t < total; it ^= 1) {
    int nxt = 0;
    if (threadIdx.x == 0) nxt = atomicAdd(w.sync, 1);  // in flight during this item's work
    if (WAVE_PREFETCH && w.pf > 0 && t + w.pf < total) wave_prefetch<C>(g, fl, w, t + w.pf, per_band, per_slot);

    const int band = t / per_band;
    const int r = t - band * per_band;
    const int k = r / per_slot;
    const int j = r - k * per_slot;
    const int yb = band * w.tyb;
    const int tb = min(w.tyb, g.ny - yb);   // rows of this (possibly ragged) band
    int *cnt = w.sync + 1 + band * (g.nz + 2);

    if (j < w.rp) {
      // ---- P item: psi of one row segment of plane p
      int p = -1;
      if (k == 0) { if (w.pre) p = g.nz; }
      else if (w.pl + (k - 1) <= w.ph) p = w.pl + (k - 1);
      if (p >= 0) {
        const int row = j / w.xb, xblk = j - row * w.xb;
        const int x = xblk * WAVE_SEG + threadIdx.x;
        if (row <= tb + 1 && x < g.nx) {
          int y = yb - 1 + row;
          y = y < 0 ? y + g.ny : (y >= g.ny ? y - g.ny : y);
          if (w.batch) psi_site<C, PSI, MASK, POL_KEEP, true>(g, fl, x, y, p);
          else psi_site<C, PSI, MASK, POL_KEEP, false>(g, fl, x, y, p);
        }
        __syncthreads();  // all psi stores of the block precede (CTA scope) thread 0's release below
        if (threadIdx.x == 0) red_release_add(cnt + p, 1);
      }
    } else if (k > 0) {
      // ---- F item: fused update of one row segment of plane zf
      const int zf = w.z_begin + (k - 1) - w.la;
      const int jf = j - w.rp;
      const int row = jf / w.xb, xblk = jf - row * w.xb;
      if (zf >= w.z_begin && zf < w.z_end && row < tb) {
        if (threadIdx.x < 3) {
          int d = zf + (int)threadIdx.x - 1;
          bool need = true;
          if (d < 1) { need = g.periodic_z_local; d = g.nz; }
          else if (d > g.nz) { need = g.periodic_z_local; d = 1; }
          if (need) {
            const int *c = cnt + d;
            while (ld_poll(c) < w.rp) __nanosleep(32);
          }
        }
        __syncthreads();
        const int x = xblk * WAVE_SEG + threadIdx.x;
        if (x < g.nx) fused_site<C, EQ, true, MASK, EXACT, true, POL_STREAM>(g, fl, P, x, yb + row, zf);
      }
    }

    if (threadIdx.x == 0) s_ticket[it] = nxt;
    __syncthreads();
    t = s_ticket[it];
  }
}

}  // namespace dlbc

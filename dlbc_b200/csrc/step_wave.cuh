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
//     P(p):  psi of plane p for rows yb-1 .. yb+TYB (the band plus a one-row halo); the two rows a band
//            shares with its predecessor are computed by the predecessor only (DLBC_WAVE_SHARE_HALO)
//     F(z):  fused pull + Shan-Chen force + collide of plane z, rows yb .. yb+TYB-1
// with F(z) scheduled LA slots behind P(z), i.e. in the slot of P(z + LA) (band-plane = TYB * nx sites,
// ~5-10 MB).
//
// Ordering without a grid barrier: every block takes a ticket (atomicAdd) when it starts and
// derives its work item from the ticket, so "earlier ticket" == "already running or finished".
// P items publish completion per (band, plane, row) with a release reduction; an F item polls
// until the nine psi row segments it reads are complete.  It only ever waits for earlier tickets,
// hence no deadlock whatever the number of resident blocks.
//
// This file holds the schedule (struct Wave, the ticket -> item mapping) and the register-path
// kernel (per-thread LDG gathers).  step_tma.cuh runs the same schedule with the populations and
// psi rows moved by TMA bulk copies into shared memory from a producer warp.
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
  int seg;             // sites per item (row segment length)
  int xb;              // tickets per row
  int alt;             // P and F tickets of a slot alternate row by row (TMA variant)
  int nsub;            // items per ticket (TMA variant: a ticket is a whole row cut into TMA_SEG-site items; else 1)
  int rp, rf;          // P / F items per slot: (tyb + 2) * xb, tyb * xb
  int slots;           // slots per band
  int nbands;
  int *sync;           // [0] ticket, [1 + ((band * (nz + 2) + plane) * (tyb + 2) + row)] completed P items of that row
};

// One work item of the launch, decoded from its ticket (block-uniform).
struct WaveItem {
  int kind;   // 0 nothing to do, 1 P (psi of a row segment), 2 F (fused update of a row segment)
  int zg;     // ghost-inclusive plane
  int y;      // lattice row (P: wrapped)
  int x0;     // first site of the segment
  int row;    // row index within the band: P 0..tb+1 (halo rows included), F 0..tb-1
  int cnt;    // index into Wave::sync of the band's counter block: counter(plane, row) = sync[cnt + plane * (tyb + 2) + row]
  int band, tb;  // band index and its height in rows (the last band may be ragged)
};

// DLBC_WAVE_SHARE_HALO: psi rows that two neighbouring bands both need (a band's P rows 0 and 1 are the previous
// band's rows tyb and tyb + 1; the last band's upper halo row is band 0's row 1) are computed ONCE, by the band
// that comes first in ticket order; the later band's tickets for them are empty and its F items wait on the
// earlier band's counters instead.  Saves 2 of tyb + 2 psi rows per band (their 19 C gathers from DRAM).
// 512^3: 19.25 -> 19.12 ms per step, porous 23.65 -> 23.46 (profiles/r02/ab_share_halo.txt).
#ifndef DLBC_WAVE_SHARE_HALO
#define DLBC_WAVE_SHARE_HALO 1
#endif

__device__ __forceinline__ WaveItem wave_decode(const Geom &g, const Wave &w, int t) {
  const int per_slot = w.rp + w.rf;
  const int per_band = w.slots * per_slot;
  const int band = t / per_band;
  const int r = t - band * per_band;
  const int k = r / per_slot;
  int j = r - k * per_slot;
  const int yb = band * w.tyb;
  const int tb = min(w.tyb, g.ny - yb);  // rows of this (possibly ragged) band
  WaveItem it;
  it.kind = 0;
  it.band = band;
  it.tb = tb;
  it.cnt = 1 + band * (g.nz + 2) * (w.tyb + 2);
  // order within a slot: all P items, then all F items; or (w.alt, one ticket per row) P row r, F row r, ...
  // so that both kinds of work are always in flight (nothing in a slot depends on anything else in it)
  int jp = j, jf = j - w.rp;
  bool is_p = j < w.rp;
  if (w.alt) {
    is_p = j >= 2 * w.rf || (j & 1) == 0;
    jp = j >= 2 * w.rf ? j - w.rf : j >> 1;
    jf = j >> 1;
  }
  if (is_p) {
    j = jp;
    int p = -1;
    if (k == 0) { if (w.pre) p = g.nz; }
    else if (w.pl + (k - 1) <= w.ph) p = w.pl + (k - 1);
    const int row = j / w.xb, xblk = j - row * w.xb;
    const bool shared = DLBC_WAVE_SHARE_HALO && w.tyb >= 2 && ((band > 0 && row <= 1) || (w.nbands > 1 && band == w.nbands - 1 && row == tb + 1));
    if (p >= 0 && row <= tb + 1 && !shared) {
      int y = yb - 1 + row;
      y = y < 0 ? y + g.ny : (y >= g.ny ? y - g.ny : y);
      it.kind = 1; it.zg = p; it.y = y; it.x0 = xblk * w.seg; it.row = row;
    }
  } else if (k > 0) {
    const int zf = w.pl + (k - 1) - w.la;  // `la` slots behind the P item of the same plane (pl < z_begin on slabs)
    const int row = jf / w.xb, xblk = jf - row * w.xb;
    if (zf >= w.z_begin && zf < w.z_end && row < tb) {
      it.kind = 2; it.zg = zf; it.y = yb + row; it.x0 = xblk * w.seg; it.row = row;
    }
  }
  return it;
}

// The i-th (0..8) completion counter an F item waits for: psi of planes zf-1..zf+1, rows y-1..y+1 (P rows
// row..row+2 of the band).  Ghost planes are written by the P items of the plane they mirror (periodic slab)
// or were received before the launch (z-slab ring): returns nullptr when there is nothing to wait for.
__device__ __forceinline__ const int *wave_wait_counter(const Geom &g, const Wave &w, const WaveItem &it, int i) {
  int d = it.zg + i / 3 - 1;
  if (d < 1) { if (!g.periodic_z_local) return nullptr; d = g.nz; }
  else if (d > g.nz) { if (!g.periodic_z_local) return nullptr; d = 1; }
  int cnt = it.cnt, prow = it.row + i % 3;  // P row index within the band: F row r reads P rows r, r + 1, r + 2
  if (DLBC_WAVE_SHARE_HALO && w.tyb >= 2) {  // (bands of one row: the previous band's rows 1, 2 would themselves be shared)
    if (it.band > 0 && prow <= 1) { cnt -= (g.nz + 2) * (w.tyb + 2); prow += w.tyb; }                    // previous band's copy
    else if (w.nbands > 1 && it.band == w.nbands - 1 && prow == it.tb + 1) { cnt = 1; prow = 1; }   // band 0's row y = 0
  }
  return w.sync + cnt + d * (w.tyb + 2) + prow;
}

// Polling load of a completion counter (relaxed: an acquire load makes ptxas emit CCTL.IVALL, an L1 wipe, at
// EVERY poll) followed, once the wait is over, by ONE acquire fence (wave_acquire): together they form the
// acquire side of the red.release.gpu the P items publish with, so the psi rows they stored are visible to
// whatever the polling thread -- and, through the block barrier that follows, the block -- reads afterwards.
// The psi loads themselves bypass L1 (ld.global.cg in shan_chen_force / fused_site, COHERENT mode): a
// 32-byte sector can straddle a row boundary, so a line another item pulled into L1 may hold last step's
// values of a row that was not complete then.
// Watchdog of every spin wait of the single-launch kernels: a protocol error must end the launch with an error
// (trap -> "unspecified launch failure" at the next synchronisation) instead of hanging the GPU.
constexpr long long WAVE_WATCHDOG_CYCLES = 20000000000ll;  // ~10 s at 2 GHz; a legitimate wait is microseconds
struct Watchdog {
  long long t0 = 0;
  __device__ __forceinline__ void tick() {
    const long long now = clock64();
    if (t0 == 0) t0 = now;
    else if (now - t0 > WAVE_WATCHDOG_CYCLES) __trap();
  }
};
// wait until *p >= target (relaxed polls; the caller adds the acquire fence)
__device__ __forceinline__ int ld_poll(const int *p);
__device__ __forceinline__ void wave_poll_until(const int *p, int target, unsigned ns) {
  Watchdog wd;
  while (ld_poll(p) < target) { __nanosleep(ns); wd.tick(); }
}
__device__ __forceinline__ int ld_poll(const int *p) {
  int v;
  asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void wave_acquire() { asm volatile("fence.acq_rel.gpu;\n" ::: "memory"); }

__device__ __forceinline__ void red_release_add(int *p, int v) {
  asm volatile("red.release.gpu.global.add.s32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}

#ifndef DLBC_WAVE_EARLY_GATHER
#define DLBC_WAVE_EARLY_GATHER 0
#endif
constexpr int WAVE_SEG = 128;  // sites per item == threads per block of the register-path kernel

// Persistent 128-thread blocks (4 per SM): each block loops over tickets; an item is a 128-site row
// segment.  The atomicAdd for the NEXT ticket is issued before the current item's work and
// consumed after it, so its L2 round trip is hidden.  A block working on ticket t only ever waits
// for items with smaller tickets, each of which is being worked on by a resident block or is the
// prefetched ticket of a block that is itself working on a still smaller one: the chain ends at
// a running item, so there is no deadlock.
template <int C, int EQ, int PSI, bool MASK, bool EXACT>
__global__ void __launch_bounds__(WAVE_SEG, (MASK ? DLBC_MASK_BLOCKS_WAVE : 4))
wave_kernel(Geom g, Fields fl, StepParams P, Wave w) {
  __shared__ int s_ticket[2];
  const bool lead = threadIdx.x == 0;
  const int total = w.nbands * w.slots * (w.rp + w.rf);
  if (lead) s_ticket[0] = atomicAdd(w.sync, 1);
  __syncthreads();
  int t = s_ticket[0];
  for (int it = 1; t < total; it ^= 1) {
    int nxt = 0;
    if (lead) nxt = atomicAdd(w.sync, 1);  // in flight during this item's work
    const WaveItem wi = wave_decode(g, w, t);
    const int x = wi.x0 + threadIdx.x;
    if (wi.kind == 1) {
      if (x < g.nx) psi_site<C, PSI, MASK, POL_KEEP, true>(g, fl, x, wi.y, wi.zg);
      __syncthreads();  // all psi stores of the block precede (CTA scope) the lead thread's release below
      if (lead) red_release_add(w.sync + wi.cnt + wi.zg * (w.tyb + 2) + wi.row, 1);
    } else if (wi.kind == 2) {
#if DLBC_WAVE_EARLY_GATHER
      // the gather of the populations depends on nothing written in this launch: issue it BEFORE the wait for psi,
      // so that the poll's round trips, the fence and the block barrier overlap the DRAM latency of the 19 C loads
      Nbr n = {};
      uint32_t bits = 0;
      double f[C][Q];
      bool go = false;
      if (x < g.nx) go = fused_gather<C, MASK, POL_STREAM>(g, fl, x, wi.y, wi.zg, n, bits, f);
#endif
      if (threadIdx.x < 32) {
        if (threadIdx.x < 9) {
          const int *c = wave_wait_counter(g, w, wi, threadIdx.x);
          if (c) wave_poll_until(c, w.xb * w.nsub, 32);
        }
        __syncwarp();    // the polling lanes leave their loops one by one: ONE converged fence for the warp instead
        wave_acquire();  // of one per straggler (1.56 fences per item measured); the warp barrier orders every poll before it
      }
      __syncthreads();
#if DLBC_WAVE_EARLY_GATHER
      if (go) fused_finish<C, EQ, true, MASK, EXACT, true, POL_STREAM>(g, fl, P, wi.zg, n, bits, f);
#else
      if (x < g.nx) fused_site<C, EQ, true, MASK, EXACT, true, POL_STREAM>(g, fl, P, x, wi.y, wi.zg);
#endif
    }
    if (lead) s_ticket[it] = nxt;
    __syncthreads();
    t = s_ticket[it];
  }
}

}  // namespace dlbc

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
// with F(z) scheduled LA slots behind P(z), i.e. in the slot of P(z + LA) (band-plane = TYB * nx sites,
// ~5-10 MB).
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
  int rowsync;         // F items wait for the 9 (plane, row) counters they read instead of 3 whole planes
  int *sync;           // [0] ticket, [1 + ((band * (nz + 2) + plane) * (tyb + 2) + row)] completed P blocks of that row
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

#ifndef DLBC_WAVE_SEG
#define DLBC_WAVE_SEG 128
#endif
constexpr int WAVE_SEG = DLBC_WAVE_SEG;  // sites per item == threads per block (64 / 256 measured slower or equal)
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

// ---- component-split site functions -----------------------------------------------------------
// One thread per (site, component) instead of one thread per site.  Why: a thread that owns all
// C components of a site holds 19 C populations (76 registers at C = 2), which pins the launch at
// 128 registers = 16 warps/SM, and the GPU then needs ~75k sites in flight to cover memory
// latency -- 46 MB of populations, which together with the P -> F lag overflows what L2 retains
// (profiles/r01_wave_experiments.md: every change that made the P items faster RAISED the DRAM
// traffic).  Split by component, a thread holds 19 populations, the launch fits 24+ warps/SM with
// fewer sites in flight, and the only cross-component coupling of the timestep -- BDist2's common
// velocity (collision.d:156-196) -- goes through 4 doubles per thread of shared memory.  A warp is
// uniform in its component (threadIdx.y), so every load/store is still a full 256-byte row piece.
// Per-site arithmetic is the same sequence of operations as fused_site/psi_site: identical bits.

// sc_force (lb_site.cuh) for the single component f1 (run-time, warp-uniform)
template <int C, bool EXACT, class PsiAt>
__device__ __forceinline__ void sc_force_one(const StepParams &P, int f1, const double (&psi1)[C], double (&F)[3],
                                             PsiAt psi_at) {
  double p1 = psi1[0];
#pragma unroll
  for (int k = 1; k < C; k++) if (k == f1) p1 = psi1[k];
#pragma unroll
  for (int f2 = 0; f2 < C; f2++) {
    const double cc = P.gcc[f1 * MAXC + f2];
    if (cc == 0.0) continue;
    if (EXACT) {
#pragma unroll
      for (int q = 1; q < Q; q++) {
        const double psi2 = psi_at(f2, q);
        const double pref = weight(q) * p1 * psi2 * cc;
        if (cx(q) == 1) F[0] -= pref; else if (cx(q) == -1) F[0] += pref;
        if (cy(q) == 1) F[1] -= pref; else if (cy(q) == -1) F[1] += pref;
        if (cz(q) == 1) F[2] -= pref; else if (cz(q) == -1) F[2] += pref;
      }
    } else {
      const double ax = psi_at(f2, 1) - psi_at(f2, 2);
      const double ay = psi_at(f2, 3) - psi_at(f2, 4);
      const double az = psi_at(f2, 5) - psi_at(f2, 6);
      const double p7 = psi_at(f2, 7), p8 = psi_at(f2, 8), p9 = psi_at(f2, 9), p10 = psi_at(f2, 10);
      const double p11 = psi_at(f2, 11), p12 = psi_at(f2, 12), p13 = psi_at(f2, 13), p14 = psi_at(f2, 14);
      const double p15 = psi_at(f2, 15), p16 = psi_at(f2, 16), p17 = psi_at(f2, 17), p18 = psi_at(f2, 18);
      const double dx = ((p7 - p8) + (p11 - p17)) + ((p12 - p18) + (p13 - p14));
      const double dy = ((p7 - p14) + (p8 - p13)) + ((p9 - p10) + (p15 - p16));
      const double dz = ((p9 - p15) + (p10 - p16)) + ((p11 - p12) + (p17 - p18));
      const double w1 = weight(1), w2 = weight(7);
      const double sx = w1 * ax + w2 * dx, sy = w1 * ay + w2 * dy, sz = w1 * az + w2 * dz;
      const double t = p1 * cc;
      F[0] -= t * sx;
      F[1] -= t * sy;
      F[2] -= t * sz;
    }
  }
}

// wall_force (step_two_pass.cuh) for the single component f1
template <int C>
__device__ __forceinline__ void wall_force_one(const Geom &g, const Fields &fl, const StepParams &P, int zg,
                                               const Nbr &n, int f1, double p1, double (&F)[3]) {
  const double wc = P.gwc[f1];
  if (wc == 0.0) return;
  const uint8_t *mp = mask_plane(fl.mask, g, zg);
  const double *__restrict__ rw = psi_plane(fl.rho_wall, g, zg, f1);
#pragma unroll
  for (int q = 0; q < Q; q++) {
    const int nb = n.at(cx(q), cy(q)) + cz(q) * g.plane;
    if (mp[nb]) {
      const double pref = weight(q) * p1 * rw[nb] * wc;
      if (cx(q) == 1) F[0] -= pref; else if (cx(q) == -1) F[0] += pref;
      if (cy(q) == 1) F[1] -= pref; else if (cy(q) == -1) F[1] += pref;
      if (cz(q) == 1) F[2] -= pref; else if (cz(q) == -1) F[2] += pref;
    }
  }
}

// psi_site for component c only
template <int C, int PSI, bool MASK, int POL>
__device__ __forceinline__ void psi_site_split(const Geom &g, const Fields &fl, int x, int y, int zg, int c) {
  const Nbr n = make_nbr(g, x, y);
  const uint8_t *ms = mask_plane(fl.mask, g, zg) + n.sxy;
  if (MASK && ms[0]) return;
  double f[Q];
  pull19<MASK, POL>(f_plane(fl.f_old, g, zg, c) + n.sxy, ms, g, n, f);
  const double v = psi_of<PSI>(density19(f));
  psi_plane(fl.psi, g, zg, c)[n.sxy] = v;
  if (g.periodic_z_local) {
    if (zg == g.nz) psi_plane(fl.psi, g, 0, c)[n.sxy] = v;
    if (zg == 1) psi_plane(fl.psi, g, g.nz + 1, c)[n.sxy] = v;
  }
}

// fused_site for component c only.  EVERY thread of the block must call it (block barrier for the
// BDist2 exchange); in_range = false marks threads beyond the end of the row.
// xch: shared memory [4][C][WAVE_SEG] (rho, m_x, m_y, m_z of every component of the block's sites).
template <int C, int EQ, bool MASK, bool EXACT, int POL>
__device__ __forceinline__ void fused_site_split(const Geom &g, const Fields &fl, const StepParams &P, int x, int y,
                                                 int zg, int c, bool in_range, double (*xch)[C][WAVE_SEG]) {
  const Nbr n = make_nbr(g, in_range ? x : 0, y);
  const uint8_t *ms = mask_plane(fl.mask, g, zg) + n.sxy;
  const bool act = in_range && !(MASK && ms[0]);
  double f[Q], rho = 0.0, m[3] = {0.0, 0.0, 0.0}, F[3] = {0.0, 0.0, 0.0};
  if (act) {
    pull19<MASK, POL>(f_plane(fl.f_old, g, zg, c) + n.sxy, ms, g, n, f);
    rho = density19(f);
    momentum19(f, m);
    double psi1[C];
#pragma unroll
    for (int k = 0; k < C; k++) psi1[k] = psi_plane(fl.psi, g, zg, k)[n.sxy];
    const double *pp = psi_plane(fl.psi, g, zg, 0) + n.sxy;  // written by P items of this launch: plain loads
    const int64_t cstride = g.sites_g;
    sc_force_one<C, EXACT>(P, c, psi1, F, [&](int f2, int q) -> double {
      const int d = n.ox[1 + cx(q)] + n.oy[1 + cy(q)] + cz(q) * g.plane;
      if (MASK && ms[d]) return psi1[f2];
      return pp[f2 * cstride + d];
    });
    if (MASK && P.any_gwc) {
      double p1 = psi1[0];
#pragma unroll
      for (int k = 1; k < C; k++) if (k == c) p1 = psi1[k];
      wall_force_one<C>(g, fl, P, zg, n, c, p1, F);
    }
  }
  double u[3] = {0.0, 0.0, 0.0};
  if (EQ == 2) {
    const int tx = threadIdx.x;
    xch[0][c][tx] = rho;
    xch[1][c][tx] = m[0];
    xch[2][c][tx] = m[1];
    xch[3][c][tx] = m[2];
    __syncthreads();
    if (act) {
      double ra[C], ma[C][3];
#pragma unroll
      for (int k = 0; k < C; k++) {
        ra[k] = xch[0][k][tx];
        ma[k][0] = xch[1][k][tx];
        ma[k][1] = xch[2][k][tx];
        ma[k][2] = xch[3][k][tx];
      }
      common_velocity<C, EXACT>(P, ma, ra, u);
    }
  }
  if (act) {
    collide_site<EQ, EXACT>(f, rho, m, F, u, P, c);
    store19<POL>(g, fl.f_new, zg, c, n.sxy, f);
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

// SPLIT: block = WAVE_SEG sites x C components (threadIdx.y = component), see above.
template <int C, bool MASK, bool SPLIT>
constexpr int wave_min_blocks() { return !SPLIT ? (MASK ? DLBC_MASK_BLOCKS : 4) * 128 / WAVE_SEG : (C <= 2 ? 3 : (C == 3 ? 2 : 1)); }

template <int C, int EQ, int PSI, bool MASK, bool EXACT, bool SPLIT>
__global__ void __launch_bounds__(WAVE_SEG * (SPLIT ? C : 1), (wave_min_blocks<C, MASK, SPLIT>()))
wave_kernel(Geom g, Fields fl, StepParams P, Wave w) {
  __shared__ int s_ticket[2];
  __shared__ double s_xch[SPLIT ? 4 : 1][SPLIT ? C : 1][SPLIT ? WAVE_SEG : 1];
  const int comp = SPLIT ? (int)threadIdx.y : 0;
  const bool lead = threadIdx.x == 0 && comp == 0;
  const int per_slot = w.rp + w.rf;
  const int per_band = w.slots * per_slot;
  const int total = w.nbands * per_band;
  if (lead) s_ticket[0] = atomicAdd(w.sync, 1);
  __syncthreads();
  int t = s_ticket[0];
  for (int it = 1; t < total; it ^= 1) {
    int nxt = 0;
    if (lead) nxt = atomicAdd(w.sync, 1);  // in flight during this item's work
    if (WAVE_PREFETCH && !SPLIT && w.pf > 0 && t + w.pf < total) wave_prefetch<C>(g, fl, w, t + w.pf, per_band, per_slot);

    const int band = t / per_band;
    const int r = t - band * per_band;
    const int k = r / per_slot;
    const int j = r - k * per_slot;
    const int yb = band * w.tyb;
    const int tb = min(w.tyb, g.ny - yb);   // rows of this (possibly ragged) band
    const int rows = w.tyb + 2;
    int *cnt = w.sync + 1 + (int64_t)band * (g.nz + 2) * rows;  // [plane][row]

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
          if constexpr (SPLIT) psi_site_split<C, PSI, MASK, POL_KEEP>(g, fl, x, y, p, comp);
          else if (w.batch) psi_site<C, PSI, MASK, POL_KEEP, true>(g, fl, x, y, p);
          else psi_site<C, PSI, MASK, POL_KEEP, false>(g, fl, x, y, p);
        }
        __syncthreads();  // all psi stores of the block precede (CTA scope) the lead thread's release below
        if (lead) red_release_add(cnt + p * rows + row, 1);
      }
    } else if (k > 0) {
      // ---- F item: fused update of one row segment of plane zf
      const int zf = w.pl + (k - 1) - w.la;  // `la` slots behind the P item of the same plane (pl < z_begin on slabs)
      const int jf = j - w.rp;
      const int row = jf / w.xb, xblk = jf - row * w.xb;
      if (zf >= w.z_begin && zf < w.z_end && row < tb) {
        // psi of planes zf-1..zf+1, rows y-1..y+1 (P rows row..row+2 of this band), all x segments of those
        // rows (xb items each): 9 counters, or -- rowsync off -- every row of the 3 planes
        const int nwait = w.rowsync ? 9 : 3 * (tb + 2);
        if (comp == 0) {
          for (int i = threadIdx.x; i < nwait; i += WAVE_SEG) {
            const int i3 = w.rowsync ? i / 3 : i / (tb + 2);
            const int rr = w.rowsync ? row + i % 3 : i % (tb + 2);
            int d = zf + i3 - 1;
            bool need = true;
            if (d < 1) { need = g.periodic_z_local; d = g.nz; }
            else if (d > g.nz) { need = g.periodic_z_local; d = 1; }
            if (need) {
              const int *c = cnt + d * rows + rr;
              while (ld_poll(c) < w.xb) __nanosleep(32);
            }
          }
        }
        __syncthreads();
        const int x = xblk * WAVE_SEG + threadIdx.x;
        if constexpr (SPLIT) fused_site_split<C, EQ, MASK, EXACT, POL_STREAM>(g, fl, P, x, yb + row, zf, comp, x < g.nx, s_xch);
        else if (x < g.nx) fused_site<C, EQ, true, MASK, EXACT, true, POL_STREAM>(g, fl, P, x, yb + row, zf);
      }
    }

    if (lead) s_ticket[it] = nxt;
    __syncthreads();
    t = s_ticket[it];
  }
}


// ---- barrier-free form ---------------------------------------------------------------------------
// Same items, same ticket order, same per-site functions as wave_kernel, but the four warps of a block
// never meet at a block barrier: in wave_kernel 20 % of the warp time was barrier stall (every warp
// waiting for the block's slowest at the psi-plane poll, at the P completion and at the ticket hand-over;
// profiles/README.md).  Here
//   * the ticket of the block's next item is fetched by warp 0 and handed over through a small ring in
//     shared memory; a warp that finishes early only waits for warp 0 to have published, and warp 0
//     only waits if it is RING items ahead of the slowest warp (s_prog), so slots are never overwritten
//     before every warp has read them;
//   * P completion is counted per warp (target: rp * warps per block), each warp releasing after its own
//     __syncwarp; F items poll the three psi planes per warp.
// Deadlock freedom is unchanged: a warp only ever waits for items with smaller tickets or for its own
// block's warp 0, which waits for nobody unless it is RING items ahead.
constexpr int WAVE_RING = 4;

template <int C, int EQ, int PSI, bool MASK, bool EXACT>
__global__ void __launch_bounds__(WAVE_SEG, (wave_min_blocks<C, MASK, false>()))
wave_free_kernel(Geom g, Fields fl, StepParams P, Wave w) {
  constexpr int NW = WAVE_SEG / 32;
  __shared__ int s_ring[WAVE_RING];
  __shared__ int s_prog[NW];  // index of the last item whose ticket warp i has read
  volatile int *ring = s_ring;
  volatile int *prog = s_prog;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int per_slot = w.rp + w.rf;
  const int per_band = w.slots * per_slot;
  const int total = w.nbands * per_band;
  if (threadIdx.x == 0) ring[0] = atomicAdd(w.sync, 1);
  else if (threadIdx.x < WAVE_RING) ring[threadIdx.x] = -1;
  if (threadIdx.x < NW) prog[threadIdx.x] = 0;
  __syncthreads();  // the only block barrier
  int t = ring[0];
  for (int n = 0; t < total; n++) {
    int nxt = 0;
    if (threadIdx.x == 0) nxt = atomicAdd(w.sync, 1);  // ticket of item n + 1, in flight during this item's work

    const int band = t / per_band;
    const int r = t - band * per_band;
    const int k = r / per_slot;
    const int j = r - k * per_slot;
    const int yb = band * w.tyb;
    const int tb = min(w.tyb, g.ny - yb);
    const int rows = w.tyb + 2;
    int *cnt = w.sync + 1 + (int64_t)band * (g.nz + 2) * rows;  // [plane][row]

    if (j < w.rp) {
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
        __syncwarp();  // the warp's psi stores precede lane 0's release
        if (lane == 0) red_release_add(cnt + p * rows + row, 1);
      }
    } else if (k > 0) {
      const int zf = w.pl + (k - 1) - w.la;  // `la` slots behind the P item of the same plane (pl < z_begin on slabs)
      const int jf = j - w.rp;
      const int row = jf / w.xb, xblk = jf - row * w.xb;
      if (zf >= w.z_begin && zf < w.z_end && row < tb) {
        if (lane < 9) {
          int d = zf + lane / 3 - 1;
          bool need = true;
          if (d < 1) { need = g.periodic_z_local; d = g.nz; }
          else if (d > g.nz) { need = g.periodic_z_local; d = 1; }
          if (need) {
            const int *c = cnt + d * rows + row + lane % 3;
            while (ld_poll(c) < w.xb * NW) __nanosleep(32);
          }
        }
        __syncwarp();
        const int x = xblk * WAVE_SEG + threadIdx.x;
        if (x < g.nx) fused_site<C, EQ, true, MASK, EXACT, true, POL_STREAM>(g, fl, P, x, yb + row, zf);
      }
    }

    // hand-over of the next ticket (item n + 1, slot (n + 1) % RING)
    const int slot = (n + 1) % WAVE_RING;
    if (threadIdx.x == 0) {
      // the slot still holds the ticket of item n + 1 - RING: every warp must have read it
      if (n + 1 >= WAVE_RING) {
#pragma unroll
        for (int i = 1; i < NW; i++) while (prog[i] < n + 1 - WAVE_RING) __nanosleep(40);
      }
      ring[slot] = nxt;
    }
    int tn;
    while ((tn = ring[slot]) <= t) __nanosleep(40);  // tickets only grow: a larger value is the new one (no busy spin:
                                                     // a spinning warp takes issue slots from the working ones)
    __syncwarp();
    if (lane == 0) prog[wid] = n + 1;
    t = tn;
  }
}

}  // namespace dlbc

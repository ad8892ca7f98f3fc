// lb_site.cuh -- per-site D3Q19 arithmetic shared by all step kernels (device code).
//
// Every function states the reference code it reproduces (paths relative to
// /root/reference/src/dlbc/).  Operation ORDER follows the reference so that, compiled with
// -fmad=false, results are bit-identical to the reference's fp64 arithmetic wherever EXACT
// arithmetic is selected (the only exception is exp(), <= 1 ulp).  Terms multiplied by a zero
// velocity component are skipped and +-1 multiplications become add/subtract: both are exact
// identities for finite values, so skipping them does not change a single bit.
#pragma once
#include <cuda_runtime.h>

namespace dlbc {

constexpr int Q = 19;
constexpr int MAXC = 4;

// lb/connectivity.d:401-418 (generateD3Q19)
__host__ __device__ constexpr int cx(int q) {
  constexpr int t[Q] = {0, 1, -1, 0, 0, 0, 0, 1, -1, 0, 0, 1, 1, 1, -1, 0, 0, -1, -1};
  return t[q];
}
__host__ __device__ constexpr int cy(int q) {
  constexpr int t[Q] = {0, 0, 0, 1, -1, 0, 0, 1, 1, 1, -1, 0, 0, -1, -1, 1, -1, 0, 0};
  return t[q];
}
__host__ __device__ constexpr int cz(int q) {
  constexpr int t[Q] = {0, 0, 0, 0, 0, 1, -1, 0, 0, 1, 1, 1, -1, 0, 0, -1, -1, 1, -1};
  return t[q];
}
// lb/connectivity.d:128-142 (generateBounce): index of -c_q
__host__ __device__ constexpr int bounce(int q) {
  constexpr int t[Q] = {0, 2, 1, 4, 3, 6, 5, 14, 13, 16, 15, 18, 17, 8, 7, 10, 9, 12, 11};
  return t[q];
}
// Storage slot of direction q inside a component's block of 19 planes (DEVICE layout only; the C ABI speaks
// the reference's q order, field_io.cu converts).  Directions that stream from the same neighbouring ROW --
// equal (c_y, c_z) -- sit in consecutive slots, so that one TMA box of 3 planes fetches all three
// (step_tma.cuh: 9 copies per item instead of 19 per component):
//   slots 0-2: q 0 1 2 (0,0) | 3-5: q 3 7 8 (+y) | 6-8: q 4 13 14 (-y) | 9-11: q 5 11 17 (+z) | 12-14: q 6 12 18 (-z)
//   15: q 9 (+y+z) | 16: q 10 (-y+z) | 17: q 15 (+y-z) | 18: q 16 (-y-z)
__host__ __device__ constexpr int qslot(int q) {
  constexpr int t[Q] = {0, 1, 2, 3, 6, 9, 12, 4, 5, 15, 16, 10, 13, 7, 8, 17, 18, 11, 14};
  return t[q];
}
// the same for a run-time q (no table in local memory): 5-bit codes
__host__ __device__ constexpr unsigned long long qslot_pack(int lo, int hi) {
  unsigned long long m = 0;
  for (int q = lo; q < hi; q++) m |= (unsigned long long)qslot(q) << (5 * (q - lo));
  return m;
}
__host__ __device__ inline int qslot_rt(int q) {
  constexpr unsigned long long a = qslot_pack(0, 12), b = qslot_pack(12, Q);
  return (int)(((q < 12 ? a >> (5 * q) : b >> (5 * (q - 12)))) & 31u);
}

// lb/connectivity.d:253-257
__host__ __device__ constexpr double weight(int q) {
  return q == 0 ? 1.0 / 3.0 : (q < 7 ? 1.0 / 18.0 : 1.0 / 36.0);
}

struct StepParams {
  int ncomp;
  double tau[MAXC];
  double omega[MAXC];        // 1.0 / tau                          (collision.d:99)
  double one_minus_omega[MAXC];
  double inv_tau[MAXC];      // == omega; used when tau_pow2 (exact) or in FAST mode
  int tau_pow2[MAXC];        // tau is a power of two: x / tau == x * (1/tau) exactly
  double gcc[MAXC * MAXC];
  double gwc[MAXC];
  double gacc[3];
  int any_gwc;
};

// lb/density.d:37-39: Phobos std.algorithm.sum on a random-access fp range is PAIRWISE:
// s(r) = len <= 2 ? direct : s(r[0..len/2]) + s(r[len/2..len]).  For 19 elements:
//   [0,9) = ((f0+f1)+(f2+f3)) + ((f4+f5)+(f6+(f7+f8)))
//   [9,19) = ((f9+f10)+(f11+(f12+f13))) + ((f14+f15)+(f16+(f17+f18)))
// Pinned by the reference's golden population -> density files (tests/test_oracle_golden.py).
__device__ __forceinline__ double density19(const double (&f)[Q]) {
  const double a = ((f[0] + f[1]) + (f[2] + f[3])) + ((f[4] + f[5]) + (f[6] + (f[7] + f[8])));
  const double b = ((f[9] + f[10]) + (f[11] + (f[12] + f[13]))) + ((f[14] + f[15]) + (f[16] + (f[17] + f[18])));
  return a + b;
}

// lb/momentum.d:39-49 / velocity.d:37-48 numerator: m[d] = sum_q f[q]*c_q[d], sequential over q
// from 0.0 (zero terms skipped, +-1 as add/sub).
__device__ __forceinline__ void momentum19(const double (&f)[Q], double (&m)[3]) {
  m[0] = 0.0; m[1] = 0.0; m[2] = 0.0;
#pragma unroll
  for (int q = 0; q < Q; q++) {
    if (cx(q) == 1) m[0] += f[q]; else if (cx(q) == -1) m[0] -= f[q];
    if (cy(q) == 1) m[1] += f[q]; else if (cy(q) == -1) m[1] -= f[q];
    if (cz(q) == 1) m[2] += f[q]; else if (cz(q) == -1) m[2] -= f[q];
  }
}

// lb/force.d:320-332
template <int PSI>
__device__ __forceinline__ double psi_of(double rho) {
  if (PSI == 0) return rho;
  return 1.0 - exp(-rho);
}

template <bool EXACT>
__device__ __forceinline__ double div_by(double x, double y, double inv_y) {
  return EXACT ? x / y : x * inv_y;
}

// lb/eqdist.d:69-100 (SecondOrder / ThirdOrder).  `m` is the raw momentum sum (velocity.d:40-45),
// divided by rho0 here (velocity.d:46).
template <int FORM, bool EXACT>
__device__ __forceinline__ void eqdist_poly(const double (&m)[3], const double (&dv)[3], double rho0,
                                            double (&dist)[Q]) {
  const double css = 1.0 / 3.0;
  const double css2 = 2.0 * css;
  const double css2p2 = css2 * css;
  const double css6p3 = 3.0 * css2p2 * css;
  const double inv_rho = EXACT ? 0.0 : 1.0 / rho0;
  double v[3];
#pragma unroll
  for (int d = 0; d < 3; d++) v[d] = dv[d] + div_by<EXACT>(m[d], rho0, inv_rho);
  const double vdotv = ((0.0 + v[0] * v[0]) + v[1] * v[1]) + v[2] * v[2];
  const double t_vv = div_by<EXACT>(vdotv, css2, 1.0 / css2);
#pragma unroll
  for (int q = 0; q < Q; q++) {
    // dotProduct (range.d:57-67): sum0 = 0; sum0 += v[d] * c[d]
    double vc = 0.0;
    if (cx(q) == 1) vc += v[0]; else if (cx(q) == -1) vc -= v[0];
    if (cy(q) == 1) vc += v[1]; else if (cy(q) == -1) vc -= v[1];
    if (cz(q) == 1) vc += v[2]; else if (cz(q) == -1) vc -= v[2];
    double poly = 1.0 + div_by<EXACT>(vc, css, 1.0 / css) + div_by<EXACT>(vc * vc, css2p2, 1.0 / css2p2) - t_vv;
    if (FORM == 1) {
      poly = poly + div_by<EXACT>(vc * vc * vc, css6p3, 1.0 / css6p3) - div_by<EXACT>(vdotv * vc, css2p2, 1.0 / css2p2);
    }
    dist[q] = rho0 * weight(q) * poly;
  }
}

// lb/eqdist.d:101-184 (BDist2, D3Q19): LB3D-compatible third-order form in dv only.
__device__ __forceinline__ void eqdist_bdist2(const double (&dv)[3], double rho0, double (&dist)[Q]) {
  const double aN1 = 1.0 / 36.0;
  const double aN0 = 1.0 / 3.0;
  const double haN1 = 0.5 * aN1;
  const double aN1_6 = aN1 / 6.0;
  const double T = 1.0 / 3.0;
  const double holdit = 1.0 / T;
  const double tux = dv[0], tuy = dv[1], tuz = dv[2];
  const double tuxt = tux * holdit, tuyt = tuy * holdit, tuzt = tuz * holdit;
  const double uke = 0.5 * (tux * tuxt + tuy * tuyt + tuz * tuzt);
  const double cst = 1.0 - uke;
  const double const1 = aN1 * cst;
  double cdotu, even, odd;
#define DLBC_PAIR(expr, qp, qm, twice)                       \
  cdotu = (expr);                                            \
  even = const1 + haN1 * cdotu * cdotu;                      \
  odd = const1 * cdotu + aN1_6 * cdotu * cdotu * cdotu;      \
  dist[qp] = twice ? 2.0 * (even + odd) : (even + odd);      \
  dist[qm] = twice ? 2.0 * (even - odd) : (even - odd);
  DLBC_PAIR(tuxt, 1, 2, true)
  DLBC_PAIR(tuyt, 3, 4, true)
  DLBC_PAIR(tuzt, 5, 6, true)
  DLBC_PAIR(tuxt + tuyt, 7, 14, false)
  DLBC_PAIR(tuxt - tuyt, 13, 8, false)
  DLBC_PAIR(tuxt + tuzt, 11, 18, false)
  DLBC_PAIR(tuxt - tuzt, 12, 17, false)
  DLBC_PAIR(tuyt + tuzt, 9, 16, false)
  DLBC_PAIR(tuyt - tuzt, 15, 10, false)
#undef DLBC_PAIR
  dist[0] = aN0 * cst;
#pragma unroll
  for (int q = 0; q < Q; q++) dist[q] *= rho0;
}

// FAST-arithmetic form of eqdist_bdist2 + BGK relaxation in one go (collision.d:100-116 with
// eqdist.d:101-184): omega * rho is folded into the polynomial constants, the odd part is
// factored as cu * (c1 + a6 * cu^2), and the relaxation is one FMA per population.  Same
// polynomial, different rounding (a few ulp; the parity tolerance is 1e-12).
__device__ __forceinline__ void bdist2_relax_fast(double (&f)[Q], const double (&dv)[3], double rho, double omega,
                                                  double om1) {
  const double aN1 = 1.0 / 36.0, aN0 = 1.0 / 3.0, haN1 = 0.5 * aN1, aN1_6 = aN1 / 6.0;
  const double holdit = 1.0 / (1.0 / 3.0);
  const double s = omega * rho;
  const double tx = dv[0] * holdit, ty = dv[1] * holdit, tz = dv[2] * holdit;
  const double cst = 1.0 - 0.5 * (dv[0] * tx + dv[1] * ty + dv[2] * tz);
  const double c1 = (s * aN1) * cst, h = s * haN1, a6 = s * aN1_6;
  const double c1x2 = 2.0 * c1, hx2 = 2.0 * h, a6x2 = 2.0 * a6;
#define DLBC_PAIR_FAST(expr, qp, qm, K1, H, A6)        \
  {                                                    \
    const double cu = (expr);                          \
    const double cu2 = cu * cu;                        \
    const double even = K1 + H * cu2;                  \
    const double odd = cu * (K1 + A6 * cu2);           \
    f[qp] = f[qp] * om1 + (even + odd);                \
    f[qm] = f[qm] * om1 + (even - odd);                \
  }
  DLBC_PAIR_FAST(tx, 1, 2, c1x2, hx2, a6x2)
  DLBC_PAIR_FAST(ty, 3, 4, c1x2, hx2, a6x2)
  DLBC_PAIR_FAST(tz, 5, 6, c1x2, hx2, a6x2)
  DLBC_PAIR_FAST(tx + ty, 7, 14, c1, h, a6)
  DLBC_PAIR_FAST(tx - ty, 13, 8, c1, h, a6)
  DLBC_PAIR_FAST(tx + tz, 11, 18, c1, h, a6)
  DLBC_PAIR_FAST(tx - tz, 12, 17, c1, h, a6)
  DLBC_PAIR_FAST(ty + tz, 9, 16, c1, h, a6)
  DLBC_PAIR_FAST(ty - tz, 15, 10, c1, h, a6)
#undef DLBC_PAIR_FAST
  f[0] = f[0] * om1 + (s * aN0) * cst;
}

// lb/collision.d:100-116: dv, equilibrium, BGK relaxation of one component at one site.
// F = Shan-Chen force on this component (force.d), m = raw momentum of f, u = BDist2 common
// velocity (collision.d:156-196; ignored otherwise).
// keep (only looked at when KEEP): this lane's site does not collide (a solid site whose populations are carried
// along so that whole sectors are written, step_two_pass.cuh): relax with omega = 0 towards the finite
// equilibrium of rho = 1, dv = 0 -- f * 1 + 0 * eq -- which returns every f[q] bit for bit (a -0.0 would come back
// as +0.0; populations are sums of non-negative terms) without a second code path for the warp to diverge into.
template <int EQ, bool EXACT, bool KEEP = false>
__device__ __forceinline__ void collide_site(double (&f)[Q], double rho, const double (&m)[3],
                                             const double (&F)[3], const double (&u)[3],
                                             const StepParams &P, int c, bool keep = false) {
  double dv[3];
  if (KEEP && keep) rho = 1.0;
  const double inv_rho = EXACT ? 0.0 : 1.0 / rho;
#pragma unroll
  for (int d = 0; d < 3; d++) {
    dv[d] = P.tau[c] * (P.gacc[d] + div_by<EXACT>(F[d], rho, inv_rho));
    if (EQ == 2) dv[d] += u[d];
    if (KEEP && keep) dv[d] = 0.0;
  }
  const double omega = (KEEP && keep) ? 0.0 : P.omega[c], om1 = (KEEP && keep) ? 1.0 : P.one_minus_omega[c];
  if (EQ == 2 && !EXACT) {
    bdist2_relax_fast(f, dv, rho, omega, om1);
    return;
  }
  double eq[Q];
  if (EQ == 2) eqdist_bdist2(dv, rho, eq);
  else eqdist_poly<EQ, EXACT>(m, dv, rho, eq);
#pragma unroll
  for (int q = 0; q < Q; q++) f[q] = f[q] * om1 + omega * eq[q];
}

// lb/force.d:238-309, fluid-fluid part: F_f1 -= sum_f2 gcc[f1][f2] * psi1[f1] * sum_q w_q psi_f2(p + c_q) c_q.
// psi_at(f2, q) returns psi_f2 at the neighbour in direction q, with the reference's substitution of
// the site's own psi when that neighbour bounces back (force.d:273) already applied.
//   EXACT: the reference's accumulation order per component f1 -- f2 outer, q = 1..18 inner,
//          prefactor = ((w_q * psi1) * psi2) * cc, one subtraction per non-zero velocity component.
//   FAST:  the same sum regrouped by weight class (6 axis links w = 1/18, 12 diagonal links
//          w = 1/36): 24 additions + 6 FMAs per f2 instead of 54 multiplies + 30 subtractions per (f1, f2).
template <int C, bool EXACT, class PsiAt>
__device__ __forceinline__ void sc_force(const StepParams &P, const double (&psi1)[C], double (&F)[C][3], PsiAt psi_at) {
#pragma unroll
  for (int f2 = 0; f2 < C; f2++) {
    bool any = false;
#pragma unroll
    for (int f1 = 0; f1 < C; f1++) any |= (P.gcc[f1 * MAXC + f2] != 0.0);
    if (!any) continue;
    if (EXACT) {
#pragma unroll
      for (int q = 1; q < Q; q++) {
        const double psi2 = psi_at(f2, q);
#pragma unroll
        for (int f1 = 0; f1 < C; f1++) {
          const double cc = P.gcc[f1 * MAXC + f2];
          if (cc == 0.0) continue;
          const double pref = weight(q) * psi1[f1] * psi2 * cc;
          if (cx(q) == 1) F[f1][0] -= pref; else if (cx(q) == -1) F[f1][0] += pref;
          if (cy(q) == 1) F[f1][1] -= pref; else if (cy(q) == -1) F[f1][1] += pref;
          if (cz(q) == 1) F[f1][2] -= pref; else if (cz(q) == -1) F[f1][2] += pref;
        }
      }
    } else {
      // axis links: 1,2 (x) 3,4 (y) 5,6 (z); the first diagonal term of each component seeds its sum
      const double ax = psi_at(f2, 1) - psi_at(f2, 2);
      const double ay = psi_at(f2, 3) - psi_at(f2, 4);
      const double az = psi_at(f2, 5) - psi_at(f2, 6);
      const double p7 = psi_at(f2, 7), p8 = psi_at(f2, 8), p9 = psi_at(f2, 9), p10 = psi_at(f2, 10);
      const double p11 = psi_at(f2, 11), p12 = psi_at(f2, 12), p13 = psi_at(f2, 13), p14 = psi_at(f2, 14);
      const double p15 = psi_at(f2, 15), p16 = psi_at(f2, 16), p17 = psi_at(f2, 17), p18 = psi_at(f2, 18);
      // c_x = +1: 7 11 12 13, -1: 8 14 17 18;  c_y = +1: 7 8 9 15, -1: 10 13 14 16;  c_z = +1: 9 10 11 17, -1: 12 15 16 18
      const double dx = ((p7 - p8) + (p11 - p17)) + ((p12 - p18) + (p13 - p14));
      const double dy = ((p7 - p14) + (p8 - p13)) + ((p9 - p10) + (p15 - p16));
      const double dz = ((p9 - p15) + (p10 - p16)) + ((p11 - p12) + (p17 - p18));
      const double w1 = weight(1), w2 = weight(7);
      const double sx = w1 * ax + w2 * dx, sy = w1 * ay + w2 * dy, sz = w1 * az + w2 * dz;
#pragma unroll
      for (int f1 = 0; f1 < C; f1++) {
        const double cc = P.gcc[f1 * MAXC + f2];
        if (cc == 0.0) continue;
        const double t = psi1[f1] * cc;
        F[f1][0] -= t * sx;
        F[f1][1] -= t * sy;
        F[f1][2] -= t * sz;
      }
    }
  }
}

}  // namespace dlbc

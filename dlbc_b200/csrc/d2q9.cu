// d2q9.cu -- the reference's 2-D lattice, Lattice!(d2q9), on the device: the same timestep as the D3Q19 path
// (main.d:174-200: advect -> density / psi -> Shan-Chen force -> collide) with the D2Q9 connectivity
// (connectivity.d:356-381: rest, 4 axis links, 4 diagonals (1,1) (1,-1) (-1,1) (-1,-1); weights 4/9, 1/9, 1/36,
// :259-263) and the D2Q9 form of the BDist2 equilibrium (eqdist.d:185-231).
//
// 2-D lattices are small (the reference's 2-D cases are 32 x 4 ... 64 x 64 ... 256 x 256): this path is built for
// parity, not for the roofline -- two plain kernels per step (psi2_kernel: streamed density and psi of every site;
// fused2_kernel: stream, force, collide, store), SoA populations f[c][q][y][x] in A/B buffers, periodic wrap in the
// kernels, one rank.  Arithmetic: the reference's operation order throughout (this unit is compiled with
// -fmad=false); DLBC_ARITH_FAST and DLBC_ARITH_EXACT are the same code here.  Solid sites keep their populations
// (advection.d:83-85); psi and rho are stored for every site (solid sites: rho of their own populations, which
// is what the wall force reads, force.d:285-307).
#include "d2q9.h"

#include <cmath>
#include <cstring>

namespace dlbc {
namespace d2 {

constexpr int Q2 = 9;
constexpr int MAXC2 = DLBC_CUDA_MAX_COMPONENTS;

__host__ __device__ constexpr int c2x(int q) { constexpr int t[Q2] = {0, 1, -1, 0, 0, 1, 1, -1, -1}; return t[q]; }
__host__ __device__ constexpr int c2y(int q) { constexpr int t[Q2] = {0, 0, 0, 1, -1, 1, -1, 1, -1}; return t[q]; }
__host__ __device__ constexpr int bounce2(int q) { constexpr int t[Q2] = {0, 2, 1, 4, 3, 8, 7, 6, 5}; return t[q]; }
__host__ __device__ constexpr double weight2(int q) { return q == 0 ? 4.0 / 9.0 : (q < 5 ? 1.0 / 9.0 : 1.0 / 36.0); }

struct Params2 {
  int nx, ny, n;  // n = nx * ny
  int ncomp, sc;
  double tau[MAXC2], omega[MAXC2], om1[MAXC2];
  double gcc[MAXC2 * MAXC2], gwc[MAXC2], gacc[2];
};

struct Lattice2 {
  Params2 p{};
  int psi_form = 0, eq_form = 0, device = 0;
  double *f[2] = {nullptr, nullptr};  // [c][q][y][x]
  int cur = 0;
  double *psi = nullptr, *rho = nullptr;  // [c][y][x], of the streamed populations of the last step
  uint8_t *mask = nullptr;                // [y][x]
  double *stage = nullptr;                // [nx][ny][9] doubles: dense AoS staging of uploads / downloads / outputs
  double *partial = nullptr;              // [4]
  int *flag = nullptr;                    // bad mask value seen
  cudaStream_t s = nullptr;
  int64_t timestep = 0, nlaunch = 0, bytes = 0;
};

// ---- per-site functions -----------------------------------------------------------------------------------
// density.d:37-39 (Phobos pairwise sum, 9 elements: 4 + 5, the 5 as 2 + 3)
__device__ __forceinline__ double density9(const double (&f)[Q2]) {
  return ((f[0] + f[1]) + (f[2] + f[3])) + ((f[4] + f[5]) + (f[6] + (f[7] + f[8])));
}
// velocity.d:37-48 numerator: sequential over q from 0.0 (the zero terms leave the sum unchanged)
__device__ __forceinline__ void momentum9(const double (&f)[Q2], double (&m)[2]) {
  m[0] = 0.0; m[1] = 0.0;
#pragma unroll
  for (int q = 0; q < Q2; q++) {
    if (c2x(q) == 1) m[0] += f[q]; else if (c2x(q) == -1) m[0] -= f[q];
    if (c2y(q) == 1) m[1] += f[q]; else if (c2y(q) == -1) m[1] -= f[q];
  }
}
template <int PSI>
__device__ __forceinline__ double psi2_of(double rho) {  // force.d:320-332
  return PSI == 0 ? rho : 1.0 - exp(-rho);
}
// periodic neighbour of (x, y) in direction (dx, dy), as a site index
__device__ __forceinline__ int nbr2(const Params2 &p, int x, int y, int dx, int dy) {
  int xx = x + dx, yy = y + dy;
  xx = xx < 0 ? p.nx - 1 : (xx >= p.nx ? 0 : xx);
  yy = yy < 0 ? p.ny - 1 : (yy >= p.ny ? 0 : yy);
  return yy * p.nx + xx;
}
// advection.d:61-89 for one component at one site (solid sites keep their populations)
__device__ __forceinline__ void pull9(const Params2 &p, const double *__restrict__ fc, const uint8_t *__restrict__ mask, int x,
                                      int y, double (&f)[Q2]) {
  const int s = y * p.nx + x;
  const bool solid = mask[s] != 0;
#pragma unroll
  for (int q = 0; q < Q2; q++) {
    const int nb = nbr2(p, x, y, -c2x(q), -c2y(q));
    int src = q * p.n + nb;
    if (q > 0 && mask[nb]) src = bounce2(q) * p.n + s;
    if (solid) src = q * p.n + s;
    f[q] = fc[src];
  }
}

// eqdist.d:69-100 (SecondOrder = 0, ThirdOrder = 1) and :185-231 (BDist2 = 2, D2Q9 constants)
template <int EQ>
__device__ __forceinline__ void eqdist9(const double (&f)[Q2], const double (&dv)[2], double rho0, double (&dist)[Q2]) {
  if (EQ == 2) {
    const double aN1 = 1.0 / 24.0, aN0 = 1.0 / 2.0, haN1 = 0.5 * aN1, aN1_6 = aN1 / 6.0;
    const double holdit = 1.0 / (1.0 / 3.0);
    const double tux = dv[0], tuy = dv[1];
    const double tuxt = tux * holdit, tuyt = tuy * holdit;
    const double uke = 0.5 * (tux * tuxt + tuy * tuyt);
    const double cst = 1.0 - uke;
    const double const1 = aN1 * cst;
#define DLBC_PAIR2(expr, qp, qm, TWO)                                   \
  {                                                                     \
    const double cu = (expr);                                           \
    const double even = const1 + haN1 * cu * cu;                        \
    const double odd = const1 * cu + aN1_6 * cu * cu * cu;              \
    dist[qp] = TWO ? 2.0 * (even + odd) : (even + odd);                 \
    dist[qm] = TWO ? 2.0 * (even - odd) : (even - odd);                 \
  }
    DLBC_PAIR2(tuxt, 1, 2, true)
    DLBC_PAIR2(tuyt, 3, 4, true)
    DLBC_PAIR2(tuxt + tuyt, 5, 8, false)
    DLBC_PAIR2(tuxt - tuyt, 6, 7, false)
#undef DLBC_PAIR2
    dist[0] = aN0 * cst;
#pragma unroll
    for (int q = 0; q < Q2; q++) dist[q] *= rho0;
    return;
  }
  const double css = 1.0 / 3.0, css2 = 2.0 * css, css2p2 = css2 * css, css6p3 = 3.0 * css2p2 * css;
  double m[2], v[2];
  momentum9(f, m);
  v[0] = dv[0] + m[0] / rho0;
  v[1] = dv[1] + m[1] / rho0;
  const double vdotv = (0.0 + v[0] * v[0]) + v[1] * v[1];
#pragma unroll
  for (int q = 0; q < Q2; q++) {
    double vc = 0.0;  // dotProduct (range.d:57-67)
    if (c2x(q) == 1) vc += v[0]; else if (c2x(q) == -1) vc -= v[0];
    if (c2y(q) == 1) vc += v[1]; else if (c2y(q) == -1) vc -= v[1];
    double poly = 1.0 + (vc / css) + (vc * vc / css2p2) - (vdotv / css2);
    if (EQ == 1) poly = poly + (vc * vc * vc / css6p3) - (vdotv * vc / css2p2);
    dist[q] = rho0 * weight2(q) * poly;
  }
}

// force.d:238-309 at one collidable site: F[f1] -= sum_f2 sum_{q=1..8} w_q psi_f1(p) psi_f2(p + c_q | p) gcc c_q, then the
// wall term over q = 0..8 with the density ON the solid neighbour
__device__ __forceinline__ void force9(const Params2 &p, const double *__restrict__ psi, const double *__restrict__ rho,
                                       const uint8_t *__restrict__ mask, int x, int y, double (&F)[MAXC2][2]) {
  const int s = y * p.nx + x;
  for (int f1 = 0; f1 < p.ncomp; f1++) {
    F[f1][0] = 0.0; F[f1][1] = 0.0;
    const double psi1 = psi[f1 * p.n + s];
    for (int f2 = 0; f2 < p.ncomp; f2++) {
      const double cc = p.gcc[f1 * MAXC2 + f2];
      if (cc == 0.0) continue;
#pragma unroll
      for (int q = 1; q < Q2; q++) {
        const int nb = nbr2(p, x, y, c2x(q), c2y(q));
        const double psi2 = mask[nb] ? psi[f2 * p.n + s] : psi[f2 * p.n + nb];
        const double pref = weight2(q) * psi1 * psi2 * cc;
        if (c2x(q) == 1) F[f1][0] -= pref; else if (c2x(q) == -1) F[f1][0] += pref;
        if (c2y(q) == 1) F[f1][1] -= pref; else if (c2y(q) == -1) F[f1][1] += pref;
      }
    }
    const double wc = p.gwc[f1];
    if (wc == 0.0) continue;
#pragma unroll
    for (int q = 0; q < Q2; q++) {
      const int nb = nbr2(p, x, y, c2x(q), c2y(q));
      if (!mask[nb]) continue;
      const double pref = weight2(q) * psi1 * rho[f1 * p.n + nb] * wc;
      if (c2x(q) == 1) F[f1][0] -= pref; else if (c2x(q) == -1) F[f1][0] += pref;
      if (c2y(q) == 1) F[f1][1] -= pref; else if (c2y(q) == -1) F[f1][1] += pref;
    }
  }
}

// ---- kernels ----------------------------------------------------------------------------------------------
// K1: density and psi of the STREAMED populations of every site (density.d:89-96 / force.d:137-147 after advection)
template <int PSI>
__global__ void __launch_bounds__(128) psi2_kernel(Params2 p, const double *__restrict__ f_old, const uint8_t *__restrict__ mask,
                                                   double *__restrict__ rho, double *__restrict__ psi) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= p.n) return;
  const int x = s % p.nx, y = s / p.nx;
  for (int c = 0; c < p.ncomp; c++) {
    double f[Q2];
    pull9(p, f_old + (size_t)c * Q2 * p.n, mask, x, y, f);
    const double r = density9(f);
    rho[c * p.n + s] = r;
    psi[c * p.n + s] = psi2_of<PSI>(r);
  }
}

// K2: stream, force, collide, store (advection.d:61-89, force.d:238-309, collision.d:90-119,156-196)
template <int EQ>
__global__ void __launch_bounds__(128) fused2_kernel(Params2 p, const double *__restrict__ f_old, double *__restrict__ f_new,
                                                     const uint8_t *__restrict__ mask, const double *__restrict__ rho,
                                                     const double *__restrict__ psi) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= p.n) return;
  const int x = s % p.nx, y = s / p.nx;
  if (mask[s]) {  // solid: new[p] = old[p]
    for (int c = 0; c < p.ncomp; c++)
#pragma unroll
      for (int q = 0; q < Q2; q++) f_new[((size_t)c * Q2 + q) * p.n + s] = f_old[((size_t)c * Q2 + q) * p.n + s];
    return;
  }
  double F[MAXC2][2];
  for (int c = 0; c < MAXC2; c++) { F[c][0] = 0.0; F[c][1] = 0.0; }
  if (p.sc) force9(p, psi, rho, mask, x, y, F);
  double u[2] = {0.0, 0.0};
  if (EQ == 2) {  // BDist2 common velocity: (sum_f m_f / tau_f) / sum_f rho_f
    double wv[2] = {0.0, 0.0}, total = 0.0;
    for (int c = 0; c < p.ncomp; c++) {
      double f[Q2], m[2];
      pull9(p, f_old + (size_t)c * Q2 * p.n, mask, x, y, f);
      momentum9(f, m);
      wv[0] += m[0] / p.tau[c];
      wv[1] += m[1] / p.tau[c];
      total += rho[c * p.n + s];
    }
    u[0] = wv[0] / total;
    u[1] = wv[1] / total;
  }
  for (int c = 0; c < p.ncomp; c++) {
    double f[Q2], eq[Q2], dv[2];
    pull9(p, f_old + (size_t)c * Q2 * p.n, mask, x, y, f);
    const double den = rho[c * p.n + s];  // == density9(f): K1 stored it from the same gather
#pragma unroll
    for (int d = 0; d < 2; d++) {
      dv[d] = p.tau[c] * (p.gacc[d] + F[c][d] / den);
      if (EQ == 2) dv[d] += u[d];
    }
    eqdist9<EQ>(f, dv, den, eq);
#pragma unroll
    for (int q = 0; q < Q2; q++) f_new[((size_t)c * Q2 + q) * p.n + s] = f[q] * p.om1[c] + p.omega[c] * eq[q];
  }
}

// layout: Field.arr order [x][y][9] (dense staging) <-> f[c][q][y][x]
__global__ void aos_to_soa2_kernel(Params2 p, const double *__restrict__ aos, int c, double *__restrict__ fa, double *__restrict__ fb) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= p.n) return;
  const int x = s % p.nx, y = s / p.nx;
#pragma unroll
  for (int q = 0; q < Q2; q++) {
    const double v = aos[((size_t)x * p.ny + y) * Q2 + q];
    fa[((size_t)c * Q2 + q) * p.n + s] = v;
    fb[((size_t)c * Q2 + q) * p.n + s] = v;
  }
}
__global__ void soa_to_aos2_kernel(Params2 p, const double *__restrict__ f, int c, double *__restrict__ aos) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= p.n) return;
  const int x = s % p.nx, y = s / p.nx;
#pragma unroll
  for (int q = 0; q < Q2; q++) aos[((size_t)x * p.ny + y) * Q2 + q] = f[((size_t)c * Q2 + q) * p.n + s];
}
__global__ void mask_in2_kernel(Params2 p, const int32_t *__restrict__ in, uint8_t *__restrict__ out, int *flag) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= p.n) return;
  const int32_t v = in[(size_t)(s % p.nx) * p.ny + s / p.nx];
  if (v != 0 && v != 1) atomicOr(flag, 1);
  out[s] = v == 1;
}
__global__ void mask_out2_kernel(Params2 p, const uint8_t *__restrict__ in, int32_t *__restrict__ out) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= p.n) return;
  out[(size_t)(s % p.nx) * p.ny + s / p.nx] = in[s];
}

// outputs (lb/io.d:60-95): density of every site; colour, velocity zero on solid sites (density.d:284-295,
// velocity.d:106-117); force as computed by the last step (zero on solid sites), all in [x][y](,[2]) order
template <int FIELD>
__global__ void output2_kernel(Params2 p, const double *__restrict__ f, const uint8_t *__restrict__ mask, const double *__restrict__ rho,
                               const double *__restrict__ psi, int c1, int c2, int have_force, double *__restrict__ out) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= p.n) return;
  const int x = s % p.nx, y = s / p.nx;
  const size_t o = (size_t)x * p.ny + y;
  double a[Q2];
#pragma unroll
  for (int q = 0; q < Q2; q++) a[q] = f[((size_t)c1 * Q2 + q) * p.n + s];
  if (FIELD == F2_DENSITY) {
    out[o] = density9(a);
  } else if (FIELD == F2_COLOUR) {
    double b[Q2];
#pragma unroll
    for (int q = 0; q < Q2; q++) b[q] = f[((size_t)c2 * Q2 + q) * p.n + s];
    out[o] = mask[s] ? 0.0 : density9(a) - density9(b);
  } else if (FIELD == F2_VELOCITY) {
    double m[2] = {0.0, 0.0};
    if (!mask[s]) {
      momentum9(a, m);
      const double r = density9(a);
      m[0] /= r; m[1] /= r;
    }
    out[o * 2] = m[0]; out[o * 2 + 1] = m[1];
  } else {
    double F[MAXC2][2];
    for (int c = 0; c < MAXC2; c++) { F[c][0] = 0.0; F[c][1] = 0.0; }
    if (have_force && p.sc && !mask[s]) force9(p, psi, rho, mask, x, y, F);
    double fx = 0.0, fy = 0.0;
    for (int c = 0; c < p.ncomp; c++) if (c == c1) { fx = F[c][0]; fy = F[c][1]; }
    out[o * 2] = fx; out[o * 2 + 1] = fy;
  }
}

// density.d:136-180 / momentum.d: sums over physical fluid sites; one block, fixed order: deterministic
__global__ void __launch_bounds__(256) reduce2_kernel(Params2 p, const double *__restrict__ f, const uint8_t *__restrict__ mask, int c,
                                                      double *__restrict__ out4) {
  __shared__ double sh[4][256];
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  for (int s = threadIdx.x; s < p.n; s += blockDim.x) {
    if (mask[s]) continue;
    double a[Q2], m[2];
#pragma unroll
    for (int q = 0; q < Q2; q++) a[q] = f[((size_t)c * Q2 + q) * p.n + s];
    momentum9(a, m);
    acc[0] += density9(a); acc[1] += m[0]; acc[2] += m[1];
  }
  for (int k = 0; k < 4; k++) sh[k][threadIdx.x] = acc[k];
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) {
    if (threadIdx.x < w) for (int k = 0; k < 4; k++) sh[k][threadIdx.x] += sh[k][threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x < 4) out4[threadIdx.x] = sh[threadIdx.x][0];
}

// ---- host side --------------------------------------------------------------------------------------------
namespace {
int fail(std::string &err, const char *what, cudaError_t e) {
  err = std::string("dlbc_cuda (D2Q9): ") + what + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")";
  return 1000 + (int)e;
}
int bad(std::string &err, const char *msg) {
  err = std::string("dlbc_cuda (D2Q9): ") + msg;
  return 1;
}
#define CK2(expr)                                        \
  do {                                                   \
    cudaError_t e_ = (expr);                             \
    if (e_ != cudaSuccess) return fail(err, #expr, e_);  \
  } while (0)

inline unsigned blocks(const Lattice2 *L) { return (unsigned)((L->p.n + 127) / 128); }

// dense device array [nx][ny][K] <-> host array [nx + 2h][ny + 2h][K]: nx rows of ny * K elements
template <typename T>
cudaError_t copy2(const Lattice2 *L, T *host, T *dev, int halo, int K, bool to_device) {
  const Params2 &p = L->p;
  const size_t width = (size_t)p.ny * K * sizeof(T), hpitch = (size_t)(p.ny + 2 * halo) * K * sizeof(T);
  T *h0 = host + ((size_t)halo * (p.ny + 2 * halo) + halo) * K;
  return to_device ? cudaMemcpy2DAsync(dev, width, h0, hpitch, width, p.nx, cudaMemcpyHostToDevice, L->s)
                   : cudaMemcpy2DAsync(h0, hpitch, dev, width, width, p.nx, cudaMemcpyDeviceToHost, L->s);
}
}  // namespace

void destroy(Lattice2 *L) {
  if (!L) return;
  cudaSetDevice(L->device);
  if (L->s) cudaStreamSynchronize(L->s);
  for (double *q : {L->f[0], L->f[1], L->psi, L->rho, L->stage, L->partial}) if (q) cudaFree(q);
  if (L->mask) cudaFree(L->mask);
  if (L->flag) cudaFree(L->flag);
  if (L->s) cudaStreamDestroy(L->s);
  delete L;
}

int create(const dlbc_cuda_config &cfg, Lattice2 **out, std::string &err) {
  *out = nullptr;
  if (cfg.nranks != 1) return bad(err, "create: a 2-D lattice runs on one rank");
  if (cfg.gn[0] < 1 || cfg.gn[1] < 1 || cfg.gn[0] > 65535 || cfg.gn[1] > 65535) return bad(err, "create: gn out of range (1..65535)");
  if (cfg.gn[2] > 1) return bad(err, "create: gn[2] must be 0 or 1 on a 2-D lattice");
  if (cfg.gn[0] * cfg.gn[1] * cfg.ncomp * Q2 >= ((int64_t)1 << 31)) return bad(err, "create: nx*ny*ncomp*9 must be below 2^31");
  Lattice2 *L = new Lattice2();
  L->device = cfg.device;
  Params2 &p = L->p;
  p.nx = (int)cfg.gn[0]; p.ny = (int)cfg.gn[1]; p.n = p.nx * p.ny;
  p.ncomp = cfg.ncomp; p.sc = cfg.enable_shan_chen ? 1 : 0;
  for (int c = 0; c < cfg.ncomp; c++) {
    p.tau[c] = cfg.tau[c];
    p.omega[c] = 1.0 / cfg.tau[c];   // collision.d:99
    p.om1[c] = 1.0 - p.omega[c];
    p.gwc[c] = cfg.gwc[c];
    for (int c2 = 0; c2 < cfg.ncomp; c2++) p.gcc[c * MAXC2 + c2] = cfg.gcc[c * DLBC_CUDA_MAX_COMPONENTS + c2];
  }
  p.gacc[0] = cfg.global_acc[0]; p.gacc[1] = cfg.global_acc[1];
  L->psi_form = cfg.psi_form; L->eq_form = cfg.eqdist_form;
  const size_t fb = (size_t)p.ncomp * Q2 * p.n * sizeof(double), sb = (size_t)p.ncomp * p.n * sizeof(double);
  const size_t stb = (size_t)p.n * Q2 * sizeof(double);
#define CKD(expr)                                        \
  do {                                                   \
    cudaError_t e_ = (expr);                             \
    if (e_ != cudaSuccess) { const int rc_ = fail(err, #expr, e_); destroy(L); return rc_; } \
  } while (0)
  CKD(cudaSetDevice(L->device));
  CKD(cudaStreamCreateWithFlags(&L->s, cudaStreamNonBlocking));
  CKD(cudaMalloc(&L->f[0], fb));
  CKD(cudaMalloc(&L->f[1], fb));
  CKD(cudaMalloc(&L->psi, sb));
  CKD(cudaMalloc(&L->rho, sb));
  CKD(cudaMalloc(&L->mask, (size_t)p.n));
  CKD(cudaMalloc(&L->stage, stb));
  CKD(cudaMalloc(&L->partial, 4 * sizeof(double)));
  CKD(cudaMalloc(&L->flag, sizeof(int)));
  CKD(cudaMemsetAsync(L->f[0], 0, fb, L->s));
  CKD(cudaMemsetAsync(L->f[1], 0, fb, L->s));
  CKD(cudaMemsetAsync(L->psi, 0, sb, L->s));
  CKD(cudaMemsetAsync(L->rho, 0, sb, L->s));
  CKD(cudaMemsetAsync(L->mask, 0, (size_t)p.n, L->s));
  CKD(cudaStreamSynchronize(L->s));
#undef CKD
  L->bytes = (int64_t)(2 * fb + 2 * sb + p.n + stb + 4 * sizeof(double) + sizeof(int));
  *out = L;
  return 0;
}

int upload_populations(Lattice2 *L, int comp, const double *aos, int halo, std::string &err) {
  if (!aos || halo < 0 || comp < 0 || comp >= L->p.ncomp) return bad(err, "upload_populations: bad argument");
  CK2(copy2(L, const_cast<double *>(aos), L->stage, halo, Q2, true));
  aos_to_soa2_kernel<<<blocks(L), 128, 0, L->s>>>(L->p, L->stage, comp, L->f[0], L->f[1]);
  CK2(cudaGetLastError());
  L->nlaunch++;
  CK2(cudaStreamSynchronize(L->s));  // the host buffer is free again on return
  return 0;
}

int download_populations(Lattice2 *L, int comp, double *aos, int halo, std::string &err) {
  if (!aos || halo < 0 || comp < 0 || comp >= L->p.ncomp) return bad(err, "download_populations: bad argument");
  soa_to_aos2_kernel<<<blocks(L), 128, 0, L->s>>>(L->p, L->f[L->cur], comp, L->stage);
  CK2(cudaGetLastError());
  L->nlaunch++;
  CK2(copy2(L, aos, L->stage, halo, Q2, false));
  CK2(cudaStreamSynchronize(L->s));
  return 0;
}

int upload_mask(Lattice2 *L, const int32_t *mask, int halo, std::string &err) {
  if (!mask || halo < 0) return bad(err, "upload_mask: bad argument");
  int32_t *st = reinterpret_cast<int32_t *>(L->stage);
  CK2(copy2(L, const_cast<int32_t *>(mask), st, halo, 1, true));
  CK2(cudaMemsetAsync(L->flag, 0, sizeof(int), L->s));
  mask_in2_kernel<<<blocks(L), 128, 0, L->s>>>(L->p, st, L->mask, L->flag);
  CK2(cudaGetLastError());
  L->nlaunch++;
  int flag = 0;
  CK2(cudaMemcpyAsync(&flag, L->flag, sizeof(int), cudaMemcpyDeviceToHost, L->s));
  CK2(cudaStreamSynchronize(L->s));
  if (flag) return bad(err, "upload_mask: values other than 0 (None) and 1 (Solid) (lb/mask.d:84-93)");
  return 0;
}

int download_mask(Lattice2 *L, int32_t *mask, int halo, std::string &err) {
  if (!mask || halo < 0) return bad(err, "download_mask: bad argument");
  int32_t *st = reinterpret_cast<int32_t *>(L->stage);
  mask_out2_kernel<<<blocks(L), 128, 0, L->s>>>(L->p, L->mask, st);
  CK2(cudaGetLastError());
  L->nlaunch++;
  CK2(copy2(L, mask, st, halo, 1, false));
  CK2(cudaStreamSynchronize(L->s));
  return 0;
}

int step(Lattice2 *L, int nsteps, std::string &err) {
  if (nsteps < 0) return bad(err, "step: nsteps < 0");
  const unsigned nb = blocks(L);
  for (int t = 0; t < nsteps; t++) {
    const double *fo = L->f[L->cur];
    double *fn = L->f[1 - L->cur];
    if (L->psi_form == 0) psi2_kernel<0><<<nb, 128, 0, L->s>>>(L->p, fo, L->mask, L->rho, L->psi);
    else psi2_kernel<1><<<nb, 128, 0, L->s>>>(L->p, fo, L->mask, L->rho, L->psi);
    switch (L->eq_form) {
      case 0: fused2_kernel<0><<<nb, 128, 0, L->s>>>(L->p, fo, fn, L->mask, L->rho, L->psi); break;
      case 1: fused2_kernel<1><<<nb, 128, 0, L->s>>>(L->p, fo, fn, L->mask, L->rho, L->psi); break;
      default: fused2_kernel<2><<<nb, 128, 0, L->s>>>(L->p, fo, fn, L->mask, L->rho, L->psi); break;
    }
    CK2(cudaGetLastError());
    L->cur ^= 1;
    L->timestep++;
    L->nlaunch += 2;
  }
  return 0;
}

int sync(Lattice2 *L, std::string &err) {
  CK2(cudaStreamSynchronize(L->s));
  return 0;
}

int download_field(Lattice2 *L, int field, int c1, int c2, double *out, int halo, std::string &err) {
  const Params2 &p = L->p;
  if (!out || halo < 0 || c1 < 0 || c1 >= p.ncomp || c2 < 0 || c2 >= p.ncomp) return bad(err, "download: bad argument");
  // psi / rho hold the last step's fields: the force of a lattice that has not stepped yet is zero (resetForce)
  const int have = L->timestep > 0;
  const unsigned nb = blocks(L);
  const double *f = L->f[L->cur];
  switch (field) {
    case F2_DENSITY: output2_kernel<F2_DENSITY><<<nb, 128, 0, L->s>>>(p, f, L->mask, L->rho, L->psi, c1, c2, have, L->stage); break;
    case F2_COLOUR: output2_kernel<F2_COLOUR><<<nb, 128, 0, L->s>>>(p, f, L->mask, L->rho, L->psi, c1, c2, have, L->stage); break;
    case F2_VELOCITY: output2_kernel<F2_VELOCITY><<<nb, 128, 0, L->s>>>(p, f, L->mask, L->rho, L->psi, c1, c2, have, L->stage); break;
    case F2_FORCE: output2_kernel<F2_FORCE><<<nb, 128, 0, L->s>>>(p, f, L->mask, L->rho, L->psi, c1, c2, have, L->stage); break;
    default: return bad(err, "download: unknown field");
  }
  CK2(cudaGetLastError());
  L->nlaunch++;
  CK2(copy2(L, out, L->stage, halo, field >= F2_VELOCITY ? 2 : 1, false));
  CK2(cudaStreamSynchronize(L->s));
  return 0;
}

int reduce(Lattice2 *L, int comp, double out4[4], std::string &err) {
  if (comp < 0 || comp >= L->p.ncomp) return bad(err, "reduction: component out of range");
  reduce2_kernel<<<1, 256, 0, L->s>>>(L->p, L->f[L->cur], L->mask, comp, L->partial);
  CK2(cudaGetLastError());
  L->nlaunch++;
  CK2(cudaMemcpyAsync(out4, L->partial, 4 * sizeof(double), cudaMemcpyDeviceToHost, L->s));
  CK2(cudaStreamSynchronize(L->s));
  return 0;
}

int64_t launches(const Lattice2 *L) { return L->nlaunch; }
int64_t device_bytes(const Lattice2 *L) { return L->bytes; }

}  // namespace d2
}  // namespace dlbc

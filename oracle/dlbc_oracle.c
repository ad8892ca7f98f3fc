/*
 * dlbc_oracle.c -- CPU restatement of DLBC's lattice-Boltzmann timestep (see dlbc_oracle.h).
 *
 * TEST INFRASTRUCTURE ONLY: parity oracle + timed CPU baseline.  Not linked into the product.
 *
 * Build: see oracle/Makefile.  Must be compiled with -ffp-contract=off and without
 * -ffast-math: the reference (DMD, SSE2 doubles) performs every fp64 operation separately
 * and in source order, and the golden files are compared to the last bit where possible.
 *
 * Citations are relative to /root/reference/src/dlbc/.
 */
#include "dlbc_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define HALO 2 /* lattice.d:102: Field!(double[q], conn, 2) */
#define MAXC ORACLE_MAX_COMPONENTS

struct oracle_lattice {
  oracle_config cfg;
  int d, q;
  int64_t n[3];     /* physical */
  int64_t nh[3];    /* with halo (nh[2] = 1 for d = 2) */
  int64_t h[3];     /* halo per dimension (0 in the unused third dimension of 2-D) */
  int64_t sites_h;  /* product of nh */
  double *fluids[MAXC];   /* [sites_h][q]  (lattice.d:102) */
  double *advection;      /* shared scratch (lattice.d:126) */
  int32_t *mask;          /* [sites_h]     (mask.d:84-93) */
  double *density[MAXC];  /* [sites_h] */
  double *psi[MAXC];      /* [sites_h] */
  double *force[MAXC];    /* [sites_h][3] (only d used) */
  double *force_distributed;
  double *weighted_velocity; /* [sites_h][3]; collision.d:144-148 keeps q doubles, uses d */
  int density_fresh, psi_fresh;
};

/* ---- connectivity tables: lb/connectivity.d:372-418 (velocities), :250-270 (weights) ---- */
static const int D3Q19_C[19][3] = {
    {0, 0, 0},  {1, 0, 0},  {-1, 0, 0}, {0, 1, 0},   {0, -1, 0}, {0, 0, 1},  {0, 0, -1},
    {1, 1, 0},  {-1, 1, 0}, {0, 1, 1},  {0, -1, 1},  {1, 0, 1},  {1, 0, -1}, {1, -1, 0},
    {-1, -1, 0}, {0, 1, -1}, {0, -1, -1}, {-1, 0, 1}, {-1, 0, -1}};
static const int D2Q9_C[9][3] = {{0, 0, 0}, {1, 0, 0},  {-1, 0, 0}, {0, 1, 0}, {0, -1, 0},
                                 {1, 1, 0}, {1, -1, 0}, {-1, 1, 0}, {-1, -1, 0}};
static double D3Q19_W[19], D2Q9_W[9];
static int D3Q19_B[19], D2Q9_B[9];
static int tables_ready = 0;

/* generateBounce, lb/connectivity.d:128-142 */
static void make_bounce(int q, const int (*c)[3], int *bounce) {
  for (int i = 0; i < q; i++)
    for (int j = 0; j < q; j++)
      if (c[i][0] + c[j][0] == 0 && c[i][1] + c[j][1] == 0 && c[i][2] + c[j][2] == 0) {
        bounce[i] = j;
        break;
      }
}

static void init_tables(void) {
  if (tables_ready) return;
  D3Q19_W[0] = 1.0 / 3.0;
  for (int i = 1; i < 7; i++) D3Q19_W[i] = 1.0 / 18.0;
  for (int i = 7; i < 19; i++) D3Q19_W[i] = 1.0 / 36.0;
  D2Q9_W[0] = 4.0 / 9.0;
  for (int i = 1; i < 5; i++) D2Q9_W[i] = 1.0 / 9.0;
  for (int i = 5; i < 9; i++) D2Q9_W[i] = 1.0 / 36.0;
  make_bounce(19, D3Q19_C, D3Q19_B);
  make_bounce(9, D2Q9_C, D2Q9_B);
  tables_ready = 1;
}

/* Phobos std.algorithm.sum on a random-access fp range is pairwise (density.d:37-39 calls
 * it): s(r) = len<=2 ? direct : s(r[0..len/2]) + s(r[len/2..len]); the outer "0.0 + s" of the
 * seed is an identity.  Pinned by the golden population -> density files. */
static inline double sum_pairwise(const double *r, int len) {
  switch (len) {
    case 0: return 0.0;
    case 1: return r[0];
    case 2: return r[0] + r[1];
    default: return sum_pairwise(r, len / 2) + sum_pairwise(r + len / 2, len - len / 2);
  }
}

static inline int64_t site_index(const oracle_lattice *L, int64_t x, int64_t y, int64_t z) {
  return (x * L->nh[1] + y) * L->nh[2] + z;
}

/* mask predicates, lb/mask.d:180-214: all four are functions of Solid-ness only */
static inline int is_fluid(int32_t m) { return m == 0; }
static inline int is_advectable(int32_t m) { return m == 0; }
static inline int is_bounce_back(int32_t m) { return m == 1; }
static inline int is_collidable(int32_t m) { return m == 0; }

/* ---- the two instantiations of the sweeps ---------------------------------------------- */
#define CONN_D 3
#define CONN_Q 19
#define CONN_C D3Q19_C
#define CONN_W D3Q19_W
#define CONN_B D3Q19_B
#define FN(name) name##_d3q19
#include "dlbc_oracle_conn.inl"
#undef CONN_D
#undef CONN_Q
#undef CONN_C
#undef CONN_W
#undef CONN_B
#undef FN

#define CONN_D 2
#define CONN_Q 9
#define CONN_C D2Q9_C
#define CONN_W D2Q9_W
#define CONN_B D2Q9_B
#define FN(name) name##_d2q9
#include "dlbc_oracle_conn.inl"
#undef CONN_D
#undef CONN_Q
#undef CONN_C
#undef CONN_W
#undef CONN_B
#undef FN

#define DISPATCH(L, name, ...)                         \
  do {                                                 \
    if ((L)->d == 3) name##_d3q19(__VA_ARGS__);        \
    else name##_d2q9(__VA_ARGS__);                     \
  } while (0)

/* ---- lifecycle ------------------------------------------------------------------------- */
oracle_lattice *oracle_create(const oracle_config *cfg) {
  init_tables();
  if (!((cfg->d == 3 && cfg->q == 19) || (cfg->d == 2 && cfg->q == 9))) return NULL;
  if (cfg->ncomp < 1 || cfg->ncomp > MAXC) return NULL;
  oracle_lattice *L = (oracle_lattice *)calloc(1, sizeof(*L));
  if (!L) return NULL;
  L->cfg = *cfg;
  L->d = cfg->d;
  L->q = cfg->q;
  L->sites_h = 1;
  for (int i = 0; i < 3; i++) {
    int used = i < cfg->d;
    L->n[i] = used ? cfg->n[i] : 1;
    L->h[i] = used ? HALO : 0;
    L->nh[i] = L->n[i] + 2 * L->h[i];
    L->sites_h *= L->nh[i];
  }
#ifdef _OPENMP
  if (cfg->nthreads > 0) omp_set_num_threads(cfg->nthreads);
#endif
  size_t S = (size_t)L->sites_h;
  for (int f = 0; f < cfg->ncomp; f++) {
    L->fluids[f] = (double *)calloc(S * L->q, sizeof(double));
    L->density[f] = (double *)calloc(S, sizeof(double));
    L->psi[f] = (double *)calloc(S, sizeof(double));
    L->force[f] = (double *)calloc(S * 3, sizeof(double));
  }
  L->advection = (double *)calloc(S * L->q, sizeof(double));
  L->mask = (int32_t *)calloc(S, sizeof(int32_t));
  L->force_distributed = (double *)calloc(S * 3, sizeof(double));
  L->weighted_velocity = (double *)calloc(S * 3, sizeof(double));
  return L;
}

void oracle_destroy(oracle_lattice *L) {
  if (!L) return;
  for (int f = 0; f < L->cfg.ncomp; f++) {
    free(L->fluids[f]);
    free(L->density[f]);
    free(L->psi[f]);
    free(L->force[f]);
  }
  free(L->advection);
  free(L->mask);
  free(L->force_distributed);
  free(L->weighted_velocity);
  free(L);
}

int64_t oracle_sites_with_halo(const oracle_lattice *L) { return L->sites_h; }
int64_t oracle_sites_physical(const oracle_lattice *L) { return L->n[0] * L->n[1] * L->n[2]; }

/* copy between a user array (with halo `halo` in the used dimensions) and an internal one */
static void copy_field(const oracle_lattice *L, void *internal, void *user, int halo,
                       size_t elem_bytes, int to_internal) {
  int64_t uh[3], un[3];
  for (int i = 0; i < 3; i++) {
    uh[i] = i < L->d ? halo : 0;
    un[i] = L->n[i] + 2 * uh[i];
  }
  for (int64_t x = 0; x < un[0]; x++)
    for (int64_t y = 0; y < un[1]; y++) {
      int64_t ix = x - uh[0] + L->h[0], iy = y - uh[1] + L->h[1], iz = L->h[2] - uh[2];
      char *pi = (char *)internal + (size_t)site_index(L, ix, iy, iz) * elem_bytes;
      char *pu = (char *)user + (size_t)((x * un[1] + y) * un[2]) * elem_bytes;
      if (to_internal) memcpy(pi, pu, (size_t)un[2] * elem_bytes);
      else memcpy(pu, pi, (size_t)un[2] * elem_bytes);
    }
}

void oracle_set_populations(oracle_lattice *L, int comp, const double *aos, int halo) {
  copy_field(L, L->fluids[comp], (void *)aos, halo, sizeof(double) * L->q, 1);
  L->density_fresh = L->psi_fresh = 0;
}
void oracle_get_populations(const oracle_lattice *L, int comp, double *aos, int halo) {
  copy_field(L, L->fluids[comp], aos, halo, sizeof(double) * L->q, 0);
}
void oracle_set_mask(oracle_lattice *L, const int32_t *mask, int halo) {
  copy_field(L, L->mask, (void *)mask, halo, sizeof(int32_t), 1);
}
void oracle_get_mask(const oracle_lattice *L, int32_t *mask, int halo) {
  copy_field(L, L->mask, mask, halo, sizeof(int32_t), 0);
}

/* ---- halo exchange: fields/parallel.d:58-103 with nc = 1 in every dimension --------------
 * x, then y, then z; each phase sends the HALO outermost physical layers to the opposite
 * halo, over the FULL extent of the other dimensions (so earlier halos are included and
 * edges/corners come out right without diagonal messages). */
static void exchange_field(const oracle_lattice *L, void *arr, size_t elem_bytes) {
  char *a = (char *)arr;
  for (int dim = 0; dim < L->d; dim++) {
    int64_t n = L->n[dim], h = L->h[dim];
    int64_t stride = 1; /* sites per unit step in `dim` */
    for (int i = dim + 1; i < 3; i++) stride *= L->nh[i];
    int64_t outer = 1;
    for (int i = 0; i < dim; i++) outer *= L->nh[i];
    size_t run = (size_t)(h * stride) * elem_bytes; /* h consecutive layers are contiguous */
    for (int64_t o = 0; o < outer; o++) {
      char *base = a + (size_t)(o * L->nh[dim] * stride) * elem_bytes;
      /* upper physical layers [n, n+h) -> lower halo [0, h) */
      memcpy(base, base + (size_t)(n * stride) * elem_bytes, run);
      /* lower physical layers [h, 2h) -> upper halo [n+h, n+2h) */
      memcpy(base + (size_t)((n + h) * stride) * elem_bytes, base + (size_t)(h * stride) * elem_bytes, run);
    }
  }
}

/* lattice.d:200-214: every @Exchange member: fluids[], mask (elec fields are unallocated) */
void oracle_exchange_halo(oracle_lattice *L) {
  for (int f = 0; f < L->cfg.ncomp; f++) exchange_field(L, L->fluids[f], sizeof(double) * L->q);
  exchange_field(L, L->mask, sizeof(int32_t));
}

/* ---- the timestep ---------------------------------------------------------------------- */
void oracle_advect_fields(oracle_lattice *L) {
  DISPATCH(L, advect_fields, L);
  /* postAdvectionHooks: markDensitiesAsStale (density.d:357), markPsiAsStale (force.d:152) */
  L->density_fresh = L->psi_fresh = 0;
}

/* force.d:165-170 */
void oracle_reset_force(oracle_lattice *L) {
  size_t bytes = (size_t)L->sites_h * 3 * sizeof(double);
  for (int f = 0; f < L->cfg.ncomp; f++) memset(L->force[f], 0, bytes);
  memset(L->force_distributed, 0, bytes);
}

/* density.d:344-352 */
static void precalculate_densities(oracle_lattice *L) {
  if (L->density_fresh) return;
  DISPATCH(L, density_fields, L);
  L->density_fresh = 1;
}

/* force.d:137-147 */
static void precalculate_psi(oracle_lattice *L) {
  precalculate_densities(L);
  if (L->psi_fresh) return;
  for (int f = 0; f < L->cfg.ncomp; f++) {
    const double *rho = L->density[f];
    double *psi = L->psi[f];
    int form = L->cfg.psi_form;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < L->sites_h; i++) psi[i] = oracle_site_psi(form, rho[i]);
  }
  L->psi_fresh = 1;
}

void oracle_add_shan_chen_force(oracle_lattice *L) {
  if (!L->cfg.enable_shan_chen) return; /* force.d:224 */
  precalculate_psi(L);
  DISPATCH(L, add_shan_chen_force, L);
}

void oracle_distribute_force(oracle_lattice *L) {
  precalculate_densities(L);
  DISPATCH(L, distribute_force, L);
}

void oracle_prepare_to_collide(oracle_lattice *L) {
  if (L->cfg.eqdist_form != ORACLE_EQ_BDIST2) return; /* collision.d:47 */
  DISPATCH(L, weighted_velocity, L); /* first loop, collision.d:168-181 */
  precalculate_densities(L);
  DISPATCH(L, weighted_velocity_divide, L); /* second loop, collision.d:186-195 */
}

void oracle_collide_fields(oracle_lattice *L) {
  precalculate_densities(L);
  DISPATCH(L, collide_fields, L);
}

/* main.d:174-200 (elec steps return immediately when elec is off) */
void oracle_step(oracle_lattice *L, int nsteps) {
  for (int t = 0; t < nsteps; t++) {
    oracle_exchange_halo(L);
    oracle_advect_fields(L);
    oracle_reset_force(L);
    oracle_add_shan_chen_force(L);
    oracle_distribute_force(L);
    oracle_prepare_to_collide(L);
    oracle_collide_fields(L);
  }
}

/* ---- outputs --------------------------------------------------------------------------- */
#define FOR_PHYSICAL(L, x, y, z, k)                                                     \
  for (int64_t x = (L)->h[0], k = 0; x < (L)->h[0] + (L)->n[0]; x++)                    \
    for (int64_t y = (L)->h[1]; y < (L)->h[1] + (L)->n[1]; y++)                         \
      for (int64_t z = (L)->h[2]; z < (L)->h[2] + (L)->n[2]; z++, k++)

void oracle_density(oracle_lattice *L, int comp, double *out) {
  FOR_PHYSICAL(L, x, y, z, k) {
    out[k] = sum_pairwise(L->fluids[comp] + site_index(L, x, y, z) * L->q, L->q);
  }
}

void oracle_colour(oracle_lattice *L, int c1, int c2, double *out) {
  FOR_PHYSICAL(L, x, y, z, k) {
    int64_t s = site_index(L, x, y, z);
    if (is_fluid(L->mask[s]))
      out[k] = sum_pairwise(L->fluids[c1] + s * L->q, L->q) - sum_pairwise(L->fluids[c2] + s * L->q, L->q);
    else
      out[k] = 0.0;
  }
}

void oracle_velocity(oracle_lattice *L, int comp, double *out) {
  FOR_PHYSICAL(L, x, y, z, k) {
    int64_t s = site_index(L, x, y, z);
    const double *pop = L->fluids[comp] + s * L->q;
    double v[3] = {0.0, 0.0, 0.0};
    if (is_fluid(L->mask[s])) oracle_site_velocity(L->d, L->q, pop, sum_pairwise(pop, L->q), v);
    for (int i = 0; i < L->d; i++) out[k * L->d + i] = v[i];
  }
}

void oracle_force(const oracle_lattice *L, int comp, double *out) {
  FOR_PHYSICAL(L, x, y, z, k) {
    int64_t s = site_index(L, x, y, z);
    for (int i = 0; i < L->d; i++) out[k * L->d + i] = L->force[comp][s * 3 + i];
  }
}

void oracle_psi(const oracle_lattice *L, int comp, double *out) {
  FOR_PHYSICAL(L, x, y, z, k) { out[k] = L->psi[comp][site_index(L, x, y, z)]; }
}

/* density.d:136-147: physical fluid sites, sequential accumulation in memory order */
double oracle_total_mass(const oracle_lattice *L, int comp) {
  double mass = 0.0;
  FOR_PHYSICAL(L, x, y, z, k) {
    int64_t s = site_index(L, x, y, z);
    if (is_fluid(L->mask[s])) mass += sum_pairwise(L->fluids[comp] + s * L->q, L->q);
  }
  return mass;
}

/* momentum.d:129-145 */
void oracle_total_momentum(const oracle_lattice *L, int comp, double *out3) {
  double mom[3] = {0.0, 0.0, 0.0};
  const int(*c)[3] = L->d == 3 ? D3Q19_C : D2Q9_C;
  FOR_PHYSICAL(L, x, y, z, k) {
    int64_t s = site_index(L, x, y, z);
    if (!is_fluid(L->mask[s])) continue;
    const double *pop = L->fluids[comp] + s * L->q;
    double pm[3] = {0.0, 0.0, 0.0};
    for (int i = 0; i < L->q; i++)
      for (int j = 0; j < L->d; j++) pm[j] += pop[i] * (double)c[i][j];
    for (int j = 0; j < L->d; j++) mom[j] += pm[j];
  }
  for (int j = 0; j < 3; j++) out3[j] = mom[j];
}

/* ---- per-site entry points (unit-test KATs) --------------------------------------------- */
double oracle_site_density(int q, const double *pop) { return sum_pairwise(pop, q); }

void oracle_site_velocity(int d, int q, const double *pop, double rho, double *vel) {
  init_tables();
  if (d == 3 && q == 19) raw_velocity_d3q19(pop, rho, vel);
  else raw_velocity_d2q9(pop, rho, vel);
}

void oracle_site_eqdist(int d, int q, int form, const double *pop, const double *dv, double rho0,
                        double *dist) {
  init_tables();
  if (d == 3 && q == 19) eq_dist_d3q19(form, pop, dv, rho0, dist);
  else eq_dist_d2q9(form, pop, dv, rho0, dist);
}

/* force.d:320-332 */
double oracle_site_psi(int form, double rho) {
  return form == ORACLE_PSI_LINEAR ? rho : (1.0 - exp(-rho));
}

void oracle_tables(int d, int q, int64_t *velocities, int64_t *bounce, double *weights, double *css) {
  init_tables();
  const int(*c)[3] = d == 3 ? D3Q19_C : D2Q9_C;
  const double *w = d == 3 ? D3Q19_W : D2Q9_W;
  const int *b = d == 3 ? D3Q19_B : D2Q9_B;
  for (int i = 0; i < q; i++) {
    for (int j = 0; j < d; j++) velocities[i * d + j] = c[i][j];
    bounce[i] = b[i];
    weights[i] = w[i];
  }
  *css = 1.0 / 3.0; /* connectivity.d:145-183 */
}

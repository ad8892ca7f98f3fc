/*
 * dlbc_oracle_conn.inl -- the per-connectivity sweeps of the oracle (included twice by
 * dlbc_oracle.c: D3Q19 and D2Q9).  TEST INFRASTRUCTURE ONLY.
 *
 * Macros provided by the includer: CONN_D, CONN_Q, CONN_C (velocities), CONN_W (weights),
 * CONN_B (bounce), FN(name).  A 2-D lattice is held as nh[2] = 1 with no halo in z and
 * c_z = 0; only the first CONN_D components of vectors are ever touched, so the arithmetic is
 * exactly the reference's 2-D arithmetic.
 */

#define D CONN_D
#define Q CONN_Q

/* velocity.d:37-48 rawVelocity: sequential over q, every term (also the zero ones) */
static inline void FN(raw_velocity)(const double *pop, double density, double *vel) {
  for (int j = 0; j < D; j++) vel[j] = 0.0;
  for (int i = 0; i < Q; i++)
    for (int j = 0; j < D; j++) vel[j] += pop[i] * (double)CONN_C[i][j];
  for (int j = 0; j < D; j++) vel[j] /= density;
}

/* range.d:57-67 dotProduct: sum0 = 0; sum0 += a[d]*b[d] */
static inline double FN(dot)(const double *a, const double *b) {
  double s = 0;
  for (int j = 0; j < D; j++) s += a[j] * b[j];
  return s;
}

/* eqdist.d:61-278 */
static inline void FN(eq_dist)(int form, const double *pop, const double *dv, double rho0, double *dist) {
  if (form == ORACLE_EQ_SECOND || form == ORACLE_EQ_THIRD) {
    const double css = 1.0 / 3.0;
    double pv[D], v[D];
    FN(raw_velocity)(pop, rho0, pv);
    for (int j = 0; j < D; j++) v[j] = dv[j] + pv[j];
    const double vdotv = FN(dot)(v, v);
    const double css2 = 2.0 * css;
    const double css2p2 = css2 * css;
    if (form == ORACLE_EQ_SECOND) {
      for (int q = 0; q < Q; q++) {
        double cvq[D];
        for (int j = 0; j < D; j++) cvq[j] = (double)CONN_C[q][j];
        const double vdotcv = FN(dot)(v, cvq);
        dist[q] = rho0 * CONN_W[q] * (1.0 + (vdotcv / css) + (vdotcv * vdotcv / css2p2) - (vdotv / css2));
      }
    } else {
      const double css6p3 = 3.0 * css2p2 * css;
      for (int q = 0; q < Q; q++) {
        double cvq[D];
        for (int j = 0; j < D; j++) cvq[j] = (double)CONN_C[q][j];
        const double vdotcv = FN(dot)(v, cvq);
        dist[q] = rho0 * CONN_W[q] *
                  (1.0 + (vdotcv / css) + (vdotcv * vdotcv / css2p2) - (vdotv / css2) +
                   (vdotcv * vdotcv * vdotcv / css6p3) - (vdotv * vdotcv / css2p2));
      }
    }
    return;
  }
  /* BDist2: eqdist.d:101-184 (D3Q19) and :185-231 (D2Q9).  Depends on dv only. */
#if CONN_D == 3
  const double aN1 = 1.0 / 36.0, aN0 = 1.0 / 3.0;
#else
  const double aN1 = 1.0 / 24.0, aN0 = 1.0 / 2.0;
#endif
  const double haN1 = 0.5 * aN1;
  const double aN1_6 = aN1 / 6.0;
  const double T = 1.0 / 3.0;
  const double holdit = 1.0 / T;
  double tu[3] = {0.0, 0.0, 0.0}, tut[3] = {0.0, 0.0, 0.0};
  for (int j = 0; j < D; j++) {
    tu[j] = dv[j];
    tut[j] = tu[j] * holdit;
  }
#if CONN_D == 3
  const double uke = 0.5 * (tu[0] * tut[0] + tu[1] * tut[1] + tu[2] * tut[2]);
#else
  const double uke = 0.5 * (tu[0] * tut[0] + tu[1] * tut[1]);
#endif
  const double cst = 1.0 - uke;
  const double const1 = aN1 * cst;
  double cdotu, even, odd;
#define PAIR(expr, qp, qm, scale)                                  \
  cdotu = (expr);                                                  \
  even = const1 + haN1 * cdotu * cdotu;                            \
  odd = const1 * cdotu + aN1_6 * cdotu * cdotu * cdotu;            \
  dist[qp] = scale(even + odd);                                    \
  dist[qm] = scale(even - odd);
#define TWICE(x) 2.0 * (x)
#define ONCE(x) (x)
#if CONN_D == 3
  PAIR(tut[0], 1, 2, TWICE)
  PAIR(tut[1], 3, 4, TWICE)
  PAIR(tut[2], 5, 6, TWICE)
  PAIR(tut[0] + tut[1], 7, 14, ONCE)
  PAIR(tut[0] - tut[1], 13, 8, ONCE)
  PAIR(tut[0] + tut[2], 11, 18, ONCE)
  PAIR(tut[0] - tut[2], 12, 17, ONCE)
  PAIR(tut[1] + tut[2], 9, 16, ONCE)
  PAIR(tut[1] - tut[2], 15, 10, ONCE)
#else
  PAIR(tut[0], 1, 2, TWICE)
  PAIR(tut[1], 3, 4, TWICE)
  PAIR(tut[0] + tut[1], 5, 8, ONCE)
  PAIR(tut[0] - tut[1], 6, 7, ONCE)
#endif
#undef PAIR
#undef TWICE
#undef ONCE
  dist[0] = aN0 * cst;
  for (int q = 0; q < Q; q++) dist[q] *= rho0;
  (void)pop;
}

/* advection.d:91-99 */
static inline int FN(is_on_edge)(const oracle_lattice *L, int64_t x, int64_t y, int64_t z) {
  if (x == 0 || x == L->nh[0] - 1) return 1;
  if (y == 0 || y == L->nh[1] - 1) return 1;
#if CONN_D == 3
  if (z == 0 || z == L->nh[2] - 1) return 1;
#else
  (void)z;
#endif
  return 0;
}

/* advection.d:61-89: pull streaming with half-way bounce-back over every site of the full
 * array except the outermost layer; one scratch field shared by all fluids, swapped in. */
static void FN(advect_fields)(oracle_lattice *L) {
  for (int f = 0; f < L->cfg.ncomp; f++) {
    const double *old = L->fluids[f];
    double *adv = L->advection;
#pragma omp parallel for schedule(static)
    for (int64_t x = 0; x < L->nh[0]; x++)
      for (int64_t y = 0; y < L->nh[1]; y++)
        for (int64_t z = 0; z < L->nh[2]; z++) {
          if (FN(is_on_edge)(L, x, y, z)) continue;
          const int64_t p = site_index(L, x, y, z);
          double *pop = adv + p * Q;
          if (is_advectable(L->mask[p])) {
            for (int vq = 0; vq < Q; vq++) {
              const int64_t nb = site_index(L, x - CONN_C[vq][0], y - CONN_C[vq][1], z - CONN_C[vq][2]);
              if (is_bounce_back(L->mask[nb])) pop[vq] = old[p * Q + CONN_B[vq]];
              else pop[vq] = old[nb * Q + vq];
            }
          } else {
            for (int vq = 0; vq < Q; vq++) pop[vq] = old[p * Q + vq];
          }
        }
    /* swap(L.fluids[f], L.advection) */
    double *tmp = L->fluids[f];
    L->fluids[f] = L->advection;
    L->advection = tmp;
  }
}

/* density.d:89-96 over field.arr: every site, halo and solid included */
static void FN(density_fields)(oracle_lattice *L) {
  for (int f = 0; f < L->cfg.ncomp; f++) {
    const double *fl = L->fluids[f];
    double *rho = L->density[f];
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < L->sites_h; i++) rho[i] = sum_pairwise(fl + i * Q, Q);
  }
}

/* force.d:238-309 */
static void FN(add_shan_chen_force)(oracle_lattice *L) {
  const int C = L->cfg.ncomp;
  for (int f1 = 0; f1 < C; f1++) {
    double *force = L->force[f1];
    for (int f2 = 0; f2 < C; f2++) {
      const double cc = L->cfg.gcc[f1 * MAXC + f2];
      if (cc == 0.0) continue;
      const double *psi1 = L->psi[f1], *psi2 = L->psi[f2];
#pragma omp parallel for schedule(static)
      for (int64_t x = L->h[0]; x < L->h[0] + L->n[0]; x++)
        for (int64_t y = L->h[1]; y < L->h[1] + L->n[1]; y++)
          for (int64_t z = L->h[2]; z < L->h[2] + L->n[2]; z++) {
            const int64_t p = site_index(L, x, y, z);
            if (!is_collidable(L->mask[p])) continue;
            const double psiden1 = psi1[p];
            for (int vq = 1; vq < Q; vq++) { /* Iota!(1, q-1): start 1, count q-1 */
              const int64_t nb = site_index(L, x + CONN_C[vq][0], y + CONN_C[vq][1], z + CONN_C[vq][2]);
              const double psiden2 = is_bounce_back(L->mask[nb]) ? psi2[p] : psi2[nb];
              const double prefactor = CONN_W[vq] * psiden1 * psiden2 * cc;
              for (int vd = 0; vd < D; vd++) force[p * 3 + vd] -= prefactor * (double)CONN_C[vq][vd];
            }
          }
    }
    /* wall interactions, force.d:285-307 */
    const double wc = L->cfg.gwc[f1];
    if (wc == 0.0) continue;
    const double *psi1 = L->psi[f1], *rho1 = L->density[f1];
#pragma omp parallel for schedule(static)
    for (int64_t x = L->h[0]; x < L->h[0] + L->n[0]; x++)
      for (int64_t y = L->h[1]; y < L->h[1] + L->n[1]; y++)
        for (int64_t z = L->h[2]; z < L->h[2] + L->n[2]; z++) {
          const int64_t p = site_index(L, x, y, z);
          if (!is_collidable(L->mask[p])) continue;
          const double psiden1 = psi1[p];
          for (int vq = 0; vq < Q; vq++) {
            const int64_t nb = site_index(L, x + CONN_C[vq][0], y + CONN_C[vq][1], z + CONN_C[vq][2]);
            if (is_bounce_back(L->mask[nb])) {
              const double prefactor = CONN_W[vq] * psiden1 * rho1[nb] * wc;
              for (int vd = 0; vd < D; vd++) force[p * 3 + vd] -= prefactor * (double)CONN_C[vq][vd];
            }
          }
        }
  }
}

/* force.d:179-197: physical sites, no mask test */
static void FN(distribute_force)(oracle_lattice *L) {
  const int C = L->cfg.ncomp;
#pragma omp parallel for schedule(static)
  for (int64_t x = L->h[0]; x < L->h[0] + L->n[0]; x++)
    for (int64_t y = L->h[1]; y < L->h[1] + L->n[1]; y++)
      for (int64_t z = L->h[2]; z < L->h[2] + L->n[2]; z++) {
        const int64_t p = site_index(L, x, y, z);
        double total = 0.0;
        for (int f = 0; f < C; f++) total += L->density[f][p];
        for (int f = 0; f < C; f++)
          for (int vd = 0; vd < D; vd++)
            L->force[f][p * 3 + vd] += L->force_distributed[p * 3 + vd] * (L->density[f][p] / total);
      }
}

/* collision.d:168-181: wv = sum_f (sum_q f_f[q] c_q) / tau_f over physical sites */
static void FN(weighted_velocity)(oracle_lattice *L) {
  const int C = L->cfg.ncomp;
#pragma omp parallel for schedule(static)
  for (int64_t x = L->h[0]; x < L->h[0] + L->n[0]; x++)
    for (int64_t y = L->h[1]; y < L->h[1] + L->n[1]; y++)
      for (int64_t z = L->h[2]; z < L->h[2] + L->n[2]; z++) {
        const int64_t p = site_index(L, x, y, z);
        double *wv = L->weighted_velocity + p * 3;
        for (int vd = 0; vd < 3; vd++) wv[vd] = 0.0;
        for (int f = 0; f < C; f++) {
          const double *pop = L->fluids[f] + p * Q;
          double per_fluid[D];
          for (int vd = 0; vd < D; vd++) per_fluid[vd] = 0.0;
          for (int vq = 0; vq < Q; vq++)
            for (int vd = 0; vd < D; vd++) per_fluid[vd] += pop[vq] * (double)CONN_C[vq][vd];
          for (int vd = 0; vd < D; vd++) wv[vd] += per_fluid[vd] / L->cfg.tau[f];
        }
      }
}

/* collision.d:186-195 */
static void FN(weighted_velocity_divide)(oracle_lattice *L) {
  const int C = L->cfg.ncomp;
#pragma omp parallel for schedule(static)
  for (int64_t x = L->h[0]; x < L->h[0] + L->n[0]; x++)
    for (int64_t y = L->h[1]; y < L->h[1] + L->n[1]; y++)
      for (int64_t z = L->h[2]; z < L->h[2] + L->n[2]; z++) {
        const int64_t p = site_index(L, x, y, z);
        double total = 0.0;
        for (int f = 0; f < C; f++) total += L->density[f][p];
        for (int vd = 0; vd < D; vd++) L->weighted_velocity[p * 3 + vd] /= total;
      }
}

/* collision.d:90-119 */
static void FN(collide_fields)(oracle_lattice *L) {
  const int form = L->cfg.eqdist_form;
  for (int f = 0; f < L->cfg.ncomp; f++) {
    const double tau = L->cfg.tau[f];
    const double omega = 1.0 / tau;
    double *fl = L->fluids[f];
    const double *rho = L->density[f];
    const double *force = L->force[f];
#pragma omp parallel for schedule(static)
    for (int64_t x = L->h[0]; x < L->h[0] + L->n[0]; x++)
      for (int64_t y = L->h[1]; y < L->h[1] + L->n[1]; y++)
        for (int64_t z = L->h[2]; z < L->h[2] + L->n[2]; z++) {
          const int64_t p = site_index(L, x, y, z);
          if (!is_collidable(L->mask[p])) continue;
          double *pop = fl + p * Q;
          const double den = rho[p];
          double dv[D], eq[Q];
          for (int vd = 0; vd < D; vd++) {
            dv[vd] = tau * (L->cfg.global_acc[vd] + force[p * 3 + vd] / den);
            if (form == ORACLE_EQ_BDIST2) dv[vd] += L->weighted_velocity[p * 3 + vd];
          }
          FN(eq_dist)(form, pop, dv, den, eq);
          for (int vq = 0; vq < Q; vq++) pop[vq] = pop[vq] * (1.0 - omega) + omega * eq[vq];
        }
  }
}

#undef D
#undef Q

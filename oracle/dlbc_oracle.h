/*
 * dlbc_oracle.h -- CPU restatement of DLBC's lattice-Boltzmann timestep.
 *
 * TEST INFRASTRUCTURE ONLY.  This is the parity oracle and the timed CPU baseline; nothing
 * under dlbc_b200/ (the product) may include, link or call it.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it.
 *
 * It follows the reference's D code sweep by sweep, in the reference's own data layout
 * (AoS populations f[x][y][z][q] with a halo of 2, int32 mask) and floating-point operation
 * order.  Parity is PINNED: tests/test_oracle_golden.py replays the reference's golden HDF5
 * outputs (merging-droplets-3d, bdist2-sc-3d, bdist2-3d, checkpoint-3d, laplace-droplet-3d,
 * poiseuille-2d) through this code.
 *
 * The reference itself (D + MPICH + parallel HDF5 + two un-vendored D dependencies) cannot
 * be built in this image, so there is no oracle/_ref.
 */
#ifndef DLBC_ORACLE_H
#define DLBC_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORACLE_MAX_COMPONENTS 8

enum { ORACLE_PSI_LINEAR = 0, ORACLE_PSI_EXPONENTIAL = 1 };                 /* force.d:48-57 */
enum { ORACLE_EQ_SECOND = 0, ORACLE_EQ_THIRD = 1, ORACLE_EQ_BDIST2 = 2 };   /* eqdist.d:23-36 */

typedef struct {
  int d;                 /* 2 (D2Q9) or 3 (D3Q19) */
  int q;                 /* 9 or 19 */
  int64_t n[3];          /* physical extents; n[2] = 1 for d = 2 */
  int ncomp;             /* lb.components */
  double tau[ORACLE_MAX_COMPONENTS];
  double gcc[ORACLE_MAX_COMPONENTS * ORACLE_MAX_COMPONENTS]; /* row-major [f1][f2] */
  double gwc[ORACLE_MAX_COMPONENTS];
  double global_acc[3];
  int psi_form;
  int eqdist_form;
  int enable_shan_chen;
  int nthreads;          /* OpenMP threads (x-slabs stand in for MPI ranks); <=0: default */
} oracle_config;

typedef struct oracle_lattice oracle_lattice;

oracle_lattice *oracle_create(const oracle_config *cfg);
void oracle_destroy(oracle_lattice *L);

/* Sizes: halo is 2 in every one of the d dimensions (lattice.d:102). */
int64_t oracle_sites_with_halo(const oracle_lattice *L);
int64_t oracle_sites_physical(const oracle_lattice *L);

/* Populations / mask in the reference's memory layout (Field.arr).  halo = 2: the full
 * array including halo; halo = 0: physical sites only (halo left untouched / stripped). */
void oracle_set_populations(oracle_lattice *L, int comp, const double *aos, int halo);
void oracle_get_populations(const oracle_lattice *L, int comp, double *aos, int halo);
void oracle_set_mask(oracle_lattice *L, const int32_t *mask, int halo);
void oracle_get_mask(const oracle_lattice *L, int32_t *mask, int halo);

/* The statements of runTimeloop (src/main.d:179-195), one entry point each. */
void oracle_exchange_halo(oracle_lattice *L);        /* lattice.d:200-214, fields/parallel.d:40-143 */
void oracle_advect_fields(oracle_lattice *L);        /* lb/advection.d:35-89 */
void oracle_reset_force(oracle_lattice *L);          /* lb/force.d:165-170 */
void oracle_add_shan_chen_force(oracle_lattice *L);  /* lb/force.d:223-309 */
void oracle_distribute_force(oracle_lattice *L);     /* lb/force.d:179-197 */
void oracle_prepare_to_collide(oracle_lattice *L);   /* lb/collision.d:41-51,156-196 */
void oracle_collide_fields(oracle_lattice *L);       /* lb/collision.d:61-119 */
void oracle_step(oracle_lattice *L, int nsteps);     /* all of the above, in order */

/* Outputs as dumpLBData writes them (lb/io.d:56-101), halo stripped, C order. */
void oracle_density(oracle_lattice *L, int comp, double *out);              /* density.d:89-96 */
void oracle_colour(oracle_lattice *L, int c1, int c2, double *out);         /* density.d:284-295 */
void oracle_velocity(oracle_lattice *L, int comp, double *out);             /* velocity.d:106-117 */
void oracle_force(const oracle_lattice *L, int comp, double *out);          /* L.force[comp] */
void oracle_psi(const oracle_lattice *L, int comp, double *out);            /* L.psi[comp] (last step) */
double oracle_total_mass(const oracle_lattice *L, int comp);                /* density.d:136-147 */
void oracle_total_momentum(const oracle_lattice *L, int comp, double *out3);/* momentum.d:129-145 */

/* Per-site functions, for the reference's in-source known-answer unit tests. */
double oracle_site_density(int q, const double *pop);                                   /* density.d:37-39 */
void oracle_site_velocity(int d, int q, const double *pop, double rho, double *vel);    /* velocity.d:37-48 */
void oracle_site_eqdist(int d, int q, int form, const double *pop, const double *dv,
                        double rho0, double *dist);                                     /* eqdist.d:61-278 */
double oracle_site_psi(int form, double rho);                                           /* force.d:320-332 */
void oracle_tables(int d, int q, int64_t *velocities, int64_t *bounce, double *weights, double *css);

#ifdef __cplusplus
}
#endif
#endif

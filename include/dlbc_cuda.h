/*
 * dlbc_cuda.h -- C ABI of libdlbc_cuda.so: DLBC's lattice-Boltzmann timestep on B200 (sm_100a).
 *
 * This is the drop-in boundary for ONE path of DLBC: the per-timestep sequence of
 * runTimeloop (reference src/main.d:179-195): exchangeHalo -> advectFields -> resetForce ->
 * addShanChenForce -> distributeForce -> prepareToCollide -> collideFields, for the D3Q19
 * lattice, N-component Shan-Chen, BGK, bounce-back walls, periodic boundaries.  Parameter files,
 * plugin registration and HDF5 output stay on the host (D) side, which hands finished arrays to
 * this library and fetches results from it; mask/geometry setup and initial conditions may come
 * from the host as arrays or be generated on the device (dlbc_cuda_init_*).
 *
 * Conventions (modelled on how the reference treats its C libraries, src/dlbc/mpi.d:16-18 and
 * src/dlbc/io/hdf5.d:54-57): plain C types only; every function returns int, 0 = success,
 * non-zero = failure with a message available from dlbc_cuda_last_error(); the library never
 * calls exit().  One host thread drives one context (the reference's ranks are single
 * threaded); contexts are not thread-safe; there is no global mutable state besides the
 * last-error string of failed dlbc_cuda_create calls.  There is NO CPU fallback: without a
 * CUDA device every compute entry point fails.
 *
 * Host array layouts are exactly the reference's Field.arr (src/dlbc/fields/field.d:114-125):
 * C order, populations AoS double[x][y][z][19], scalars [x][y][z], vectors [x][y][z][3], mask
 * int32 (enum Mask: None = 0, Solid = 1; src/dlbc/lb/mask.d:84-93), with `halo` extra sites on
 * every side of every dimension (2 from the D host, 0 is accepted for convenience).  The
 * library keeps populations on the device as fp64 structure-of-arrays f[c][q][z][y][x].
 *
 * Decomposition: z-slabs.  Rank r of nranks owns global z in [r*gn[2]/nranks, (r+1)*gn[2]/nranks)
 * and all of x and y, i.e. the reference's M.nc = [1, 1, nranks], M.c = [0, 0, r]
 * (src/dlbc/parallel.d:190-237).  One process (or thread) per GPU, one context each.
 */
#ifndef DLBC_CUDA_H
#define DLBC_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DLBC_CUDA_ABI_VERSION 3
#define DLBC_CUDA_MAX_COMPONENTS 4
#define DLBC_CUDA_NCCL_ID_BYTES 128

/* enum PsiForm, src/dlbc/lb/force.d:48-57 */
enum { DLBC_PSI_LINEAR = 0, DLBC_PSI_EXPONENTIAL = 1 };
/* enum EqDistForm, src/dlbc/lb/eqdist.d:23-36 */
enum { DLBC_EQDIST_SECOND_ORDER = 0, DLBC_EQDIST_THIRD_ORDER = 1, DLBC_EQDIST_BDIST2 = 2 };
/* arithmetic mode */
enum {
  DLBC_ARITH_FAST = 0,  /* divisions by constants/shared denominators become reciprocal multiplies
                           (<= 1 ulp per operation; within the 1e-12 parity tolerance) */
  DLBC_ARITH_EXACT = 1  /* every fp64 operation as the reference performs it (IEEE divisions, same
                           order): bit-identical to the CPU restatement except for exp() */
};
/* step kernel variants (evidence for the design choice; results are identical bit for bit) */
enum {
  DLBC_KERNEL_AUTO = 0,       /* WAVE on one rank and on thick z-slabs; TWO_PASS on thin slabs and without
                                 Shan-Chen (measured choice, profiles/README.md) */
  DLBC_KERNEL_TWO_PASS = 1,   /* psi kernel + fused pull/force/collide kernel (psi via L1/L2) */
  DLBC_KERNEL_WAVE = 2,       /* psi sites + fused sites in ONE launch, ticket-ordered band by band so that the
                                 second gather of every population is an L2 hit; per-thread register gathers */
  DLBC_KERNEL_TMA = 3         /* the same schedule, warp-specialised: producer warps stage population rows and
                                 psi rows into shared memory with tensor-map TMA copies (cp.async.bulk.tensor +
                                 mbarrier), consumer warps compute from shared memory.  No mask path, even nx,
                                 at most two components: everything else must use AUTO or WAVE */
};

/* Replaces: the module-level parameters the timestep reads -- lattice.gn (lattice.d:28),
 * lb.components, lb.tau (lb/lb.d), lb.force.{globalAcc,enableShanChen,gcc,gwc,psiForm}
 * (lb/force.d:26-43), lb.eqdist.eqDistForm (lb/eqdist.d:18), parallel.nc (parallel.d:34). */
typedef struct dlbc_cuda_config {
  int32_t abi_version;            /* DLBC_CUDA_ABI_VERSION */
  int32_t d, q;                   /* 3, 19 (connectivity.d:21-32 default gconn) or 2, 9 (version(D2Q9), :27-29): a
                                     2-D context (csrc/d2q9.cu) runs on one rank and serves create / destroy,
                                     upload / download of populations and mask ([x][y][9], [x][y]), step, sync,
                                     timestep, download_density / colour / velocity / force ([x][y] and
                                     [x][y][2]), total_mass, total_momentum (z = 0), launch_count, device_bytes,
                                     kernel_variant; every other entry point fails on it with a message.
                                     Arithmetic there is always the reference's operation order */
  int64_t gn[3];                  /* global physical lattice size (gn[2]: 0 or 1 when d = 2) */
  int32_t ncomp;                  /* 1..DLBC_CUDA_MAX_COMPONENTS */
  double tau[DLBC_CUDA_MAX_COMPONENTS];
  double gcc[DLBC_CUDA_MAX_COMPONENTS * DLBC_CUDA_MAX_COMPONENTS]; /* row-major [f1][f2] */
  double gwc[DLBC_CUDA_MAX_COMPONENTS];
  double global_acc[3];
  int32_t psi_form;
  int32_t eqdist_form;
  int32_t enable_shan_chen;
  int32_t arithmetic;             /* DLBC_ARITH_* */
  int32_t kernel_variant;         /* DLBC_KERNEL_* */
  int32_t device;                 /* CUDA device ordinal for this context */
  int32_t rank, nranks;           /* z-slab index / count; nranks = 1: periodic in-process */
  uint8_t nccl_id[DLBC_CUDA_NCCL_ID_BYTES]; /* from dlbc_cuda_nccl_unique_id on rank 0, broadcast
                                               by the host (MPI_Bcast in D); unused if nranks = 1 */
} dlbc_cuda_config;

typedef struct dlbc_cuda_ctx dlbc_cuda_ctx;

/* ---- lifecycle ------------------------------------------------------------------------------
 * create replaces LType(gn, components, ...), prepareLBFields, prepareForce (main.d:119-124,
 * lb/init.d:60-72, lb/force.d:67-107): validates the parameters as prepareForce does, selects
 * the device, allocates the device fields.  Collective over all ranks when nranks > 1. */
int dlbc_cuda_abi_version(void);
int dlbc_cuda_device_count(int *count);
int dlbc_cuda_nccl_unique_id(uint8_t id[DLBC_CUDA_NCCL_ID_BYTES]);
int dlbc_cuda_create(const dlbc_cuda_config *cfg, dlbc_cuda_ctx **out);
int dlbc_cuda_destroy(dlbc_cuda_ctx *ctx);                 /* endMpi-time teardown */
const char *dlbc_cuda_last_error(const dlbc_cuda_ctx *ctx); /* ctx may be NULL (create failures) */

/* Pinned host memory for the arrays handed to upload/download (optional; any host memory works,
 * pinned memory makes the copies asynchronous and faster). */
int dlbc_cuda_host_alloc(void **ptr, size_t bytes);
int dlbc_cuda_host_free(void *ptr);

/* ---- state transfer -------------------------------------------------------------------------
 * upload_populations replaces the result of initFluids / readCheckpoint (lb/init.d:77-136,
 * io/checkpoint.d:80-103) for L.fluids[comp]; download_populations is what
 * dumpField("population-...") and dumpCheckpoint read (lb/io.d:58-62).  `aos` is this rank's
 * local block, Field.arr._data.ptr when halo = 2.  Only physical sites are transferred; the
 * library maintains its own ghost planes.  The *_xrange forms move x-layers
 * [x_begin, x_begin + x_count) only (`aos` then points at the first of those layers), so a host
 * can stream a lattice that does not fit in its memory. */
int dlbc_cuda_upload_populations(dlbc_cuda_ctx *ctx, int comp, const double *aos, int halo);
int dlbc_cuda_download_populations(dlbc_cuda_ctx *ctx, int comp, double *aos, int halo);
int dlbc_cuda_upload_populations_xrange(dlbc_cuda_ctx *ctx, int comp, const double *aos, int halo,
                                        int64_t x_begin, int64_t x_count);
int dlbc_cuda_download_populations_xrange(dlbc_cuda_ctx *ctx, int comp, double *aos, int halo,
                                          int64_t x_begin, int64_t x_count);
/* upload_mask replaces the result of initMask (lb/mask.d:134-172); values must be 0 or 1. */
int dlbc_cuda_upload_mask(dlbc_cuda_ctx *ctx, const int32_t *mask, int halo);
int dlbc_cuda_download_mask(dlbc_cuda_ctx *ctx, int32_t *mask, int halo);

/* ---- geometry and initial conditions generated on the device ---------------------------------
 * For lattices whose state should never exist on the host (1024^3 is 326 GB per copy).  Bit-identical to what
 * the reference's initialisers produce on this rank's block:
 * init_mask: initMask (lb/mask.d:134-172; fields/init.d:377-467) for MaskInit None / Box / Tube / Plates / Walls;
 *   `axis` = lb.mask.initAxis, `wall_offset` = wallOffset (Plates), `thickness` = wallThickness and
 *   has_wall[2 d + side] = dimensionHasWall[d][side] (Walls; may be NULL otherwise).
 * init_populations: initEqDist / initEqDistRandom / initEqDistPerturb / initEqDistPerturbFrac
 *   (fields/init.d:52-82) with the context's eqDistForm.  Random kinds draw one Phobos uniform() per site of the
 *   rank's FULL array including its halo of 2, in memory order (x slowest, z fastest), from Mt19937(seed) after
 *   discarding `skip` outputs -- seed = randomSeed[f] (+ rank if shiftSeedByRank) with skip = 0 for a reseeding
 *   fluid (lb/init.d:109-118), else random.seed (+ rank) with skip = the draws earlier fluids consumed.
 * init_wall: initEqDistWall (fields/init.d:469-477, lb/init.d:121) for every component: populations of solid
 *   sites := (1, 0, ..., 0).  Call after the mask and the populations are in place. */
enum { DLBC_MASK_NONE = 0, DLBC_MASK_BOX = 1, DLBC_MASK_TUBE = 2, DLBC_MASK_PLATES = 3, DLBC_MASK_WALLS = 4 };
enum { DLBC_INIT_EQDIST = 0, DLBC_INIT_EQDIST_RANDOM = 1, DLBC_INIT_EQDIST_PERTURB = 2, DLBC_INIT_EQDIST_PERTURB_FRAC = 3 };
int dlbc_cuda_init_mask(dlbc_cuda_ctx *ctx, int kind, int axis, int wall_offset, int thickness, const int32_t has_wall[6]);
int dlbc_cuda_init_populations(dlbc_cuda_ctx *ctx, int comp, int kind, double density, double perturb, uint32_t seed,
                               int64_t skip);
/* init_spheres: initEqDistSphere / TwoSpheres / Cylinder (fields/init.d:84-259): density1 within `radius` of the
 * nearest of the 1 or 2 centres (global coordinates, centres[3 k + d] = nc n / 2 + initOffset -+ initSeparation / 2),
 * density2 outside, a linear transition `thickness` wide (init.d:346-366); skip_axis >= 0 makes it a cylinder
 * along that axis (-1: spheres).  The *Frac variants are the same call with the host's scaled radius / offsets.
 * init_axis_profile: a density that depends on the coordinate along `axis` only (initEqDistLamellae,
 * init.d:261-313: the host evaluates the gn[axis] profile values). */
int dlbc_cuda_init_spheres(dlbc_cuda_ctx *ctx, int comp, int ncentres, const double *centres, int skip_axis, double radius,
                           double thickness, double density1, double density2);
int dlbc_cuda_init_axis_profile(dlbc_cuda_ctx *ctx, int comp, int axis, const double *values);
int dlbc_cuda_init_wall(dlbc_cuda_ctx *ctx);

/* ---- the hot path ---------------------------------------------------------------------------
 * step replaces the statements of runTimeloop, main.d:179-195, nsteps times.  Asynchronous:
 * returns after enqueueing on the context's streams; sync waits (and reports device errors). */
int dlbc_cuda_step(dlbc_cuda_ctx *ctx, int nsteps);
int dlbc_cuda_sync(dlbc_cuda_ctx *ctx);
int64_t dlbc_cuda_timestep(const dlbc_cuda_ctx *ctx); /* steps done since create */

/* ---- outputs (dumpLBData, lb/io.d:56-101) -- state after the last completed step -------------
 * density: densityField (density.d:89-96); colour: colourField (density.d:284-295), zero on
 * solid sites; velocity: rawVelocityField (velocity.d:106-117), zero on solid sites; force:
 * L.force[comp] of the last step (zero before the first step and on solid sites).  Physical
 * sites only; with halo > 0 the halo of `out` is left untouched. */
int dlbc_cuda_download_density(dlbc_cuda_ctx *ctx, int comp, double *out, int halo);
int dlbc_cuda_download_colour(dlbc_cuda_ctx *ctx, int comp1, int comp2, double *out, int halo);
int dlbc_cuda_download_velocity(dlbc_cuda_ctx *ctx, int comp, double *out, int halo);
int dlbc_cuda_download_force(dlbc_cuda_ctx *ctx, int comp, double *out, int halo);
/* pressure: density.d:306-326 evaluated on every site, P = css * sum_i (rho_i + sum_j 0.5 gcc[i][j] psi_i psi_j).
 * corrected_velocity: velocity.d:177-185, raw velocity + tau * F / rho with L.force of the last step; zero on
 * solid sites like velocityField.  volume_averaged_density: laplace.d:236-272, the mean density of every
 * component over the fluid sites of the cube offset +- half_width (global coordinates, not wrapped, summed in
 * the reference's site order) -- with total_mass the only lattice data dumpLaplace (laplace.d:55-222) needs, so
 * Laplace's surface tension is evaluated without downloading a field.  out: ncomp doubles. */
int dlbc_cuda_download_pressure(dlbc_cuda_ctx *ctx, double *out, int halo);
int dlbc_cuda_download_corrected_velocity(dlbc_cuda_ctx *ctx, int comp, double *out, int halo);
int dlbc_cuda_volume_averaged_density(dlbc_cuda_ctx *ctx, const int64_t offset[3], int half_width, double *out);
/* globalTotalMass (density.d:172-180) / globalTotalMomentum (momentum.d:172-180): sums over
 * physical fluid sites of all ranks; deterministic tree reduction (not the reference's
 * sequential order, so equal only to summation round-off). */
int dlbc_cuda_total_mass(dlbc_cuda_ctx *ctx, int comp, double *out);
int dlbc_cuda_total_momentum(dlbc_cuda_ctx *ctx, int comp, double out[3]);
/* mass_momentum_no_alloc: NO COUNTERPART IN THE REFERENCE -- added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  The same sums as total_mass / total_momentum (same kernels' grid and tree order: bit-identical to a
 * single-rank context on the same data), but context-free and allocation-free: the call performs no cudaMalloc,
 * cudaFree or any other allocating routine; it enqueues two kernels on `stream` (a cudaStream_t, NULL = the
 * default stream) of the CURRENT device and returns without synchronising.  All pointers are DEVICE pointers:
 *   d_pops[ncomp][19][nz][ny][nx]  populations, x fastest, directions in connectivity.d order
 *   d_mask[nz][ny][nx]             uint8, nonzero = not a fluid site (skipped); NULL = every site is fluid
 *   d_out[ncomp][4]                mass, momentum x, y, z of every component
 * The per-block partial sums live in one file-scope __device__ array of DLBC_CUDA_NO_ALLOC_SCRATCH_BYTES
 * (csrc/no_alloc.cu) that every GPU loading the library holds until the process exits.  Limits: ncomp <=
 * DLBC_CUDA_NO_ALLOC_MAX_COMPONENTS and nx * ny * nz <= DLBC_CUDA_NO_ALLOC_MAX_SITES; larger sizes are rejected
 * (non-zero return, message in err[0..err_cap), which may be NULL).  Every call on a device shares the array and
 * it keeps the last call's contents (nothing needs clearing: each call overwrites every partial it reads), so
 * two calls whose kernels can OVERLAP on the GPU (different streams without a dependency) RACE: serialise them. */
#define DLBC_CUDA_NO_ALLOC_MAX_COMPONENTS 4
#define DLBC_CUDA_NO_ALLOC_MAX_SITES ((int64_t)1 << 31)
#define DLBC_CUDA_NO_ALLOC_SCRATCH_BYTES (DLBC_CUDA_NO_ALLOC_MAX_COMPONENTS * 4 * 1184 * 8) /* 151552 */
int dlbc_cuda_mass_momentum_no_alloc(int ncomp, int64_t nx, int64_t ny, int64_t nz, const double *d_pops,
                                     const uint8_t *d_mask, double *d_out, void *stream, char *err, size_t err_cap);

/* ---- introspection / measurement --------------------------------------------------------------
 * tables: the device's copy of connectivity.d:401-418 / :198-212 / :253-257 (velocities[19][3],
 * bounce[19], weights[19], css).  launch_count: kernels launched by this context so far.
 * timed_steps: run nsteps between two CUDA events recorded on the context's compute stream and
 * return the elapsed milliseconds (synchronous).  kernel_times: run one step with events around
 * every kernel; fills names/ms for up to `cap` kernels and returns how many in *n. */
int dlbc_cuda_tables(int64_t *velocities, int64_t *bounce, double *weights, double *css);
int dlbc_cuda_launch_count(const dlbc_cuda_ctx *ctx, int64_t *count);
/* site_function: the per-site device functions every step kernel is built from (csrc/lb_site.cuh, EXACT
 * instantiation) on n sites, for the reference's in-source known-answer tests -- density (density.d:37-39,50-61),
 * raw velocity (velocity.d:37-48,59-74; in2 = n pre-computed densities or NULL), psi (force.d:320-332,370-384; in =
 * n densities, form = DLBC_PSI_*), equilibrium distribution (eqdist.d:61-184,317-360; in2 = dv[n][3], form =
 * DLBC_EQDIST_*).  in = pop[n][19] except for psi.  Host pointers; no context needed. */
enum { DLBC_SITE_DENSITY = 0, DLBC_SITE_VELOCITY = 1, DLBC_SITE_PSI = 2, DLBC_SITE_EQDIST = 3 };
int dlbc_cuda_site_function(int device, int fn, int form, int n, const double *in, const double *in2, double *out);
int dlbc_cuda_timed_steps(dlbc_cuda_ctx *ctx, int nsteps, double *elapsed_ms);
int dlbc_cuda_kernel_times(dlbc_cuda_ctx *ctx, int cap, const char **names, double *ms, int *n);
int dlbc_cuda_device_bytes(const dlbc_cuda_ctx *ctx, int64_t *bytes);
/* kernel_variant: the DLBC_KERNEL_* the next step will run (what AUTO resolved to).
 * set_option: tuning knobs of the single-launch schedules, for measurements and for tests that must reach
 * every scheduling regime on small lattices; 0 restores the default.  Names: "wave_tyb" (band height in
 * rows), "wave_la" (lag of the fused items behind the psi items, in band-planes, >= 2).  Results do not
 * depend on either.  Unknown names fail. */
int dlbc_cuda_kernel_variant(const dlbc_cuda_ctx *ctx, int *variant);
int dlbc_cuda_set_option(dlbc_cuda_ctx *ctx, const char *name, int64_t value);

#ifdef __cplusplus
}
#endif
#endif /* DLBC_CUDA_H */

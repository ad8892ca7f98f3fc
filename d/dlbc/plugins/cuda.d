// Written in the D programming language.

/**
   DLBC plugin: run the lattice-Boltzmann timestep on a B200 through libdlbc_cuda.so.

   UNCOMPILED SOURCE.  There is no D compiler (dmd/ldc2/gdc), MPICH or HDF5 in the image this
   repository is built in, so this module has never been compiled; it is the binding a DLBC
   maintainer would drop into `src/dlbc/plugins/` (see INTEGRATION.md).  The C ABI it binds,
   `include/dlbc_cuda.h`, is exercised instead by the ctypes host layer (`dlbc_b200/lib.py`,
   `dlbc_b200/lattice.py`) and by the test suite.

   How it plugs in (reference mechanisms only):
   $(UL
     $(LI module constructor registers an init function: `plugins/README.md:19-23`, `hooks.d:34-37`)
     $(LI `@("param")` variables below are picked up by `get-plugin-modules.sh` / `parameters.d:367-370`)
     $(LI upload happens in a `postLbInitHooks` callback (`lb/init.d:128-131`), i.e. after
          `initMask`/`initFluids` or `readCheckpoint` filled `L.fluids[]` and `L.mask`)
     $(LI `runTimeloop` (`main.d:174-200`) calls its eight statements directly and cannot be hooked,
          so `main.d` calls `runTimeloopCuda` below instead when `cuda.enable` is set - a two-line
          change shown in INTEGRATION.md)
   )
*/
module dlbc.plugins.cuda;

import dlbc.hooks;
import dlbc.lattice;
import dlbc.lb.lb;
import dlbc.lb.eqdist: eqDistForm;
import dlbc.lb.force: enableShanChen, gcc, gwc, globalAcc, psiForm;
import dlbc.lb.init: postLbInitHooks;
import dlbc.logging;
import dlbc.parallel;
import dlbc.mpi;

/// Run the timestep on the GPU.
@("param") bool enable = false;
/// CUDA device for rank r is `deviceOffset + r % devicesPerNode`.
@("param") int devicesPerNode = 8;
/// Ditto
@("param") int deviceOffset = 0;
/// 0 = FAST arithmetic (reciprocal multiplies, <= 1e-12 relative), 1 = EXACT (reference operation order).
@("param") int arithmetic = 0;
/// 0 = let the library choose the kernel variant.
@("param") int kernelVariant = 0;
/// Generate mask and initial populations on the device (dlbc_cuda_init_*) instead of uploading L.mask / L.fluids:
/// for lattices whose state should never exist on the host.  Supported: maskInit None/Box/Tube/Plates/Walls and
/// fluidInit EqDist, EqDistRandom, EqDistPerturb(Frac), EqDistSphere(Frac), EqDistTwoSpheres(Frac),
/// EqDistCylinder(Frac); the draws are those of THIS rank's block, as lb/init.d:109-125 makes them.
@("param") bool deviceInit = false;

// ---- extern(C) binding of include/dlbc_cuda.h (same style as src/dlbc/mpi.d:16-18) ----------------
@trusted nothrow @nogc extern(C) {
  enum DLBC_CUDA_ABI_VERSION = 3;
  enum DLBC_CUDA_MAX_COMPONENTS = 4;
  enum DLBC_CUDA_NCCL_ID_BYTES = 128;

  struct dlbc_cuda_config {
    int abi_version;
    int d, q;
    long[3] gn;
    int ncomp;
    double[DLBC_CUDA_MAX_COMPONENTS] tau;
    double[DLBC_CUDA_MAX_COMPONENTS * DLBC_CUDA_MAX_COMPONENTS] gcc;
    double[DLBC_CUDA_MAX_COMPONENTS] gwc;
    double[3] global_acc;
    int psi_form;
    int eqdist_form;
    int enable_shan_chen;
    int arithmetic;
    int kernel_variant;
    int device;
    int rank, nranks;
    ubyte[DLBC_CUDA_NCCL_ID_BYTES] nccl_id;
  }
  struct dlbc_cuda_ctx;

  int dlbc_cuda_abi_version();
  int dlbc_cuda_device_count(int* count);
  int dlbc_cuda_nccl_unique_id(ubyte* id);
  int dlbc_cuda_create(const(dlbc_cuda_config)* cfg, dlbc_cuda_ctx** ctx);
  int dlbc_cuda_destroy(dlbc_cuda_ctx* ctx);
  const(char)* dlbc_cuda_last_error(const(dlbc_cuda_ctx)* ctx);
  int dlbc_cuda_upload_populations(dlbc_cuda_ctx* ctx, int comp, const(double)* aos, int halo);
  int dlbc_cuda_download_populations(dlbc_cuda_ctx* ctx, int comp, double* aos, int halo);
  int dlbc_cuda_upload_mask(dlbc_cuda_ctx* ctx, const(int)* mask, int halo);
  int dlbc_cuda_step(dlbc_cuda_ctx* ctx, int nsteps);
  int dlbc_cuda_sync(dlbc_cuda_ctx* ctx);
  int dlbc_cuda_total_mass(dlbc_cuda_ctx* ctx, int comp, double* result);
  // no counterpart in the reference: allocation-free, context-free sums from device pointers
  int dlbc_cuda_mass_momentum_no_alloc(int ncomp, long nx, long ny, long nz, const(double)* dPops, const(ubyte)* dMask,
                                       double* dOut, void* stream, char* err, size_t errCap);
  int dlbc_cuda_download_density(dlbc_cuda_ctx* ctx, int comp, double* result, int halo);
  int dlbc_cuda_download_colour(dlbc_cuda_ctx* ctx, int comp1, int comp2, double* result, int halo);
  int dlbc_cuda_download_velocity(dlbc_cuda_ctx* ctx, int comp, double* result, int halo);
  int dlbc_cuda_download_force(dlbc_cuda_ctx* ctx, int comp, double* result, int halo);
  int dlbc_cuda_download_pressure(dlbc_cuda_ctx* ctx, double* result, int halo);
  int dlbc_cuda_download_corrected_velocity(dlbc_cuda_ctx* ctx, int comp, double* result, int halo);
  int dlbc_cuda_volume_averaged_density(dlbc_cuda_ctx* ctx, const(long)* offset, int halfWidth, double* result);
  // geometry / initial conditions generated on the device (lb/mask.d:134-172, fields/init.d:52-82,84-313,469-477)
  enum { DLBC_MASK_NONE, DLBC_MASK_BOX, DLBC_MASK_TUBE, DLBC_MASK_PLATES, DLBC_MASK_WALLS }  // = enum MaskInit order minus File
  enum { DLBC_INIT_EQDIST, DLBC_INIT_EQDIST_RANDOM, DLBC_INIT_EQDIST_PERTURB, DLBC_INIT_EQDIST_PERTURB_FRAC }
  int dlbc_cuda_init_mask(dlbc_cuda_ctx* ctx, int kind, int axis, int wallOffset, int thickness, const(int)* hasWall);
  int dlbc_cuda_init_populations(dlbc_cuda_ctx* ctx, int comp, int kind, double density, double perturb, uint seed, long skip);
  int dlbc_cuda_init_spheres(dlbc_cuda_ctx* ctx, int comp, int ncentres, const(double)* centres, int skipAxis,
                             double radius, double thickness, double density1, double density2);
  int dlbc_cuda_init_axis_profile(dlbc_cuda_ctx* ctx, int comp, int axis, const(double)* values);
  int dlbc_cuda_init_wall(dlbc_cuda_ctx* ctx);
}

private __gshared dlbc_cuda_ctx* ctx;

shared static this() {
  pluginRegister.registerPlugin(&initializePlugin);
}

void initializePlugin() {
  if ( enable ) {
    postLbInitHooks.registerFunction(&uploadLattice!(Lattice!gconn), "cuda-upload");
  }
}

private void check(int e, string what) {
  import std.string: fromStringz;
  if ( e != 0 ) {
    // same convention as the HDF5 calls, io/hdf5.d:54-57
    writeLogF("%s failed: %s", what, dlbc_cuda_last_error(ctx).fromStringz);
  }
}

/**
   Create the device context for this rank's z-slab and upload mask and populations.
   Requires the z-slab decomposition the library implements: `parallel.nc = [1, 1, M.size]`.
*/
void uploadLattice(T)(ref T L) if ( isLattice!T ) {
  alias conn = T.lbconn;
  static assert((conn.d == 3 && conn.q == 19) || (conn.d == 2 && conn.q == 9),
                "libdlbc_cuda implements D3Q19 and D2Q9 (a version(D2Q9) build: one rank).");
  static if ( conn.d == 3 ) {
    if ( M.nc[0] != 1 || M.nc[1] != 1 ) {
      writeLogF("cuda.enable requires parallel.nc = [1, 1, %d] (z-slabs).", M.size);
    }
  }
  else {
    if ( M.size != 1 ) writeLogF("cuda.enable on a 2-D lattice runs on one rank.");
  }
  dlbc_cuda_config cfg;
  cfg.abi_version = DLBC_CUDA_ABI_VERSION;
  cfg.d = conn.d; cfg.q = conn.q;
  foreach(immutable i; 0..conn.d) cfg.gn[i] = L.gn[i];
  cfg.ncomp = cast(int) L.fluids.length;
  foreach(immutable f; 0..L.fluids.length) {
    cfg.tau[f] = tau[f];
    cfg.gwc[f] = enableShanChen ? gwc[f] : 0.0;
    foreach(immutable f2; 0..L.fluids.length) {
      cfg.gcc[f * DLBC_CUDA_MAX_COMPONENTS + f2] = enableShanChen ? gcc[f][f2] : 0.0;
    }
  }
  foreach(immutable i; 0..conn.d) cfg.global_acc[i] = globalAcc[i];
  cfg.psi_form = cast(int) psiForm;        // enum order = DLBC_PSI_*   (lb/force.d:48-57)
  cfg.eqdist_form = cast(int) eqDistForm;  // enum order = DLBC_EQDIST_* (lb/eqdist.d:23-36)
  cfg.enable_shan_chen = enableShanChen ? 1 : 0;
  cfg.arithmetic = arithmetic;
  cfg.kernel_variant = kernelVariant;
  cfg.device = deviceOffset + M.rank % devicesPerNode;
  static if ( conn.d == 3 ) {
    cfg.rank = M.c[2];
    cfg.nranks = M.nc[2];
  }
  else {
    cfg.rank = 0;
    cfg.nranks = 1;
  }
  if ( M.size > 1 ) {
    if ( M.isRoot ) check(dlbc_cuda_nccl_unique_id(cfg.nccl_id.ptr), "dlbc_cuda_nccl_unique_id");
    MPI_Bcast(cfg.nccl_id.ptr, DLBC_CUDA_NCCL_ID_BYTES, MPI_BYTE, M.root, M.comm);
  }
  check(dlbc_cuda_create(&cfg, &ctx), "dlbc_cuda_create");
  // Field.arr._data.ptr is the [nx+4][ny+4][nz+4]([19]) block ([nx+4][ny+4]([9]) in 2-D) the library expects with halo = 2.
  check(dlbc_cuda_upload_mask(ctx, cast(const(int)*) L.mask.arr._data.ptr, 2), "dlbc_cuda_upload_mask");
  foreach(immutable f, ref field; L.fluids) {
    check(dlbc_cuda_upload_populations(ctx, cast(int) f, cast(const(double)*) field.arr._data.ptr, 2),
          "dlbc_cuda_upload_populations");
  }
}

/**
   Replacement for runTimeloop (main.d:174-200): the eight per-step statements become one call;
   populations come back to the host only when an output or checkpoint is due.
*/
void runTimeloopCuda(T)(ref T L) if ( isLattice!T ) {
  import dlbc.io.io: dumpData, nextOutputTimestep;  // nextOutputTimestep: see INTEGRATION.md
  import dlbc.timers;
  while ( timestep < timesteps ) {
    immutable int n = cast(int) (nextOutputTimestep(timestep, timesteps) - timestep);
    check(dlbc_cuda_step(ctx, n), "dlbc_cuda_step");
    timestep += n;
    foreach(immutable f, ref field; L.fluids) {
      check(dlbc_cuda_download_populations(ctx, cast(int) f, cast(double*) field.arr._data.ptr, 2),
            "dlbc_cuda_download_populations");
    }
    L.markDensitiesAsStale(); // density/psi caches refer to the old populations (lb/density.d)
    L.dumpData(timestep);
  }
  check(dlbc_cuda_destroy(ctx), "dlbc_cuda_destroy");
}

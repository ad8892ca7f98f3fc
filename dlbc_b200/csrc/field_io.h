// field_io.h -- launchers of the layout-conversion, output and reduction kernels (field_io.cu).
// None of these is on the per-step hot path.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include "lattice.cuh"

namespace dlbc {

constexpr int REDUCE_BLOCKS = 1184;  // 8 per SM on a 148-SM B200; fixed so reductions are deterministic

// Staged device arrays are dense and halo-free (the halo of the host's Field.arr is stripped or
// skipped by the strided host<->device copy): `aos` = [xc][ny][nz][19] for x-layers x0..x0+xc.
// soa_a / soa_b / soa: the whole population buffer f[zg][c][q][y][x]; c selects the component.
cudaError_t launch_aos_to_soa(const Geom &g, const double *aos, int xc, int x0, int c, double *soa_a, double *soa_b,
                              cudaStream_t s);
cudaError_t launch_soa_to_aos(const Geom &g, const double *soa, int xc, int x0, int c, double *aos, cudaStream_t s);

// mask: dense int32 [nx][ny][nz] <-> uint8 [zg][y][x]
cudaError_t launch_mask_in(const Geom &g, const int32_t *in, uint8_t *out, int *flags, cudaStream_t s);
cudaError_t launch_mask_out(const Geom &g, const uint8_t *in, int32_t *out, cudaStream_t s);

// ghost planes when the slab is its own z neighbour (nranks == 1), fields [narrays][zg][y][x]
cudaError_t launch_ghost_fill(const Geom &g, double *f, int narrays, cudaStream_t s);
cudaError_t launch_ghost_fill_mask(const Geom &g, uint8_t *m, cudaStream_t s);
// neighbourhood solid flags (Fields::nb) of the physical planes from a mask whose ghost planes are valid
cudaError_t launch_neighbour_bits(const Geom &g, const uint8_t *m, uint32_t *nb, cudaStream_t s);

// copy the populations of solid sites (all components, ghost planes included) from src to dst
cudaError_t launch_sync_solid(const Geom &g, const uint8_t *m, const double *src, double *dst, cudaStream_t s);

// derived outputs into a dense, halo-free device array [nx][ny][nz][K]
// kind 0 density(c1), 1 colour(c1 - c2), 2 raw velocity(c1), 3 copy of an SoA vector field [3][sites_g]
cudaError_t launch_output(const Geom &g, int kind, const double *f, const uint8_t *mask, int c1, int c2, double *out,
                          cudaStream_t s);
// pressure (density.d:306-326), corrected velocity (velocity.d:177-185) and the cube sums behind
// volumeAveragedDensity (laplace.d:236-272): see field_io.cu
cudaError_t launch_pressure(const Geom &g, const double *f, const StepParams &P, int psi_form, double *out, cudaStream_t s);
cudaError_t launch_corrected_velocity(const Geom &g, const double *f, const uint8_t *mask, int c, double tau, const double *F,
                                      double *out, cudaStream_t s);
cudaError_t launch_cube_density(const Geom &g, const double *f, const uint8_t *mask, int ncomp, int64_t z_first,
                                const int64_t lo[3], const int64_t hi[3], double *out, cudaStream_t s);

// lb_site.cuh's per-site functions on n sites (known-answer tests): see field_io.cu
cudaError_t launch_site_function(int fn, int form, int n, const double *in, const double *in2, double *out, cudaStream_t s);

// initial conditions / geometry generated on the device (fields/init.d:52-82,377-477; random.d:19-26)
cudaError_t launch_mt19937(uint32_t seed, int64_t skip, int64_t n, uint32_t *out, cudaStream_t s);
cudaError_t launch_init_populations(const Geom &g, int kind, double density, double a, double b_below, const double *eq,
                                    const uint32_t *raw, int c, double *fa, double *fb, cudaStream_t s);
cudaError_t launch_init_profile(const Geom &g, int kind, int ncentres, const double *centres, int skip_axis, int axis, double radius,
                                double width, double negval, double posval, const double *profile, int64_t z_first,
                                const double *eq, int c, double *fa, double *fb, cudaStream_t s);
cudaError_t launch_init_wall(const Geom &g, const uint8_t *m, double *fa, double *fb, cudaStream_t s);
cudaError_t launch_init_mask(const Geom &g, int kind, int axis, int wall_offset, int thickness, const int has_wall[6],
                             const int64_t gn[3], int64_t z_first, uint8_t *m, int *nsolid, cudaStream_t s);

// own-site density of every site incl. ghosts (SoA -> SoA), for the wall term
cudaError_t launch_site_density(const Geom &g, const double *f, int c, double *rho, cudaStream_t s);

// mass and momentum of one component over physical fluid sites: partial[4*REDUCE_BLOCKS + k], k = 0..3
cudaError_t launch_reduce(const Geom &g, const double *f, int c, const uint8_t *mask, double *partial, cudaStream_t s);

}  // namespace dlbc

// d2q9.h -- host interface of the 2-D (D2Q9) lattice (d2q9.cu).  dlbc_cuda.cu routes the entry points of a context
// created with d = 2, q = 9 here; every function returns 0 or an error code with the message in `err`.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <string>

#include "../../include/dlbc_cuda.h"

namespace dlbc {
namespace d2 {

struct Lattice2;  // device arrays + parameters of one 2-D lattice

enum Field2 { F2_DENSITY = 0, F2_COLOUR = 1, F2_VELOCITY = 2, F2_FORCE = 3 };

int create(const dlbc_cuda_config &cfg, Lattice2 **out, std::string &err);
void destroy(Lattice2 *L);
// aos: Field.arr of Lattice!(d2q9): [nx + 2 halo][ny + 2 halo][9]; mask: [nx + 2 halo][ny + 2 halo] int32
int upload_populations(Lattice2 *L, int comp, const double *aos, int halo, std::string &err);
int download_populations(Lattice2 *L, int comp, double *aos, int halo, std::string &err);
int upload_mask(Lattice2 *L, const int32_t *mask, int halo, std::string &err);
int download_mask(Lattice2 *L, int32_t *mask, int halo, std::string &err);
int step(Lattice2 *L, int nsteps, std::string &err);
int sync(Lattice2 *L, std::string &err);
// scalar fields [nx + 2 halo][ny + 2 halo], vector fields [..][..][2]; only the physical sites are written
int download_field(Lattice2 *L, int field, int c1, int c2, double *out, int halo, std::string &err);
int reduce(Lattice2 *L, int comp, double out4[4], std::string &err);  // mass, momentum x, y, 0
int64_t launches(const Lattice2 *L);
int64_t device_bytes(const Lattice2 *L);

}  // namespace d2
}  // namespace dlbc

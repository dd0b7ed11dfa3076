// tools/open_bc_emulate.cpp -- DEVELOPMENT TOOL, not part of the product and not a fallback.
//
// Steps through the SOURCE of open_bc_kernel (ludwig_b200/csrc/open_bc_kernels.cuh) on the
// host over the grid api.cu launches, for tools/check_open_bc_on_host.py to compare with the
// oracle. Nothing in ludwig_b200/ links this.
//
//   g++ -O1 -ffp-contract=off -w -std=c++17 -I/usr/local/cuda/include -shared -fPIC \
//       -o tools/_open_bc_emulate.so tools/open_bc_emulate.cpp

#include <cstring>
#include <vector>

#include <cuda_runtime.h>

static uint3 blockIdx, threadIdx;
static dim3 blockDim, gridDim;
#ifndef __launch_bounds__
#define __launch_bounds__(...)
#endif

#include "../ludwig_b200/csrc/open_bc_kernels.cuh"

static Layout make_layout(const int nlocal[3], int zalign) {
  Layout L;
  for (int d = 0; d < 3; d++) { L.n[d] = nlocal[d]; L.h[d] = 1; L.na[d] = nlocal[d] + 2; }
  L.z0 = zalign - L.h[2];
  i64 row = L.z0 + L.na[2];
  row = (row + zalign - 1)/zalign*zalign;
  L.s1 = row;
  L.s0 = L.s1*L.na[1];
  L.nsd = L.s0*L.na[0] + zalign;
  return L;
}

// host arrays in the reference layout over the one-site-halo box: rho[nsite], u[nsite][3],
// f[nsite][nvel] (in/out); cv[nvel][3], wv[nvel]
extern "C" int open_bc_emulate(const int nlocal[3], int nvel, const int * cv, const double * wv,
			       int flow_dir, const double u0[3], double rho0,
			       const double * rho, const double * u, double * f) {
  const Layout L = make_layout(nlocal, 16);
  std::vector<double> drho(L.nsd, 0.0), du(3*L.nsd, 0.0), df((size_t) nvel*L.nsd, 0.0);
  for (int i0 = 0; i0 < L.na[0]; i0++) for (int i1 = 0; i1 < L.na[1]; i1++) for (int i2 = 0; i2 < L.na[2]; i2++) {
    const i64 index = ((i64) i0*L.na[1] + i1)*L.na[2] + i2;
    const i64 site = i0*L.s0 + i1*L.s1 + L.z0 + i2;
    drho[site] = rho[index];
    for (int a = 0; a < 3; a++) du[a*L.nsd + site] = u[3*index + a];
    for (int p = 0; p < nvel; p++) df[(i64) p*L.nsd + site] = f[index*nvel + p];
  }
  OpenTab T;
  memset(&T, 0, sizeof(T));
  T.L = L; T.nvel = nvel; T.id = flow_dir;
  T.da = (T.id == 0) ? 1 : 0;
  T.db = (T.id == 2) ? 1 : 2;
  T.has_in = T.has_out = 1;
  for (int a = 0; a < 3; a++) { T.lo[a] = 0; T.hi[a] = L.n[a] - 1; T.u0[a] = u0[a]; }
  T.rho0 = rho0;
  for (int p = 0; p < nvel; p++) { for (int a = 0; a < 3; a++) T.cv[p][a] = (signed char) cv[3*p + a]; T.wv[p] = wv[p]; }
  dim3 block(128, 1, 1);
  dim3 grid((L.n[T.db] + 127)/128, L.n[T.da], 2);
  gridDim = grid; blockDim = block;
  for (unsigned bz = 0; bz < grid.z; bz++) for (unsigned by = 0; by < grid.y; by++) for (unsigned bx = 0; bx < grid.x; bx++)
    for (unsigned tx = 0; tx < block.x; tx++) {
      blockIdx.x = bx; blockIdx.y = by; blockIdx.z = bz;
      threadIdx.x = tx; threadIdx.y = 0; threadIdx.z = 0;
      open_bc_kernel(T, drho.data(), du.data(), df.data());
    }
  for (int i0 = 0; i0 < L.na[0]; i0++) for (int i1 = 0; i1 < L.na[1]; i1++) for (int i2 = 0; i2 < L.na[2]; i2++) {
    const i64 index = ((i64) i0*L.na[1] + i1)*L.na[2] + i2;
    const i64 site = i0*L.s0 + i1*L.s1 + L.z0 + i2;
    for (int p = 0; p < nvel; p++) f[index*nvel + p] = df[(i64) p*L.nsd + site];
  }
  return 0;
}

// tools/solid_grad_emulate.cpp -- DEVELOPMENT TOOL, not part of the product and not a fallback.
// Steps through the SOURCE of grad27_solid_kernel (ludwig_b200/csrc/solid_grad_kernels.cuh)
// on the host for tools/check_solid_grad_on_host.py.
//   g++ -O1 -ffp-contract=off -w -std=c++17 -I/usr/local/cuda/include -shared -fPIC \
//       -o tools/_solid_grad_emulate.so tools/solid_grad_emulate.cpp

#include <cstring>
#include <vector>

#include <cuda_runtime.h>

static uint3 blockIdx, threadIdx;
static dim3 blockDim, gridDim;
template <class T> static inline T __ldg(const T * p) { return *p; }
#ifndef __launch_bounds__
#define __launch_bounds__(...)
#endif

#include "../ludwig_b200/csrc/solid_grad_kernels.cuh"

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

// phi[nsite], grad[nsite][3], delsq[nsite] in the reference layout over the one-site-halo box
extern "C" int solid_grad_emulate(const int nlocal[3], const int walls[3], const double * phi,
				  double * grad, double * delsq) {
  const Layout L = make_layout(nlocal, 16);
  std::vector<double> dphi(L.nsd, 0.0), dgrad(3*L.nsd, 0.0), ddelsq(L.nsd, 0.0);
  for (int i0 = 0; i0 < L.na[0]; i0++) for (int i1 = 0; i1 < L.na[1]; i1++) for (int i2 = 0; i2 < L.na[2]; i2++)
    dphi[i0*L.s0 + i1*L.s1 + L.z0 + i2] = phi[((i64) i0*L.na[1] + i1)*L.na[2] + i2];
  SolidGradTab T;
  memset(&T, 0, sizeof(T));
  T.L = L;
  for (int d = 0; d < 3; d++) { T.wrap[d] = walls[d] ? 0 : 1; T.wall_lo[d] = T.wall_hi[d] = walls[d]; }
  dim3 block(128, 1, 1);
  dim3 grid((L.n[2] + 127)/128, L.n[1], L.n[0]);
  gridDim = grid; blockDim = block;
  for (unsigned bz = 0; bz < grid.z; bz++) for (unsigned by = 0; by < grid.y; by++) for (unsigned bx = 0; bx < grid.x; bx++)
    for (unsigned tx = 0; tx < block.x; tx++) {
      blockIdx.x = bx; blockIdx.y = by; blockIdx.z = bz;
      threadIdx.x = tx; threadIdx.y = 0; threadIdx.z = 0;
      grad27_solid_kernel(T, dphi.data(), dgrad.data(), ddelsq.data());
    }
  for (int i0 = 0; i0 < L.na[0]; i0++) for (int i1 = 0; i1 < L.na[1]; i1++) for (int i2 = 0; i2 < L.na[2]; i2++) {
    const i64 index = ((i64) i0*L.na[1] + i1)*L.na[2] + i2;
    const i64 site = i0*L.s0 + i1*L.s1 + L.z0 + i2;
    for (int a = 0; a < 3; a++) grad[3*index + a] = dgrad[a*L.nsd + site];
    delsq[index] = ddelsq[site];
  }
  return 0;
}

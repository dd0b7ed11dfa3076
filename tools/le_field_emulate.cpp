// tools/le_field_emulate.cpp -- DEVELOPMENT TOOL, not part of the product and not a fallback.
//
// Steps through the SOURCE of le_phi_buffer_kernel and le_grad_fix_kernel
// (ludwig_b200/csrc/le_field_kernels.cuh) and of grad7_kernel / grad27_kernel
// (common_kernels.cuh) on the host, thread by thread over the grids api.cu launches, so
// that the order parameter across Lees-Edwards planes can be compared with the oracle in a
// container without a GPU (tools/check_le_field_on_host.py). Nothing in ludwig_b200/
// links this.
//
//   g++ -O1 -ffp-contract=off -w -std=c++17 -I/usr/local/cuda/include -shared -fPIC \
//       -o tools/_le_field_emulate.so tools/le_field_emulate.cpp

#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include <cuda_runtime.h>

static uint3 blockIdx, threadIdx;
static dim3 blockDim, gridDim;
template <class T> static inline T __ldg(const T * p) { return *p; }
static inline double __dmul_rn(double a, double b) { return a*b; }
static inline double __dadd_rn(double a, double b) { return a + b; }
using std::min;
using std::max;
#ifndef __launch_bounds__
#define __launch_bounds__(...)
#endif

#include "../ludwig_b200/csrc/le_field_kernels.cuh"

// the layout of b200lb create_impl (api.cu) for a three-dimensional lattice
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

template <class F> static void for_grid(dim3 grid, dim3 block, F body) {
  gridDim = grid; blockDim = block;
  for (unsigned bz = 0; bz < grid.z; bz++) for (unsigned by = 0; by < grid.y; by++) for (unsigned bx = 0; bx < grid.x; bx++) {
    blockIdx.x = bx; blockIdx.y = by; blockIdx.z = bz;
    for (unsigned ty = 0; ty < block.y; ty++) for (unsigned tx = 0; tx < block.x; tx++) {
      threadIdx.x = tx; threadIdx.y = ty; threadIdx.z = 0;
      body();
    }
  }
}

// What le_grad_fix (api.cu) does after the stencil kernel has run: the sites next to a
// plane get grad and delsq again. grad[nsite][3], delsq[nsite] come in holding the
// plane-unaware values (host layout, halo-inclusive) and are updated in place; buf_out
// [2 nplanes][nall_y][nall_z] receives the buffer planes. wrap[d]: direction d periodic
// within the rank (neighbours wrap; else phi's halo is read).
extern "C" int emu_le_grad_fix(const int nlocal[3], const int wrap[3], int nplanes, const int * loc,
			       double uy, int oscillatory, int period, int nt0, double t, int npt,
			       const double * phi, double * grad, double * delsq, double * buf_out) {
  const Layout L = make_layout(nlocal, 16);
  std::vector<double> dphi(L.nsd, 0.0), dgrad(3*L.nsd, 0.0), ddelsq(L.nsd, 0.0);
  for (int i0 = 0; i0 < L.na[0]; i0++) for (int i1 = 0; i1 < L.na[1]; i1++) for (int i2 = 0; i2 < L.na[2]; i2++) {
    const i64 index = ((i64) i0*L.na[1] + i1)*L.na[2] + i2;
    const i64 site = i0*L.s0 + i1*L.s1 + L.z0 + i2;
    dphi[site] = phi[index];
    for (int a = 0; a < 3; a++) dgrad[a*L.nsd + site] = grad[3*index + a];
    ddelsq[site] = delsq[index];
  }
  LePhiTab T;
  memset(&T, 0, sizeof(T));
  T.L = L;
  T.w0 = wrap[0]; T.w1 = wrap[1]; T.w2 = wrap[2];
  le_phi_tab_fill(T, nplanes, loc, uy, oscillatory, period, nt0, t);
  std::vector<double> buf((size_t) T.nbuf*L.s0, -999.0);
  dim3 block(128, 1, 1);
  dim3 gbuf((L.n[2] + 2 + 127)/128, L.n[1] + 2, T.nbuf);
  for_grid(gbuf, block, [&]() { le_phi_buffer_kernel(T, dphi.data(), buf.data()); });
  dim3 gfix((L.n[2] + 127)/128, L.n[1], T.nbuf);
  if (npt == 27) for_grid(gfix, block, [&]() { le_grad_fix_kernel<27>(T, dphi.data(), buf.data(), dgrad.data(), ddelsq.data()); });
  else for_grid(gfix, block, [&]() { le_grad_fix_kernel<7>(T, dphi.data(), buf.data(), dgrad.data(), ddelsq.data()); });
  for (int i0 = 0; i0 < L.na[0]; i0++) for (int i1 = 0; i1 < L.na[1]; i1++) for (int i2 = 0; i2 < L.na[2]; i2++) {
    const i64 index = ((i64) i0*L.na[1] + i1)*L.na[2] + i2;
    const i64 site = i0*L.s0 + i1*L.s1 + L.z0 + i2;
    for (int a = 0; a < 3; a++) grad[3*index + a] = dgrad[a*L.nsd + site];
    delsq[index] = ddelsq[site];
  }
  for (int k = 0; k < T.nbuf; k++) for (int i1 = 0; i1 < L.na[1]; i1++) for (int i2 = 0; i2 < L.na[2]; i2++) {
    buf_out[((i64) k*L.na[1] + i1)*L.na[2] + i2] = buf[(i64) k*L.s0 + i1*L.s1 + L.z0 + i2];
  }
  return 0;
}

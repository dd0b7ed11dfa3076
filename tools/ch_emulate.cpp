// tools/ch_emulate.cpp -- DEVELOPMENT TOOL, not part of the product and not a fallback.
//
// Steps through the SOURCE of ch_kernel (ludwig_b200/csrc/ch_kernels.cu) on the host,
// thread by thread over the same grid the launcher uses, so that its indexing and
// arithmetic can be compared with the oracle in a container without a GPU
// (tools/check_ch_kernel_on_host.py). Nothing in ludwig_b200/ links this.
//
//   g++ -O1 -ffp-contract=off -w -I/usr/local/cuda/include -shared -fPIC -DCH_EXACT=1 \
//       -o tools/_ch_emulate_exact.so tools/ch_emulate.cpp        (and -DCH_EXACT=0: _fast.so)

#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include <cuda_runtime.h>

static uint3 blockIdx, threadIdx;
template <class T> static inline T __ldg(const T * p) { return *p; }
static inline double __dmul_rn(double a, double b) { return a*b; }
static inline double __dadd_rn(double a, double b) { return a + b; }
static inline double __fma_rn(double a, double b, double c) { return std::fma(a, b, c); }
using std::min;
#ifndef __launch_bounds__
#define __launch_bounds__(...)
#endif

#define B200LB_KERNEL_SOURCE_ONLY
#include "../ludwig_b200/csrc/ch_kernels.cu"

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

static void to_device(const Layout & L, const double * host, int nf, std::vector<double> & dev) {
  dev.assign((size_t) L.nsd*nf, 0.0);
  for (int i0 = 0; i0 < L.na[0]; i0++) for (int i1 = 0; i1 < L.na[1]; i1++) for (int i2 = 0; i2 < L.na[2]; i2++) {
    const i64 index = ((i64) i0*L.na[1] + i1)*L.na[2] + i2;
    const i64 site = i0*L.s0 + i1*L.s1 + L.z0 + i2;
    for (int a = 0; a < nf; a++) dev[a*L.nsd + site] = host[index*nf + a];
  }
}

static void to_host(const Layout & L, const std::vector<double> & dev, int nf, double * host) {
  for (int i0 = 0; i0 < L.na[0]; i0++) for (int i1 = 0; i1 < L.na[1]; i1++) for (int i2 = 0; i2 < L.na[2]; i2++) {
    const i64 index = ((i64) i0*L.na[1] + i1)*L.na[2] + i2;
    const i64 site = i0*L.s0 + i1*L.s1 + L.z0 + i2;
    for (int a = 0; a < nf; a++) host[index*nf + a] = dev[a*L.nsd + site];
  }
}

template <int METHOD, int ORDER>
static void run(const ChParams & P) {
  const Layout & L = P.L;
  const unsigned gx = (L.n[2] + CH_TZ - 1)/CH_TZ, gy = (L.n[1] + CH_TY - 1)/CH_TY, gz = (L.n[0] + P.nx - 1)/P.nx;
  for (unsigned bz = 0; bz < gz; bz++) for (unsigned by = 0; by < gy; by++) for (unsigned bx = 0; bx < gx; bx++)
    for (unsigned ty = 0; ty < CH_TY; ty++) for (unsigned tx = 0; tx < CH_TZ; tx++) {
      blockIdx.x = bx; blockIdx.y = by; blockIdx.z = bz;
      threadIdx.x = tx; threadIdx.y = ty; threadIdx.z = 0;
      ch_kernel<CH_EXACT, METHOD, ORDER>(P);
    }
}

template <int METHOD>
static int run_order(const ChParams & P, int order) {
  switch (order) {
  case 0: run<METHOD, 0>(P); return 0;
  case 1: run<METHOD, 1>(P); return 0;
  case 2: run<METHOD, 2>(P); return 0;
  case 3: run<METHOD, 3>(P); return 0;
  case 4: run<METHOD, 4>(P); return 0;
  }
  return -1;
}

// Host arrays in the reference layout over the one-site-halo box: phi[nsite], grad[nsite][3],
// delsq[nsite], u[nsite][3], force[nsite][3] (in/out), phi_out[nsite] (out).
// wrap = 0 reads the halo planes as they are (what a decomposed direction does).
extern "C" int ch_emulate(const int nlocal[3], const int wrap[3], int method, int order,
			  const double prm[4], const double gmu[3], const double fconst[3],
			  int accumulate, const double * phi, const double * grad,
			  const double * delsq, const double * u, double * force, double * phi_out) {
  Layout L = make_layout(nlocal, 16);
  std::vector<double> dphi, dgrad, ddelsq, du, dforce, dout;
  to_device(L, phi, 1, dphi); to_device(L, grad, 3, dgrad); to_device(L, delsq, 1, ddelsq);
  to_device(L, u, 3, du); to_device(L, force, 3, dforce); to_device(L, phi, 1, dout);
  ChParams P;
  memset(&P, 0, sizeof(P));
  P.L = L;
  for (int d = 0; d < 3; d++) { P.wrap[d] = wrap[d]; P.gmu[d] = gmu[d]; P.fconst[d] = fconst[d]; }
  P.phi = dphi.data(); P.grad = dgrad.data(); P.delsq = ddelsq.data(); P.u = du.data();
  P.force = dforce.data(); P.phi_out = dout.data();
  P.A = prm[0]; P.B = prm[1]; P.kappa = prm[2]; P.mobility = prm[3];
  P.accumulate = accumulate; P.order = order; P.method = method; P.nx = 3;
  int rc = -1;
  if (method == 0 || method == CH_FORCE_NONE) rc = run_order<0>(P, order);
  if (method == CH_FORCE_STRESS_DIVERGENCE) rc = run_order<CH_FORCE_STRESS_DIVERGENCE>(P, order);
  if (method == CH_FORCE_PHI_GRADMU) rc = run_order<CH_FORCE_PHI_GRADMU>(P, order);
  to_host(L, dforce, 3, force);
  to_host(L, dout, 1, phi_out);
  return rc;
}

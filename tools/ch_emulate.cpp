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

static dim3 blockDim, gridDim;
static inline void __syncthreads() {}
using std::max;
#undef __shared__
#define __shared__ static

#define B200LB_KERNEL_SOURCE_ONLY
#include "../ludwig_b200/csrc/le_field_kernels.cuh"
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
  P.skipz = (nlocal[2] == 1) ? 1 : 0;
  int rc = -1;
  if (method == 0 || method == CH_FORCE_NONE) rc = run_order<0>(P, order);
  if (method == CH_FORCE_STRESS_DIVERGENCE) rc = run_order<CH_FORCE_STRESS_DIVERGENCE>(P, order);
  if (method == CH_FORCE_PHI_GRADMU) rc = run_order<CH_FORCE_PHI_GRADMU>(P, order);
  to_host(L, dforce, 3, force);
  to_host(L, dout, 1, phi_out);
  return rc;
}


// ---------------------------------------------------------------------------
// The same with Lees-Edwards planes: what ch_run + ch_le_run (api.cu) launch, in their
// order. phi with its periodic halo; grad, delsq: the plane-aware gradient of the
// interior (what grad*_kernel + le_grad_fix leave); u: the last collision's velocity.
// loc[p]: 1-based x of the layer below plane p. t = physics_control_time.
// The plane reduction (ch_le_reduce: shared memory + barriers) is restated as a loop.
// ---------------------------------------------------------------------------

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

template <int ORDER>
static void le_passes(const ChParams & P, const ChLeParams & Q, int do_force) {
  const Layout & L = P.L;
  dim3 block(128, 1, 1);
  dim3 grid((L.n[2] + 127)/128, L.n[1], Q.nplanes);
  for_grid(grid, block, [&]() { ch_le_pass_a<CH_EXACT, ORDER>(P, Q, do_force); });
  if (do_force) {
    for (int p = 0; p < Q.nplanes; p++) for (int a = 0; a < 3; a++) {
      const double * scr = Q.scr + (i64) 8*p*L.s0;
      double part[256];
      const int n = L.n[1]*L.n[2];
      for (int tid = 0; tid < 256; tid++) {
	double acc = 0.0;
	for (int i = tid; i < n; i += 256) {
	  const int jc = 1 + i/L.n[2], kc = 1 + i % L.n[2];
	  const i64 col = (i64) jc*L.s1 + L.z0 + kc;
	  acc += - scr[(2 + a)*L.s0 + col] + scr[(5 + a)*L.s0 + col];
	}
	part[tid] = acc;
      }
      for (int w = 128; w > 0; w >>= 1) for (int tid = 0; tid < w; tid++) part[tid] += part[tid + w];
      Q.fbar[3*p + a] = part[0];
    }
  }
  for_grid(grid, block, [&]() { ch_le_pass_b<CH_EXACT, ORDER>(P, Q, do_force); });
}

extern "C" int ch_le_emulate(const int nlocal[3], int nplanes, const int * loc, double uy,
			     int oscillatory, int period, int nt0, double t, int npt, int order,
			     const double prm[4], const double gmu[3],
			     const double * phi, const double * grad, const double * delsq,
			     const double * u, double * force, double * phi_out) {
  Layout L = make_layout(nlocal, 16);
  std::vector<double> dphi, dgrad, ddelsq, du, dforce, dout;
  to_device(L, phi, 1, dphi); to_device(L, grad, 3, dgrad); to_device(L, delsq, 1, ddelsq);
  to_device(L, u, 3, du); to_device(L, force, 3, dforce); to_device(L, phi, 1, dout);
  ChParams P;
  memset(&P, 0, sizeof(P));
  P.L = L;
  for (int d = 0; d < 3; d++) { P.wrap[d] = 1; P.gmu[d] = gmu[d]; P.fconst[d] = 0.0; }
  P.phi = dphi.data(); P.grad = dgrad.data(); P.delsq = ddelsq.data(); P.u = du.data();
  P.force = dforce.data(); P.phi_out = dout.data();
  P.A = prm[0]; P.B = prm[1]; P.kappa = prm[2]; P.mobility = prm[3];
  P.accumulate = 0; P.order = order; P.method = CH_FORCE_STRESS_DIVERGENCE; P.nx = 3;
  P.skipz = (nlocal[2] == 1) ? 1 : 0;
  if (run_order<CH_FORCE_STRESS_DIVERGENCE>(P, order)) return -1;

  const int np = nplanes;
  std::vector<double> buf((size_t) 22*np*L.s0 + 3*np, 0.0);
  double * bphi = buf.data();
  double * bgrad = bphi + (size_t) 4*np*L.s0;
  double * bdelsq = bgrad + (size_t) 6*np*L.s0;
  double * bux = bdelsq + (size_t) 2*np*L.s0;
  double * scr = bux + (size_t) 2*np*L.s0;
  double * fbar = scr + (size_t) 8*np*L.s0;
  LePhiTab T;
  memset(&T, 0, sizeof(T));
  T.L = L; T.w0 = T.w1 = T.w2 = 1;
  dim3 block(128, 1, 1);
  le_phi_tab_fill4(T, np, loc, uy, oscillatory, period, nt0, t);
  for_grid(dim3((L.n[2] + 2 + 127)/128, L.n[1] + 2, T.nbuf), block, [&]() { le_phi_buffer_kernel(T, dphi.data(), bphi); });
  LeBufGradTab G;
  memset(&G, 0, sizeof(G));
  G.L = L; G.w1 = 1; G.w2 = 1; G.nplanes = np;
  for (int p = 0; p < np; p++) G.loc[p] = loc[p];
  dim3 gplane((L.n[2] + 127)/128, L.n[1], 2*np);
  if (npt == 27) for_grid(gplane, block, [&]() { le_buffer_grad_kernel<27>(G, dphi.data(), bphi, bgrad, bdelsq); });
  else for_grid(gplane, block, [&]() { le_buffer_grad_kernel<7>(G, dphi.data(), bphi, bgrad, bdelsq); });
  le_u_tab_fill(T, np, loc, uy, oscillatory, period, nt0, t);
  for_grid(gplane, block, [&]() { le_u_buffer_kernel(T, du.data(), bux); });
  ChLeParams Q;
  memset(&Q, 0, sizeof(Q));
  Q.nplanes = np;
  for (int p = 0; p < np; p++) Q.loc[p] = loc[p];
  Q.bphi = bphi; Q.bgrad = bgrad; Q.bdelsq = bdelsq; Q.bux = bux; Q.scr = scr; Q.fbar = fbar;
  {
    const int ny = L.n[1];
    double dy = fmod(+(t*uy), 1.0*ny);
    Q.jdy_up = (int) floor(dy); Q.fr_up = dy - Q.jdy_up;
    dy = fmod(-(t*uy), 1.0*ny);
    Q.jdy_dn = (int) floor(dy); Q.fr_dn = dy - Q.jdy_dn;
    Q.ra = 0.5/(1.0*nlocal[1]*nlocal[2]);
  }
  switch (order) {
  case 1: le_passes<1>(P, Q, 1); break;
  case 2: le_passes<2>(P, Q, 1); break;
  case 3: le_passes<3>(P, Q, 1); break;
  case 4: le_passes<4>(P, Q, 1); break;
  default: return -1;
  }
  to_host(L, dforce, 3, force);
  to_host(L, dout, 1, phi_out);
  return 0;
}

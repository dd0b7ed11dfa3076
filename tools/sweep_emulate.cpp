// tools/sweep_emulate.cpp -- DEVELOPMENT TOOL, not part of the product and not a fallback.
//
// Steps through the SOURCE of phi_from_g_kernel and lb_sweep_kernel
// (ludwig_b200/csrc/aux_kernels.cuh, site_kernels.cuh) on the host, thread by thread over
// the grid the launchers use, to check in a container without a GPU that the fused binary
// sweep that takes the first moment of g from phi_from_g (GM bit 3, option g_moment)
// writes exactly what the sweep that pulls g itself writes. Nothing in ludwig_b200/
// links this.
//
//   g++ -O1 -ffp-contract=off -w -std=c++17 -I/usr/local/cuda/include \
//       -o tools/_sweep_emulate tools/sweep_emulate.cpp && tools/_sweep_emulate

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include <cuda_runtime.h>

static uint3 blockIdx, threadIdx;
static dim3 blockDim, gridDim;
template <class T> static inline T __ldg(const T * p) { return *p; }
static inline double __dmul_rn(double a, double b) { return a*b; }
static inline double __dadd_rn(double a, double b) { return a + b; }
static inline double __fma_rn(double a, double b, double c) { return std::fma(a, b, c); }
static inline int __double2loint(double d) { long long b; memcpy(&b, &d, 8); return (int) (b & 0xffffffffll); }
static inline int __double2hiint(double d) { long long b; memcpy(&b, &d, 8); return (int) ((b >> 32) & 0xffffffffll); }
static inline double __hiloint2double(int hi, int lo) {
  unsigned long long b = ((unsigned long long) (unsigned int) hi << 32) | (unsigned int) lo;
  double d; memcpy(&d, &b, 8); return d;
}
// the in-kernel diagnostics (DIAG) are not stepped through here: one thread at a time
static inline double __shfl_xor_sync(unsigned, double v, int) { return v; }
static inline double __shfl_down_sync(unsigned, double v, int) { return v; }
static inline void __syncthreads() {}
using std::min;
using std::max;
#ifndef __launch_bounds__
#define __launch_bounds__(...)
#endif
#undef __shared__
#define __shared__ static
#undef __constant__
#define __constant__

#include "../ludwig_b200/csrc/layout.cuh"
#include "../ludwig_b200/csrc/models.cuh"
#include "../ludwig_b200/csrc/model_tables.h"

static GhostTab c_ghost;

#include "../ludwig_b200/csrc/site_kernels.cuh"

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

static unsigned long long rng_state = 88172645463325252ull;
static double rnd() {
  rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17;
  return (double) (rng_state >> 11)/9007199254740992.0;
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

template <class M, int GM>
static int run_case(const int nlocal[3], const int wrap[3], bool lazy_grad) {
  constexpr int Q = M::NVEL;
  const Layout L = make_layout(nlocal, 16);
  const size_t nsd = (size_t) L.nsd;

  std::vector<double> src(2*Q*nsd), phi(nsd), jphi(3*nsd, -777.0), grad(3*nsd), delsq(nsd), force(3*nsd);
  std::vector<double> dst0(2*Q*nsd, -1.0), dst1(2*Q*nsd, -1.0), u0(3*nsd, -2.0), u1(3*nsd, -2.0);
  std::vector<double> noise(2*nsd), noise0, noise1;
  for (size_t k = 0; k < src.size(); k++) {
    const int p = (int) ((k/nsd) % Q);
    src[k] = (k/nsd < (size_t) Q ? 1.0 : 0.1)*M::wv(p)*(1.0 + 0.2*(rnd() - 0.5));
  }
  for (auto & x : grad) x = 0.05*(rnd() - 0.5);
  for (auto & x : delsq) x = 0.05*(rnd() - 0.5);
  for (auto & x : force) x = 1.0e-3*(rnd() - 0.5);
  for (auto & x : noise) { unsigned long long b = (unsigned long long) (rnd()*1.8e19) | 1ull; memcpy(&x, &b, 8); }

  build_ghost_tab<M>(c_ghost);

  // phi_impl (model_impl.cuh)
  {
    dim3 block(128, 2, 1);
    dim3 grid((L.n[2] + 127)/128, (L.n[1] + 1)/2, L.n[0]);
    for_grid(grid, block, [&]() {
      phi_from_g_kernel<M, true, true>(L, wrap[0], wrap[1], wrap[2], src.data() + Q*nsd, phi.data(), jphi.data());
    });
  }
  // the stencil of the gradient that is not materialised reads the halo of phi
  for (size_t k = 0; k < nsd; k++) if (phi[k] == 0.0) phi[k] = 0.2*(rnd() - 0.5);

  KParams P;
  memset(&P, 0, sizeof(P));
  P.L = L;
  P.src = src.data();
  P.force = force.data();
  P.phi = phi.data();
  if (!lazy_grad) { P.grad = grad.data(); P.delsq = delsq.data(); }
  for (int d = 0; d < 3; d++) { P.lo[d] = 0; P.hi[d] = L.n[d]; P.wrap[d] = wrap[d]; }
  P.rtau_shear = 1.0/(0.5 + 3.0*0.1); P.rtau_bulk = 1.0/(0.5 + 3.0*0.07);
  P.fg[0] = 1.0e-4; P.fg[1] = -2.0e-4; P.fg[2] = 3.0e-4;
  P.rtau2 = 2.0/(1.0 + 6.0*0.15); P.A = -0.0625; P.B = 0.0625; P.kappa = 0.04;
  for (int k = 0; k < B200LB_MAXGHOST; k++) P.gkeep[k] = 0.3 + 0.02*k;
  P.nv_diag = 1.1e-3; P.nv_off = 0.8e-3; P.nv_bulk = 0.9e-3;
  for (int k = 0; k < B200LB_MAXNOISE; k++) P.nv_ghost[k] = 1.0e-3*(1 + k);
  P.nghost_draw = 1;
  P.nlv_big = sqrt(2.0 + sqrt(2.0)); P.nlv_small = sqrt(2.0 - sqrt(2.0));

  // launch_one (model_impl.cuh), binary: 128 threads per block, one site per thread
  const int bx = 64, by = 2;
  dim3 block(bx, by, 1);
  dim3 grid((L.n[2] + bx - 1)/bx, (L.n[1] + by - 1)/by, L.n[0]);

  noise0 = noise; noise1 = noise;
  P.dst = dst0.data(); P.u = u0.data(); P.noise = noise0.data(); P.jphi = nullptr;
  for_grid(grid, block, [&]() { lb_sweep_kernel<M, 1, true, true, true, GM, false, false>(P); });
  P.dst = dst1.data(); P.u = u1.data(); P.noise = noise1.data(); P.jphi = jphi.data();
  for_grid(grid, block, [&]() { lb_sweep_kernel<M, 1, true, true, true, GM | 8, false, false>(P); });

  size_t ndiff = 0, nwritten = 0;
  for (size_t k = 0; k < dst0.size(); k++) {
    if (memcmp(&dst0[k], &dst1[k], 8) != 0) ndiff++;
    if (dst0[k] != -1.0) nwritten++;
  }
  for (size_t k = 0; k < u0.size(); k++) if (memcmp(&u0[k], &u1[k], 8) != 0) ndiff++;
  for (size_t k = 0; k < noise0.size(); k++) if (memcmp(&noise0[k], &noise1[k], 8) != 0) ndiff++;
  const size_t nint = (size_t) L.n[0]*L.n[1]*L.n[2];
  const bool ok = (ndiff == 0 && nwritten == 2*Q*nint);
  printf("%s Q=%d GM=%d local %dx%dx%d wrap %d%d%d %s: %zu values differ, %zu of %zu written\n",
	 ok ? "ok  " : "FAIL", Q, GM, nlocal[0], nlocal[1], nlocal[2], wrap[0], wrap[1], wrap[2],
	 lazy_grad ? "stencil gradient" : "stored gradient", ndiff, nwritten, 2*Q*nint);
  return ok ? 0 : 1;
}

// The propagation of the binary fluid with flat no-slip walls folded in (WALL = true,
// BINARY = true, COLLIDE = false: what materialise() launches for the slabs next to walls,
// here over the whole box): f'(r, p) = f(r - c_p, p), or, where r - c_p lies inside a wall,
// the population that left r the other way, f(r, Q - p) (wall_bbl, src/wall.c:997-1107,
// both distributions, walls at rest). Pure data movement: identical to the bits.
template <class M>
static int run_wall_case(const int nlocal[3], const int walls[3]) {
  constexpr int Q = M::NVEL;
  const Layout L = make_layout(nlocal, 16);
  const size_t nsd = (size_t) L.nsd;
  std::vector<double> src(2*Q*nsd), dst(2*Q*nsd, -1.0);
  for (auto & x : src) x = rnd();
  KParams P;
  memset(&P, 0, sizeof(P));
  P.L = L;
  P.src = src.data(); P.dst = dst.data();
  for (int d = 0; d < 3; d++) {
    P.lo[d] = 0; P.hi[d] = L.n[d];
    P.wrap[d] = walls[d] ? 0 : 1;
    P.wall_lo[d] = P.wall_hi[d] = walls[d];
  }
  const int bx = 64, by = 2;
  dim3 block(bx, by, 1);
  dim3 grid((L.n[2] + bx - 1)/bx, (L.n[1] + by - 1)/by, L.n[0]);
  for_grid(grid, block, [&]() { lb_sweep_kernel<M, 1, true, false, true, 0, false, true>(P); });
  size_t ndiff = 0;
  for (int n = 0; n < 2; n++) for (int p = 0; p < Q; p++)
    for (int i0 = 1; i0 <= L.n[0]; i0++) for (int i1 = 1; i1 <= L.n[1]; i1++) for (int i2 = 1; i2 <= L.n[2]; i2++) {
      int c[3] = {i0, i1, i2}, s[3];
      bool inwall = false;
      for (int d = 0; d < 3; d++) {
	s[d] = c[d] - dcv<M>(p, d);
	if (walls[d]) { if (s[d] < 1 || s[d] > L.n[d]) inwall = true; }
	else { if (s[d] < 1) s[d] += L.n[d]; if (s[d] > L.n[d]) s[d] -= L.n[d]; }
      }
      const i64 here = (i64) c[0]*L.s0 + (i64) c[1]*L.s1 + L.z0 + c[2];
      const i64 from = (i64) s[0]*L.s0 + (i64) s[1]*L.s1 + L.z0 + s[2];
      const double want = (inwall && p > 0) ? src[(size_t) (n*Q + (Q - p))*nsd + here] : src[(size_t) (n*Q + p)*nsd + from];
      if (memcmp(&want, &dst[(size_t) (n*Q + p)*nsd + here], 8) != 0) ndiff++;
    }
  const bool ok = (ndiff == 0);
  printf("%s Q=%d propagation with walls %d%d%d, binary, local %dx%dx%d: %zu values differ\n",
	 ok ? "ok  " : "FAIL", Q, walls[0], walls[1], walls[2], nlocal[0], nlocal[1], nlocal[2], ndiff);
  return ok ? 0 : 1;
}

int main(void) {
  int bad = 0;
  {
    const int wsz[2][3] = {{6, 5, 7}, {4, 4, 70}};
    const int ww[4][3] = {{1, 1, 1}, {0, 1, 1}, {1, 0, 0}, {0, 0, 1}};
    for (int s = 0; s < 2; s++) for (int w = 0; w < 4; w++) {
      bad += run_wall_case<D3Q19>(wsz[s], ww[w]);
      bad += run_wall_case<D3Q27>(wsz[s], ww[w]);
    }
  }
  const int sizes[3][3] = {{6, 5, 7}, {4, 4, 70}, {3, 9, 16}};
  const int wraps[3][3] = {{1, 1, 1}, {0, 0, 0}, {1, 0, 1}};
  for (int s = 0; s < 3; s++) for (int w = 0; w < 3; w++) {
    bad += run_case<D3Q19, 0>(sizes[s], wraps[w], false);
    bad += run_case<D3Q19, 1>(sizes[s], wraps[w], true);
    bad += run_case<D3Q19, 2>(sizes[s], wraps[w], false);
    bad += run_case<D3Q19, 3>(sizes[s], wraps[w], false);
    bad += run_case<D3Q27, 0>(sizes[s], wraps[w], true);
    bad += run_case<D3Q27, 1>(sizes[s], wraps[w], false);
  }
  printf(bad ? "%d case(s) FAILED\n" : "all cases identical (%d failures)\n", bad);
  return bad ? 1 : 0;
}

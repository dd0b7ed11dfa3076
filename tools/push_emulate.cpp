// tools/push_emulate.cpp -- DEVELOPMENT TOOL, not part of the product and not a fallback.
//
// Steps through the SOURCE of lb_sweep_kernel's peer-memory variants (PUSH = true,
// ludwig_b200/csrc/site_kernels.cuh) on the host, thread by thread over the grid the
// launchers use, with the lattice as its own neighbour in every decomposed direction
// (what option "loopback" + b200lb_p2p_import(nranks = 1) sets up), and checks that three
// steps -- a collision in place, then two fused stream-and-collide sweeps -- leave exactly
// what the same sweeps leave when every direction wraps within the rank. Run for the
// row-end halo words (8-byte and whole-sector stores) and for the packed z-face planes
// (KParams::zoff / zread). Nothing in ludwig_b200/ links this.
//
//   g++ -O1 -ffp-contract=off -w -std=c++17 -I/usr/local/cuda/include \
//       -o tools/_push_emulate tools/push_emulate.cpp && tools/_push_emulate

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
#include "../ludwig_b200/csrc/zpack_kernels.cuh"

// create_impl (api.cu)
static Layout make_layout(const int nlocal[3], int zalign, bool twod) {
  Layout L;
  for (int d = 0; d < 3; d++) { L.n[d] = nlocal[d]; L.h[d] = (twod && d == 0) ? 0 : 1; L.na[d] = nlocal[d] + 2*L.h[d]; }
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

// launch_one (model_impl.cuh)
template <int VEC> static void geometry(const Layout & L, dim3 & grid, dim3 & block) {
  const int nt2 = (L.n[2] + VEC - 1)/VEC;
  int bx = 128;
  while (bx > 1 && bx/2 >= nt2) bx /= 2;
  int by = 256/bx;
  while (by > 1 && by/2 >= L.n[1]) by /= 2;
  block = dim3(bx, by, 1);
  grid = dim3((nt2 + bx - 1)/bx, (L.n[1] + by - 1)/by, L.n[0]);
}

// zpack_geom (api.cu)
static void zgeom(const Layout & L, int arr, KParams & P) {
  const i64 na0 = L.na[0], na1 = L.na[1];
  const i64 w = (arr == 2) ? 4 : 1;
  P.zvec = (arr == 2) ? 1 : 0;
  P.zsy = w; P.zsx = w*na1;
  P.zsp = w*na0*na1;
}

// mode 0: 8-byte stores into the row-end halo words; 1: whole-sector stores where the layout
// allows; 2, 3: packed planes in arrangement mode - 1 (option zpack)
template <class M, int VEC, int GM>
static int run_case(const int nlocal[3], const int dec[3], int mode, int zalign) {
  constexpr int Q = M::NVEL;
  const bool twod = (M::NDIM == 2);
  const Layout L = make_layout(nlocal, zalign, twod);
  const size_t nsd = (size_t) L.nsd;
  const size_t zpack = (size_t) 2*9*L.na[0]*L.na[1]*4 + 16;
  const size_t nlat = Q*nsd + zpack;

  std::vector<double> init(nlat, 0.0), force(3*nsd);
  for (size_t k = 0; k < Q*nsd; k++) {
    const int p = (int) (k/nsd);
    init[k] = M::wv(p)*(1.0 + 0.2*(rnd() - 0.5));
  }
  for (auto & x : force) x = 1.0e-3*(rnd() - 0.5);
  build_ghost_tab<M>(c_ghost);

  KParams P0;
  memset(&P0, 0, sizeof(P0));
  P0.L = L;
  P0.force = force.data();
  for (int d = 0; d < 3; d++) { P0.lo[d] = 0; P0.hi[d] = L.n[d]; }
  P0.rtau_shear = 1.0/(0.5 + 3.0*0.1); P0.rtau_bulk = 1.0/(0.5 + 3.0*0.07);
  P0.fg[0] = 1.0e-4; P0.fg[1] = -2.0e-4; P0.fg[2] = 3.0e-4;
  for (int k = 0; k < B200LB_MAXGHOST; k++) P0.gkeep[k] = 0.3 + 0.02*k;

  dim3 grid, block;
  geometry<VEC>(L, grid, block);

  // reference: every direction wraps within the rank
  std::vector<double> ra(init), rb(nlat, -1.0);
  {
    KParams P = P0;
    for (int d = 0; d < 3; d++) P.wrap[d] = (L.h[d] == 1) ? 1 : 0;
    P.src = ra.data(); P.dst = ra.data();
    for_grid(grid, block, [&]() { lb_sweep_kernel<M, VEC, false, true, false, GM, false, false, false, 0>(P); });
    P.src = ra.data(); P.dst = rb.data();
    for_grid(grid, block, [&]() { lb_sweep_kernel<M, VEC, true, true, false, GM, false, false, false, 0>(P); });
    P.src = rb.data(); P.dst = ra.data();
    for_grid(grid, block, [&]() { lb_sweep_kernel<M, VEC, true, true, false, GM, false, false, false, 0>(P); });
  }

  // peer-memory sweeps, the lattice as its own neighbour in the decomposed directions
  std::vector<double> a(init), b(nlat, -1.0);
  {
    KParams P = P0;
    for (int d = 0; d < 3; d++) P.wrap[d] = (L.h[d] == 1 && !dec[d]) ? 1 : 0;
    P.push = 1;
    if (mode == 1) {
      const i64 lo = L.z0 + L.h[2] - 1, hi = L.z0 + L.h[2] + L.n[2];
      P.zsec = (L.h[2] == 1 && lo >= 3 && (lo + 1) % 4 == 0 && hi % 4 == 0 && hi + 3 < L.s1
		&& L.s1 % 4 == 0 && L.s0 % 4 == 0 && L.nsd % 4 == 0) ? 1 : 0;
    }
    const bool zuse = (mode >= 2) && dec[2];
    if (zuse) { P.zoff = (i64) Q*L.nsd; P.zwrite = 1; zgeom(L, mode - 1, P); }
    auto peers = [&](double * lat) {
      for (int k = 0; k < 27; k++) {
	const int dir[3] = {k/9 - 1, (k/3)%3 - 1, k%3 - 1};
	bool ok = (k != 13);
	for (int d = 0; d < 3; d++) if (dir[d] != 0 && !(dec[d] && L.h[d] == 1)) ok = false;
	P.peer[k] = ok ? lat : nullptr;
      }
    };
    P.src = a.data(); P.dst = a.data(); peers(a.data()); P.zread = 0;
    for_grid(grid, block, [&]() { lb_sweep_kernel<M, VEC, false, true, false, GM, false, false, true, 0>(P); });
    P.src = a.data(); P.dst = b.data(); peers(b.data()); P.zread = zuse ? 1 : 0;
    for_grid(grid, block, [&]() { lb_sweep_kernel<M, VEC, true, true, false, GM, false, false, true, 0>(P); });
    P.src = b.data(); P.dst = a.data(); peers(a.data()); P.zread = zuse ? 1 : 0;
    for_grid(grid, block, [&]() { lb_sweep_kernel<M, VEC, true, true, false, GM, false, false, true, 0>(P); });
  }

  size_t ndiff = 0, nint = 0;
  for (int p = 0; p < Q; p++)
    for (int i0 = 0; i0 < L.n[0]; i0++) for (int i1 = 0; i1 < L.n[1]; i1++) for (int i2 = 0; i2 < L.n[2]; i2++) {
      const size_t at = (size_t) p*nsd + (size_t) (i0 + L.h[0])*L.s0 + (size_t) (i1 + L.h[1])*L.s1 + L.z0 + L.h[2] + i2;
      nint++;
      if (memcmp(&a[at], &ra[at], 8) != 0) ndiff++;
    }
  const bool ok = (ndiff == 0);
  printf("%s Q=%d VEC=%d GM=%d local %dx%dx%d zalign %d decomposed %d%d%d %s: %zu of %zu values differ\n",
	 ok ? "ok  " : "FAIL", Q, VEC, GM, nlocal[0], nlocal[1], nlocal[2], zalign, dec[0], dec[1], dec[2],
	 mode == 0 ? "halo words" : (mode == 1 ? "halo sectors" : (mode == 2 ? "packed planes 1" : "packed planes 2")), ndiff, nint);
  return ok ? 0 : 1;
}

// A collision in place that pushes into the packed planes, zhalo_unpack (api.cu) and then a
// propagation on its own, which reads the row-end halo words: against the same with the
// halo words pushed directly.
template <class M, int VEC>
static int run_unpack_case(const int nlocal[3], const int dec[3], int arr) {
  constexpr int Q = M::NVEL;
  const Layout L = make_layout(nlocal, 16, M::NDIM == 2);
  const size_t nsd = (size_t) L.nsd;
  const size_t nlat = Q*nsd + (size_t) 2*9*L.na[0]*L.na[1]*4 + 16;
  std::vector<double> init(nlat, 0.0), force(3*nsd, 0.0);
  for (size_t k = 0; k < Q*nsd; k++) init[k] = M::wv((int) (k/nsd))*(1.0 + 0.2*(rnd() - 0.5));
  build_ghost_tab<M>(c_ghost);
  KParams P0;
  memset(&P0, 0, sizeof(P0));
  P0.L = L;
  for (int d = 0; d < 3; d++) { P0.lo[d] = 0; P0.hi[d] = L.n[d]; P0.wrap[d] = (L.h[d] == 1 && !dec[d]) ? 1 : 0; }
  P0.rtau_shear = 1.0/(0.5 + 3.0*0.1); P0.rtau_bulk = 1.0/(0.5 + 3.0*0.07);
  dim3 grid, block;
  geometry<VEC>(L, grid, block);
  std::vector<double> out[2];
  for (int packed = 0; packed < 2; packed++) {
    std::vector<double> a(init), b(nlat, -1.0);
    KParams P = P0;
    P.push = 1;
    if (packed) { P.zoff = (i64) Q*L.nsd; P.zwrite = 1; zgeom(L, arr, P); }
    for (int k = 0; k < 27; k++) {
      const int dir[3] = {k/9 - 1, (k/3)%3 - 1, k%3 - 1};
      bool ok = (k != 13);
      for (int d = 0; d < 3; d++) if (dir[d] != 0 && !(dec[d] && L.h[d] == 1)) ok = false;
      P.peer[k] = ok ? a.data() : nullptr;
    }
    P.src = a.data(); P.dst = a.data();
    for_grid(grid, block, [&]() { lb_sweep_kernel<M, VEC, false, true, false, 0, false, false, true, 0>(P); });
    if (packed) {
      ZPackTab T;
      memset(&T, 0, sizeof(T));
      int n[2] = {0, 0};
      for (int p = 0; p < Q; p++) {
	if (dcv<M>(p, 2) ==  1) T.pidx[0][n[0]++] = p;
	if (dcv<M>(p, 2) == -1) T.pidx[1][n[1]++] = p;
      }
      T.np = n[0];
      const int x0 = (dec[0] && L.h[0]) ? 0 : L.h[0], x1 = L.h[0] + L.n[0] + ((dec[0] && L.h[0]) ? L.h[0] : 0);
      const int y0 = dec[1] ? 0 : L.h[1], y1 = L.h[1] + L.n[1] + (dec[1] ? L.h[1] : 0);
      dim3 ub(256, 1, 1), ug((y1 - y0 + 255)/256, x1 - x0, 2*T.np);
      for_grid(ug, ub, [&]() { zhalo_unpack_kernel(L, T, a.data(), (i64) Q*L.nsd, P.zsx, P.zsy, P.zsp, x0, x1, y0, y1); });
    }
    KParams R = P0;
    R.src = a.data(); R.dst = b.data();
    for_grid(grid, block, [&]() { lb_sweep_kernel<M, VEC, true, false, false, 0, false, false, false, 0>(R); });
    out[packed].assign(b.begin(), b.begin() + Q*nsd);
  }
  size_t ndiff = 0;
  for (size_t k = 0; k < out[0].size(); k++) if (memcmp(&out[0][k], &out[1][k], 8) != 0) ndiff++;
  printf("%s Q=%d VEC=%d local %dx%dx%d decomposed %d%d%d arrangement %d unpack + propagation: %zu values differ\n",
	 ndiff == 0 ? "ok  " : "FAIL", Q, VEC, nlocal[0], nlocal[1], nlocal[2], dec[0], dec[1], dec[2], arr, ndiff);
  return ndiff == 0 ? 0 : 1;
}

int main(void) {
  int bad = 0;
  const int sizes[4][3] = {{6, 5, 8}, {4, 4, 70}, {3, 9, 16}, {2, 2, 2}};
  const int decs[5][3] = {{1, 1, 1}, {0, 0, 1}, {1, 0, 1}, {0, 1, 1}, {1, 1, 0}};
  for (int s = 0; s < 4; s++) for (int w = 0; w < 5; w++) for (int mode = 0; mode < 4; mode++) {
    bad += run_case<D3Q19, 2, 0>(sizes[s], decs[w], mode, 16);
    bad += run_case<D3Q19, 1, 1>(sizes[s], decs[w], mode, 16);
    bad += run_case<D3Q27, 1, 0>(sizes[s], decs[w], mode, 16);
  }
  {
    // two-dimensional: device axes (z, x, y), the unit axis first and without halo
    const int s2[2][3] = {{1, 6, 8}, {1, 5, 33}};
    const int d2[3][3] = {{0, 1, 1}, {0, 0, 1}, {0, 1, 0}};
    for (int s = 0; s < 2; s++) for (int w = 0; w < 3; w++) for (int mode = 0; mode < 4; mode++)
      bad += run_case<D2Q9, 1, 0>(s2[s], d2[w], mode, 16);
  }
  for (int s = 0; s < 4; s++) for (int w = 0; w < 4; w++) for (int arr = 1; arr <= 2; arr++) {
    bad += run_unpack_case<D3Q19, 2>(sizes[s], decs[w], arr);
    bad += run_unpack_case<D3Q27, 1>(sizes[s], decs[w], arr);
  }
  for (int arr = 1; arr <= 2; arr++) {
    const int s2[3] = {1, 6, 8}, d2[3] = {0, 1, 1};
    bad += run_unpack_case<D2Q9, 1>(s2, d2, arr);
  }
  printf(bad ? "%d case(s) FAILED\n" : "all cases identical (%d failures)\n", bad);
  return bad ? 1 : 0;
}

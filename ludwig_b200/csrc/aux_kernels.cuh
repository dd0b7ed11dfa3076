// aux_kernels.cuh -- model-dependent helpers: order parameter and diagnostics.

#pragma once

#include "layout.cuh"
#include "models.cuh"

// ---------------------------------------------------------------------------
// First moment sum_p c_p g_p of one site's order-parameter populations, physical
// components, summed over the pairs (p, Q - p). The binary collision needs nothing
// else of the incoming g (src/collision.c:956-1023 rebuilds g from phi, this flux,
// u and mu), so whoever holds the pulled g may hand the flux on instead.
// ---------------------------------------------------------------------------

template <class M>
__device__ __forceinline__ void
g_first_moment(const double (&g)[M::NVEL], double (&j)[M::NDIM]) {
  constexpr int Q = M::NVEL;
  #pragma unroll
  for (int a = 0; a < M::NDIM; a++) j[a] = 0.0;
  #pragma unroll
  for (int p = 1; p <= (Q - 1)/2; p++) {
    const double d = g[p] - g[Q - p];
    #pragma unroll
    for (int a = 0; a < M::NDIM; a++) {
      if (M::cv(p, a) ==  1) j[a] += d;
      if (M::cv(p, a) == -1) j[a] -= d;
    }
  }
}

// ---------------------------------------------------------------------------
// phi = sum_p g_p, p in index order (src/phi_lb_coupler.c:72-102), interior.
// PULL: the lattice holds post-collision values whose propagation is still
// pending, so the sum runs over g_p(r - c_p): identical to propagating first.
// MOM: also store the first moment of the same (pulled) populations, jphi[NDIM][nsd].
// The next fused stream-and-collide sweep then reads phi and these three values
// instead of pulling the Q populations of g a second time.
// ---------------------------------------------------------------------------

template <class M, bool PULL, bool MOM>
__global__ void __launch_bounds__(256)
phi_from_g_kernel(const Layout L, const int w0, const int w1, const int w2,
		  const double * __restrict__ g, double * __restrict__ phi,
		  double * __restrict__ jphi) {

  const int i2 = blockIdx.x*blockDim.x + threadIdx.x;
  const int i1 = blockIdx.y*blockDim.y + threadIdx.y;
  const int i0 = blockIdx.z;
  if (i2 >= L.n[2] || i1 >= L.n[1]) return;
  const i64 site = (i64)(i0 + L.h[0])*L.s0 + (i64)(i1 + L.h[1])*L.s1 + L.z0 + L.h[2] + i2;

  double v[M::NVEL];
  #pragma unroll
  for (int p = 0; p < M::NVEL; p++) {
    i64 s = site;
    if (PULL) {
      // r - c_p, wrapping along axes that are periodic within this rank
      if (dcv<M>(p, 0) ==  1) s += (w0 && i0 == 0) ? (i64) (L.n[0] - 1)*L.s0 : -L.s0;
      if (dcv<M>(p, 0) == -1) s += (w0 && i0 == L.n[0] - 1) ? -(i64) (L.n[0] - 1)*L.s0 : L.s0;
      if (dcv<M>(p, 1) ==  1) s += (w1 && i1 == 0) ? (i64) (L.n[1] - 1)*L.s1 : -L.s1;
      if (dcv<M>(p, 1) == -1) s += (w1 && i1 == L.n[1] - 1) ? -(i64) (L.n[1] - 1)*L.s1 : L.s1;
      if (dcv<M>(p, 2) ==  1) s += (w2 && i2 == 0) ? L.n[2] - 1 : -1;
      if (dcv<M>(p, 2) == -1) s += (w2 && i2 == L.n[2] - 1) ? -(L.n[2] - 1) : 1;
    }
    v[p] = __ldg(g + (i64) p*L.nsd + s);
  }
  double sum = 0.0;
  #pragma unroll
  for (int p = 0; p < M::NVEL; p++) sum += v[p];
  phi[site] = sum;
  if (MOM) {
    double j[M::NDIM];
    g_first_moment<M>(v, j);
    #pragma unroll
    for (int a = 0; a < M::NDIM; a++) jphi[a*L.nsd + site] = j[a];
  }
}

// ---------------------------------------------------------------------------
// Conserved-quantity diagnostics over the interior (restates the sums of
// stats_distribution_print / distribution_gm_kernel,
// src/stats_distribution.c:55-118, :281-362): warp-shuffle reduction, one
// partial per block, then one block folds the partials in a fixed order
// (deterministic; no atomics).
// out: sum rho, sum rho^2, min rho, max rho, g[3], sum phi
// ---------------------------------------------------------------------------

__device__ __forceinline__ double warp_sum(double v) {
  #pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_min(double v) {
  #pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
  #pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ void block_reduce8(double (&q)[8], double * out) {
  __shared__ double sm[8][8];
  const int tid = threadIdx.y*blockDim.x + threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const int nwarp = (blockDim.x*blockDim.y + 31) >> 5;
  #pragma unroll
  for (int k = 0; k < 8; k++) {
    double r = (k == 2) ? warp_min(q[k]) : ((k == 3) ? warp_max(q[k]) : warp_sum(q[k]));
    if (lane == 0) sm[k][warp] = r;
  }
  __syncthreads();
  if (tid < 8) {
    const int k = tid;
    double r = sm[k][0];
    for (int w = 1; w < nwarp; w++) {
      r = (k == 2) ? fmin(r, sm[k][w]) : ((k == 3) ? fmax(r, sm[k][w]) : r + sm[k][w]);
    }
    out[k] = r;
  }
}

template <class M, int NDIST>
__global__ void __launch_bounds__(256)
stats_kernel(const Layout L, const double * __restrict__ f, double * __restrict__ partial) {

  const int i2 = blockIdx.x*blockDim.x + threadIdx.x;
  const int i1 = blockIdx.y*blockDim.y + threadIdx.y;
  const int i0 = blockIdx.z;
  double q[8] = {0.0, 0.0, 1.0e300, -1.0e300, 0.0, 0.0, 0.0, 0.0};

  if (i2 < L.n[2] && i1 < L.n[1]) {
    const i64 site = (i64)(i0 + L.h[0])*L.s0 + (i64)(i1 + L.h[1])*L.s1 + L.z0 + L.h[2] + i2;
    double rho = 0.0, g[3] = {0.0, 0.0, 0.0};
    #pragma unroll
    for (int p = 0; p < M::NVEL; p++) {
      const double fp = __ldg(f + (i64) p*L.nsd + site);
      rho += fp;
      #pragma unroll
      for (int a = 0; a < M::NDIM; a++) {
	if (M::cv(p, a) ==  1) g[a] += fp;
	if (M::cv(p, a) == -1) g[a] -= fp;
      }
    }
    q[0] = rho; q[1] = rho*rho; q[2] = rho; q[3] = rho;
    q[4] = g[0]; q[5] = g[1]; q[6] = g[2];
    if (NDIST == 2) {
      double ph = 0.0;
      #pragma unroll
      for (int p = 0; p < M::NVEL; p++) ph += __ldg(f + (i64)(M::NVEL + p)*L.nsd + site);
      q[7] = ph;
    }
  }
  const i64 b = ((i64) blockIdx.z*gridDim.y + blockIdx.y)*gridDim.x + blockIdx.x;
  block_reduce8(q, partial + 8*b);
}


// ---------------------------------------------------------------------------
// rho and u of the state that entered the last collision, recomputed from the
// post-collision populations still sitting (unstreamed) at each site: mass is a
// collision invariant and the collision adds the full force to the momentum, so
//   rho = sum_p f*_p,   rho u = sum_p c_p f*_p - F/2      (src/collision.c:355-382, :523-525)
// Used when the collision did not store them (option hydro_write = 0) and somebody
// asks: identical to the stored values to round-off.
// ---------------------------------------------------------------------------

template <class M, bool BINARY>
__global__ void __launch_bounds__(256)
hydro_from_fstar_kernel(const Layout L, const double * __restrict__ f,
			const double * __restrict__ force, const double fg0,
			const double fg1, const double fg2,
			double * __restrict__ rho, double * __restrict__ u,
			const int mode, const int w0, const int w1, const int w2) {

  // mode: where the post-collision population p of a site is kept. 0: its slot at the
  // site; single-lattice streaming: 1 = the slot of its opposite at the site (after an
  // even step), 2 = its slot at the site it streams to (after an odd step)
  const int i2 = blockIdx.x*blockDim.x + threadIdx.x;
  const int i1 = blockIdx.y*blockDim.y + threadIdx.y;
  const int i0 = blockIdx.z;
  if (i2 >= L.n[2] || i1 >= L.n[1]) return;
  const i64 site = (i64)(i0 + L.h[0])*L.s0 + (i64)(i1 + L.h[1])*L.s1 + L.z0 + L.h[2] + i2;
  // offsets that reach coordinate +1 / -1, wrapping around axes periodic within the rank
  i64 op[3] = {L.s0, L.s1, 1}, om[3] = {-L.s0, -L.s1, -1};
  if (mode == 2) {
    if (w0 && i0 == L.n[0] - 1) op[0] = -(i64)(L.n[0] - 1)*L.s0;
    if (w0 && i0 == 0) om[0] = (i64)(L.n[0] - 1)*L.s0;
    if (w1 && i1 == L.n[1] - 1) op[1] = -(i64)(L.n[1] - 1)*L.s1;
    if (w1 && i1 == 0) om[1] = (i64)(L.n[1] - 1)*L.s1;
    if (w2 && i2 == L.n[2] - 1) op[2] = -(i64)(L.n[2] - 1);
    if (w2 && i2 == 0) om[2] = (i64)(L.n[2] - 1);
  }

  double r = 0.0, g[3] = {0.0, 0.0, 0.0};
  #pragma unroll
  for (int p = 0; p < M::NVEL; p++) {
    i64 at = (i64) p*L.nsd + site;
    if (mode == 1 && p > 0) at = (i64)(M::NVEL - p)*L.nsd + site;
    if (mode == 2) {
      #pragma unroll
      for (int d = 0; d < 3; d++) {
	if (dcv<M>(p, d) ==  1) at += op[d];
	if (dcv<M>(p, d) == -1) at += om[d];
      }
    }
    const double fp = __ldg(f + at);
    r += fp;
    #pragma unroll
    for (int a = 0; a < M::NDIM; a++) {
      if (M::cv(p, a) ==  1) g[a] += fp;
      if (M::cv(p, a) == -1) g[a] -= fp;
    }
  }
  double F[3] = {fg0, fg1, fg2};
  if (force != nullptr) {
    #pragma unroll
    for (int a = 0; a < M::NDIM; a++) F[a] += force[a*L.nsd + site];
  }
  const double rr = 1.0/r;
  if (!BINARY) rho[site] = r;          // the binary collision never stores rho
  #pragma unroll
  for (int a = 0; a < 3; a++) {
    u[a*L.nsd + site] = (a < M::NDIM) ? rr*(g[a] - 0.5*F[a]) : 0.0;
  }
}

// solid_grad_kernels.cuh -- the 27-point gradient next to flat walls
// (grad_3d_27pt_solid_d2, src/gradient_3d_27pt_solid.c:131-325; "fd_gradient_calculation
// 3d_27pt_solid") with neutral wetting (C = H = 0): differences to the 26 neighbours, the
// links that end inside a wall contributing the boundary gradient -(C phi_b + H)/kappa,
// which is zero; delsq = (1/9) sum, grad_a = (1/18) sum c_a over the links in the order of
// the reference's table (bs_cv: x slowest, z fastest). A neighbour is solid when its
// coordinate along a wall direction of this rank lies outside the lattice.
//
// One thread per site, neighbours through L1/L2 (8 B read + 32 B written per site, as
// grad27_kernel); the binary fluid between walls is not a headline configuration.

#pragma once

#include "layout.cuh"

struct SolidGradTab {
  Layout L;
  int wrap[3];                    // axis periodic within this rank: neighbours wrap
  int wall_lo[3], wall_hi[3];     // this rank holds the low / high wall of the axis
};

__global__ void __launch_bounds__(128)
grad27_solid_kernel(const SolidGradTab T, const double * __restrict__ phi,
		    double * __restrict__ grad, double * __restrict__ delsq) {
  const Layout & L = T.L;
  const int i2 = blockIdx.x*blockDim.x + threadIdx.x;
  const int i1 = blockIdx.y;
  const int i0 = blockIdx.z;
  if (i2 >= L.n[2]) return;
  const int c[3] = {i0 + L.h[0], i1 + L.h[1], i2 + L.h[2]};
  const i64 site = (i64) c[0]*L.s0 + (i64) c[1]*L.s1 + L.z0 + c[2];
  const double ph0 = __ldg(phi + site);
  const double r9 = (1.0/9.0), r18 = (1.0/18.0);

  // per axis and step (-1, 0, +1): offset in doubles, and whether the step ends in a wall
  i64 off[3][3];
  bool sol[3][3];
  const i64 str[3] = {L.s0, L.s1, 1};
  #pragma unroll
  for (int d = 0; d < 3; d++) {
    #pragma unroll
    for (int s = -1; s <= 1; s++) {
      int t = c[d] + s;
      sol[d][s + 1] = (s == -1 && c[d] == 1 && T.wall_lo[d]) || (s == 1 && c[d] == L.n[d] && T.wall_hi[d]);
      if (T.wrap[d]) { if (t < 1) t += L.n[d]; else if (t > L.n[d]) t -= L.n[d]; }
      off[d][s + 1] = (i64) (t - c[d])*str[d];
    }
  }
  double dphi = 0.0, gn[3] = {0.0, 0.0, 0.0};
  #pragma unroll
  for (int a = -1; a <= 1; a++) {
    #pragma unroll
    for (int b = -1; b <= 1; b++) {
      #pragma unroll
      for (int q = -1; q <= 1; q++) {
	if (a == 0 && b == 0 && q == 0) continue;
	const bool solid = sol[0][a + 1] || sol[1][b + 1] || sol[2][q + 1];
	// a link into a wall carries -(C phi_b + H)/kappa = -0 with neutral wetting
	const double gt = solid ? -0.0 : (__ldg(phi + site + off[0][a + 1] + off[1][b + 1] + off[2][q + 1]) - ph0);
	dphi += gt;
	gn[0] += gt*a; gn[1] += gt*b; gn[2] += gt*q;
      }
    }
  }
  delsq[site] = r9*dphi;
  grad[site]           = r18*gn[0];
  grad[L.nsd + site]   = r18*gn[1];
  grad[2*L.nsd + site] = r18*gn[2];
}

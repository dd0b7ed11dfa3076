// open_bc_kernels.cuh -- open boundaries for the fluid: inflow / outflow of type "rhou"
// (src/lb_bc_inflow_rhou.c, src/lb_bc_outflow_rhou.c; chosen in src/lb_bc_open_rt.c:48-170)
// along the one direction that has no walls. Between lb_halo and lb_propagation
// (src/ludwig.c:820-831) the reference
//   update: gives the layer of halo sites in front of the inlet the density of the fluid
//           sites behind it and the prescribed velocity u0, and the layer behind the outlet
//           the prescribed density rho0 and the velocity of the fluid sites in front of it;
//   impose: sets every population the propagation will carry from such a halo site into
//           the fluid to the equilibrium of the halo site's (rho, u) (host loops over link
//           lists, :478-531 of either file).
// Here one thread owns one halo site of either layer and writes its inward populations
// straight into the lattice; nothing else moves. The flow direction is not periodic, so
// the pull of the sweep that follows reads these planes as they are.
//
// Three-dimensional models: device axes are (x, y, z).

#pragma once

#include "layout.cuh"

struct OpenTab {
  Layout L;
  int nvel;
  int id;                         // flow axis
  int da, db;                     // the two axes spanning the face, db fastest (device axes)
  int has_in, has_out;            // this rank holds the inlet / the outlet
  int lo[3], hi[3];               // interior extent of the WHOLE system in local coordinates
                                  // (links whose halo end lies outside belong to the walls)
  signed char cv[27][3];
  double wv[27];
  double u0[3];
  double rho0;
};

// grid (tiles along db, n_da, 2), block (128): blockIdx.z = 0 inlet, 1 outlet
__global__ void __launch_bounds__(128)
open_bc_kernel(const OpenTab T, const double * __restrict__ rho, const double * __restrict__ u,
	       double * __restrict__ f) {
  const Layout & L = T.L;
  const int side = blockIdx.z;
  if (side == 0 ? !T.has_in : !T.has_out) return;
  const int ib = blockIdx.x*blockDim.x + threadIdx.x;      // interior, 0-based
  const int ia = blockIdx.y;
  if (ib >= L.n[T.db]) return;
  const int dir = (side == 0) ? -1 : +1;

  int c[3];                                                // halo-inclusive coordinates
  c[T.da] = ia + L.h[T.da];
  c[T.db] = ib + L.h[T.db];
  c[T.id] = (side == 0) ? L.h[T.id] : L.n[T.id] + L.h[T.id] - 1;     // the fluid site
  const i64 fluid = (i64) c[0]*L.s0 + (i64) c[1]*L.s1 + L.z0 + c[2];
  c[T.id] += dir;                                                    // the halo site
  const i64 halo = (i64) c[0]*L.s0 + (i64) c[1]*L.s1 + L.z0 + c[2];

  double rh, uh[3];
  if (side == 0) {
    rh = rho[fluid];
    uh[0] = T.u0[0]; uh[1] = T.u0[1]; uh[2] = T.u0[2];
  }
  else {
    rh = T.rho0;
    uh[0] = u[fluid]; uh[1] = u[L.nsd + fluid]; uh[2] = u[2*L.nsd + fluid];
  }
  const double cs2 = 1.0/3.0, rcs2 = 3.0;
  for (int q = 1; q < T.nvel; q++) {
    if (T.cv[q][T.id] != -dir) continue;                   // into the fluid
    // the fluid end of the link must be a fluid site (else nobody pulls this population)
    const int ea = ia + T.cv[q][T.da], eb = ib + T.cv[q][T.db];
    if (ea < T.lo[T.da] || ea > T.hi[T.da] || eb < T.lo[T.db] || eb > T.hi[T.db]) continue;
    double udotc = 0.0, sdotq = 0.0;
    #pragma unroll
    for (int a = 0; a < 3; a++) {
      udotc += uh[a]*T.cv[q][a];
      #pragma unroll
      for (int b = 0; b < 3; b++) {
	const double d_ab = (a == b);
	const double s_ab = T.cv[q][a]*T.cv[q][b] - cs2*d_ab;
	sdotq += s_ab*uh[a]*uh[b];
      }
    }
    f[(i64) q*L.nsd + halo] = rh*T.wv[q]*(1.0 + rcs2*udotc + 0.5*rcs2*rcs2*sdotq);
  }
}

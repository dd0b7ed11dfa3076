// zpack_kernels.cuh -- the packed z-face halo of peer-memory sweeps, unpacked on demand.

#pragma once

#include "layout.cuh"

// ---------------------------------------------------------------------------
// Packed halo of the contiguous axis (KParams::zoff, site_kernels.cuh): while peer-memory
// sweeps follow one another the populations that crossed a z face stay in the dense planes
// behind the lattice. Anything else that reads halo sites (a propagation on its own, a halo
// exchange that copies edges on, a download) first gets them where it expects them: the
// row-end halo words. Rare, so one thread per value and no attempt at coalescing the
// scattered side.
//   x0..x1, y0..y1: halo-inclusive columns that hold pushed values (the apron of a
//   direction only when that direction is decomposed too)
// ---------------------------------------------------------------------------

struct ZPackTab {
  int np;              // populations per face
  int pidx[2][9];      // face 0: c_z = +1 (low halo), face 1: c_z = -1 (high halo), index order
};

__global__ void __launch_bounds__(256)
zhalo_unpack_kernel(const Layout L, const ZPackTab T, double * __restrict__ lat, const i64 zoff,
		    const i64 zsx, const i64 zsy, const i64 zsp,
		    const int x0, const int x1, const int y0, const int y1) {
  const int ya = y0 + blockIdx.x*blockDim.x + threadIdx.x;
  const int xa = x0 + blockIdx.y;
  const int pl = blockIdx.z;                 // face*np + k
  if (ya >= y1 || xa >= x1) return;
  const int face = pl/T.np, k = pl - face*T.np;
  const int p = T.pidx[face][k];
  const double val = lat[zoff + (i64) pl*zsp + (i64) xa*zsx + (i64) ya*zsy];
  const i64 zw = (face == 0) ? (L.h[2] - 1) : (L.h[2] + L.n[2]);
  lat[(i64) p*L.nsd + (i64) xa*L.s0 + (i64) ya*L.s1 + L.z0 + zw] = val;
}

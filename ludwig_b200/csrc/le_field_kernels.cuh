// le_field_kernels.cuh -- the order parameter of the binary fluid across Lees-Edwards
// planes: field_leesedwards (src/field.c:500-600) and the plane-aware x neighbours of the
// gradient stencils (lees_edw_ic_to_buff as called from
// src/gradient_3d_7pt_fluid.c:264-265 and src/gradient_3d_27pt_fluid.c:262-265).
//
// The reference keeps, behind the real x-planes of phi, two buffer planes per plane,
//   buffer 2p     = the sites x = loc(p)     as seen from x = loc(p) + 1,
//   buffer 2p + 1 = the sites x = loc(p) + 1 as seen from x = loc(p),
// displaced along y by what the plane has slid so far (-/+ u_y t, four-point Lagrange
// interpolation at the fractional part), and lets the stencil of a site next to a plane
// take its x - 1 / x + 1 neighbours from them. Here grad7_kernel / grad27_kernel first run
// over the whole lattice as if there were no planes (their rolling windows stay as they
// are); le_phi_buffer_kernel then fills the buffer planes and le_grad_fix_kernel
// recomputes the 2 x nplanes x-planes of sites next to a plane with them -- same
// operations in the same order as the main kernels, so the values are the reference's bit
// for bit. Cost: 2 nplanes planes out of n_x; y is not decomposed (as for the
// distributions, b200lb_le_commit), z may be.
//
// Binary fluid = three-dimensional models: device axes are (x, y, z).

#pragma once

#include <math.h>

#include "layout.cuh"

#ifndef B200LB_LE_MAXPLANE
#define B200LB_LE_MAXPLANE 32
#endif

struct LePhiTab {
  Layout L;
  int w0, w1, w2;                     // device axis periodic within this rank: neighbours wrap
  int nbuf;                           // 2 x local planes
  int xsrc[2*B200LB_LE_MAXPLANE];     // halo-inclusive x of the real plane buffer k is taken from
  int xsite[2*B200LB_LE_MAXPLANE];    // halo-inclusive x of the sites that see buffer k ...
  int dir[2*B200LB_LE_MAXPLANE];      // ... as their x + dir neighbour (-1 or +1)
  int jdy[2*B200LB_LE_MAXPLANE];      // integer part of the displacement
  double c[2*B200LB_LE_MAXPLANE][4];  // Lagrange weights of the four sites j0 .. j3
};

// Displacements at time t = physics_control_time (src/physics.c:651-659: t_start +
// t_current - 1, NOT the step number the distributions use) and the interpolation
// weights, evaluated on the host in the reference's association
// (src/field.c:553-556, 577-580; lees_edw_buffer_displacement, src/leesedwards.c:633-655).
// loc[p]: local 1-based x of plane p (the plane sits between loc and loc + 1).
static inline void le_phi_tab_fill(LePhiTab & T, int nplanes_local, const int * loc,
				   double uy, int oscillatory, int period, int nt0, double t) {
  const double r6 = (1.0/6.0);
  const int ny = T.L.n[1];
  T.nbuf = 2*nplanes_local;
  for (int k = 0; k < T.nbuf; k++) {
    const int p = k/2;
    const double tle = ((t < 0.0) ? 0.0 : t) - 1.0*nt0;
    double dy = tle*uy*((k % 2 < 1) ? -1 : +1);
    if (oscillatory) {
      const double omega = 2.0*4.0*atan(1.0)/period;
      dy = uy*sin(omega*tle)/omega;
    }
    dy = fmod(dy, 1.0*ny);
    const int jdy = (int) floor(dy);
    const double fr = 1.0 - (dy - jdy);
    T.xsrc[k]  = loc[p] + (k % 2);
    T.xsite[k] = loc[p] + 1 - (k % 2);
    T.dir[k]   = (k % 2) ? +1 : -1;
    T.jdy[k]   = jdy;
    T.c[k][0] = -r6*fr*(fr-1.0)*(fr-2.0);
    T.c[k][1] = 0.5*(fr*fr-1.0)*(fr-2.0);
    T.c[k][2] = 0.5*fr*(fr+1.0)*(fr-2.0);
    T.c[k][3] = r6*fr*(fr*fr-1.0);
  }
}

// buf[nbuf][s0]: every buffer plane has the layout of an x-plane of phi.
// grid (z tiles, n_y + 2, nbuf), block (128): jc = blockIdx.y in 0 .. n_y + 1 (the y halo
// rows are periodic images by construction), kc over the interior and, where z is not
// periodic within this rank, its two halo columns (phi's z halo is current then).
__global__ void __launch_bounds__(128)
le_phi_buffer_kernel(const LePhiTab T, const double * __restrict__ phi, double * __restrict__ buf) {
  const Layout & L = T.L;
  const int k = blockIdx.z;
  const int jc = blockIdx.y;
  const int kc = blockIdx.x*blockDim.x + threadIdx.x + (T.w2 ? 1 : 0);
  if (kc > L.n[2] + (T.w2 ? 0 : 1)) return;
  const int ny = L.n[1];
  // src/field.c:564-567
  const int j0 = 1 + (jc - T.jdy[k] - 3 + 2*ny) % ny;
  const int j1 = 1 + j0 % ny;
  const int j2 = 1 + j1 % ny;
  const int j3 = 1 + j2 % ny;
  const double * pl = phi + (i64) T.xsrc[k]*L.s0 + L.z0 + kc;
  const double d0 = __ldg(pl + (i64) j0*L.s1), d1 = __ldg(pl + (i64) j1*L.s1);
  const double d2 = __ldg(pl + (i64) j2*L.s1), d3 = __ldg(pl + (i64) j3*L.s1);
  // c0 d0 + c1 d1 - c2 d2 + c3 d3, every product and sum separately rounded
  double v = __dadd_rn(__dmul_rn(T.c[k][0], d0), __dmul_rn(T.c[k][1], d1));
  v = __dadd_rn(v, -__dmul_rn(T.c[k][2], d2));
  v = __dadd_rn(v, __dmul_rn(T.c[k][3], d3));
  buf[(i64) k*L.s0 + (i64) jc*L.s1 + L.z0 + kc] = v;
}

// One site of either stencil from the three x-planes it spans (plm, plc, plp: the planes at
// x - 1, x, x + 1, each laid out like an x-plane of phi; col: the site's offset in a plane;
// ymo .. zpo: offsets to its y and z neighbours). NPT = 7 or 27; the arithmetic is
// grad7_kernel's / grad27_kernel's (common_kernels.cuh), term by term.
template <int NPT>
__device__ __forceinline__ void
le_stencil_site(const double * __restrict__ plm, const double * __restrict__ plc,
		const double * __restrict__ plp, i64 col, i64 ymo, i64 ypo, i64 zmo, i64 zpo,
		double (&g)[3], double & dsq) {
  if (NPT == 7) {
    const double c  = __ldg(plc + col);
    const double xp = __ldg(plp + col), xm = __ldg(plm + col);
    const double yp = __ldg(plc + col + ypo), ym = __ldg(plc + col + ymo);
    const double zp = __ldg(plc + col + zpo), zm = __ldg(plc + col + zmo);
    g[0] = 0.5*(xp - xm);
    g[1] = 0.5*(yp - ym);
    g[2] = 0.5*(zp - zm);
    dsq = xp + xm + yp + ym + zp + zm - __dmul_rn(6.0, c);
  }
  else {
    const i64 yo[3] = {ymo, 0, ypo}, zo[3] = {zmo, 0, zpo};
    double m[3][3], c[3][3], p[3][3];
    #pragma unroll
    for (int j = 0; j < 3; j++) {
      #pragma unroll
      for (int q = 0; q < 3; q++) {
	const i64 o = col + yo[j] + zo[q];
	m[j][q] = __ldg(plm + o); c[j][q] = __ldg(plc + o); p[j][q] = __ldg(plp + o);
      }
    }
    const double r9 = 1.0/9.0;
    const double hr9 = 0.5*r9;
    double gx = p[0][0] - m[0][0];
    double gy = m[2][0] - m[0][0];
    double gz = m[0][2] - m[0][0];
    double d2 = m[0][0];
    #pragma unroll
    for (int q = 1; q < 9; q++) {
      const int a = q/3, b = q - 3*a;
      gx = gx + p[a][b];  gx = gx - m[a][b];
      d2 = d2 + m[a][b];
    }
    #pragma unroll
    for (int q = 1; q < 3; q++) { gy = gy + m[2][q]; gy = gy - m[0][q]; }
    #pragma unroll
    for (int q = 0; q < 3; q++) { gy = gy + c[2][q]; gy = gy - c[0][q]; }
    #pragma unroll
    for (int q = 0; q < 3; q++) { gy = gy + p[2][q]; gy = gy - p[0][q]; }
    #pragma unroll
    for (int j = 1; j < 3; j++) { gz = gz + m[j][2]; gz = gz - m[j][0]; }
    #pragma unroll
    for (int j = 0; j < 3; j++) { gz = gz + c[j][2]; gz = gz - c[j][0]; }
    #pragma unroll
    for (int j = 0; j < 3; j++) { gz = gz + p[j][2]; gz = gz - p[j][0]; }
    #pragma unroll
    for (int q = 0; q < 9; q++) { if (q != 4) d2 = d2 + c[q/3][q - 3*(q/3)]; }
    #pragma unroll
    for (int q = 0; q < 9; q++) d2 = d2 + p[q/3][q - 3*(q/3)];
    d2 = d2 - __dmul_rn(26.0, c[1][1]);
    g[0] = hr9*gx;
    g[1] = hr9*gy;
    g[2] = hr9*gz;
    dsq  = r9*d2;
  }
}

// grid (z tiles, n_y, nbuf), block (128): the sites x = xsite[k] again, with the plane
// x + dir[k] taken from buffer k.
template <int NPT>
__global__ void __launch_bounds__(128)
le_grad_fix_kernel(const LePhiTab T, const double * __restrict__ phi,
		   const double * __restrict__ buf,
		   double * __restrict__ grad, double * __restrict__ delsq) {
  const Layout & L = T.L;
  const int k = blockIdx.z;
  const int c1 = blockIdx.y + 1;                               // halo-inclusive y, z
  const int c2 = blockIdx.x*blockDim.x + threadIdx.x + 1;
  if (c2 > L.n[2]) return;
  const int x = T.xsite[k];

  int xo = x - T.dir[k];                     // the neighbour plane on the other side: real
  if (T.w0) xo = (xo == 0) ? L.n[0] : ((xo == L.n[0] + 1) ? 1 : xo);
  const double * pb = buf + (i64) k*L.s0;
  const double * po = phi + (i64) xo*L.s0;
  const double * plm = (T.dir[k] < 0) ? pb : po;               // x - 1
  const double * plp = (T.dir[k] < 0) ? po : pb;               // x + 1
  const double * plc = phi + (i64) x*L.s0;

  const i64 col = (i64) c1*L.s1 + L.z0 + c2;
  const i64 zmo = (T.w2 && c2 == 1)      ? (i64) (L.n[2] - 1) : -1;
  const i64 zpo = (T.w2 && c2 == L.n[2]) ? -(i64) (L.n[2] - 1) : 1;
  const i64 ymo = (T.w1 && c1 == 1)      ? (i64) (L.n[1] - 1)*L.s1 : -L.s1;
  const i64 ypo = (T.w1 && c1 == L.n[1]) ? -(i64) (L.n[1] - 1)*L.s1 : L.s1;
  const i64 site = (i64) x*L.s0 + col;

  double g[3], dsq;
  le_stencil_site<NPT>(plm, plc, plp, col, ymo, ypo, zmo, zpo, g, dsq);
  grad[site]           = g[0];
  grad[L.nsd + site]   = g[1];
  grad[2*L.nsd + site] = g[2];
  delsq[site]          = dsq;
}

// ---------------------------------------------------------------------------
// The finite-difference order parameter ("free_energy symmetric") across the planes. The
// reference's halo is two sites wide there: FOUR buffer planes of phi per plane,
//   b = 0, 1: the real planes loc - 1, loc   displaced by -u_y t (seen from above),
//   b = 2, 3: the real planes loc + 1, loc + 2 displaced by +u_y t (seen from below)
// (le_phi_tab_fill4 + le_phi_buffer_kernel: buffer 4p + b), the gradient evaluated IN the
// first buffer plane of either side (grad_3d_*_fluid_le, nextra = 1:
// src/gradient_3d_7pt_fluid.c:325-460, src/gradient_3d_27pt_fluid.c:377-560) from the
// plane triples (b0, b1, real loc + 1) and (real loc, b2, b3), and buffer planes of the
// velocity at time t + 1, linearly interpolated (hydro_lees_edwards, src/hydro.c:412-500;
// only u_x is ever read there, and the velocity jump is along y).
// ---------------------------------------------------------------------------

// as le_phi_tab_fill, four buffers per plane: at most B200LB_LE_MAXPLANE/2 planes
static inline void le_phi_tab_fill4(LePhiTab & T, int nplanes_local, const int * loc,
				    double uy, int oscillatory, int period, int nt0, double t) {
  const double r6 = (1.0/6.0);
  const int ny = T.L.n[1];
  T.nbuf = 4*nplanes_local;
  for (int k = 0; k < T.nbuf; k++) {
    const int p = k/4, b = k % 4;
    const double tle = ((t < 0.0) ? 0.0 : t) - 1.0*nt0;
    double dy = tle*uy*((b < 2) ? -1 : +1);
    if (oscillatory) {
      const double omega = 2.0*4.0*atan(1.0)/period;
      dy = uy*sin(omega*tle)/omega;
    }
    dy = fmod(dy, 1.0*ny);
    const int jdy = (int) floor(dy);
    const double fr = 1.0 - (dy - jdy);
    T.xsrc[k]  = loc[p] - 1 + b;
    T.xsite[k] = 0;
    T.dir[k]   = 0;
    T.jdy[k]   = jdy;
    T.c[k][0] = -r6*fr*(fr-1.0)*(fr-2.0);
    T.c[k][1] = 0.5*(fr*fr-1.0)*(fr-2.0);
    T.c[k][2] = 0.5*fr*(fr+1.0)*(fr-2.0);
    T.c[k][3] = r6*fr*(fr*fr-1.0);
  }
}

// velocity buffers: entry k = 2p + s, s = 0: real plane loc (buffer b = 1), s = 1: real
// plane loc + 1 (b = 2); displacement at time t + 1; c[k][0] = fr, c[k][1] = 1 - fr
static inline void le_u_tab_fill(LePhiTab & T, int nplanes_local, const int * loc,
				 double uy, int oscillatory, int period, int nt0, double t) {
  const int ny = T.L.n[1];
  T.nbuf = 2*nplanes_local;
  for (int k = 0; k < T.nbuf; k++) {
    const int p = k/2, s = k % 2;
    const double t1 = t + 1.0;
    const double tle = ((t1 < 0.0) ? 0.0 : t1) - 1.0*nt0;
    double dy = tle*uy*((s == 0) ? -1 : +1);
    if (oscillatory) {
      const double omega = 2.0*4.0*atan(1.0)/period;
      dy = uy*sin(omega*tle)/omega;
    }
    dy = fmod(dy, 1.0*ny);
    const int jdy = (int) floor(dy);
    const double fr = dy - jdy;
    T.xsrc[k] = loc[p] + s;
    T.jdy[k]  = jdy;
    T.c[k][0] = fr;
    T.c[k][1] = 1.0 - fr;
  }
}

// bux[nbuf][s0] from the x component of u (one plane of the [3][nsd] array); interior (y, z)
// grid (z tiles, n_y, nbuf), block (128)
__global__ void __launch_bounds__(128)
le_u_buffer_kernel(const LePhiTab T, const double * __restrict__ ux, double * __restrict__ bux) {
  const Layout & L = T.L;
  const int k = blockIdx.z;
  const int jc = blockIdx.y + 1;
  const int kc = blockIdx.x*blockDim.x + threadIdx.x + 1;
  if (kc > L.n[2]) return;
  const int ny = L.n[1];
  const int j1 = 1 + (jc - T.jdy[k] - 2 + 2*ny) % ny;
  const int j2 = 1 + j1 % ny;
  const double * pl = ux + (i64) T.xsrc[k]*L.s0 + L.z0 + kc;
  // ule_x + u(j1) fr + u(j2) (1 - fr), ule_x = 0 (src/hydro.c:483-487)
  const double v = __dadd_rn(__dadd_rn(0.0, __dmul_rn(__ldg(pl + (i64) j1*L.s1), T.c[k][0])),
			     __dmul_rn(__ldg(pl + (i64) j2*L.s1), T.c[k][1]));
  bux[(i64) k*L.s0 + (i64) jc*L.s1 + L.z0 + kc] = v;
}

// gradient and Laplacian in the first buffer plane of either side of every plane:
// entry 2p + s, s = 0: buffer b = 1 from (b0, b1, real loc + 1), s = 1: buffer b = 2 from
// (real loc, b2, b3). bgrad[nplanes*2][3][s0], bdelsq[nplanes*2][s0]; interior (y, z).
// loc[p]: halo-inclusive x of the layer below plane p. bphi: le_phi_tab_fill4's buffers.
struct LeBufGradTab {
  Layout L;
  int w1, w2;
  int nplanes;
  int loc[B200LB_LE_MAXPLANE];
};

template <int NPT>
__global__ void __launch_bounds__(128)
le_buffer_grad_kernel(const LeBufGradTab T, const double * __restrict__ phi,
		      const double * __restrict__ bphi,
		      double * __restrict__ bgrad, double * __restrict__ bdelsq) {
  const Layout & L = T.L;
  const int e = blockIdx.z;                  // 2p + s
  const int p = e/2, s = e % 2;
  const int c1 = blockIdx.y + 1;
  const int c2 = blockIdx.x*blockDim.x + threadIdx.x + 1;
  if (c2 > L.n[2]) return;
  const double * b = bphi + (i64) 4*p*L.s0;
  const double * plm = (s == 0) ? b : phi + (i64) T.loc[p]*L.s0;
  const double * plc = (s == 0) ? b + L.s0 : b + 2*L.s0;
  const double * plp = (s == 0) ? phi + (i64) (T.loc[p] + 1)*L.s0 : b + 3*L.s0;
  const i64 col = (i64) c1*L.s1 + L.z0 + c2;
  const i64 zmo = (T.w2 && c2 == 1)      ? (i64) (L.n[2] - 1) : -1;
  const i64 zpo = (T.w2 && c2 == L.n[2]) ? -(i64) (L.n[2] - 1) : 1;
  const i64 ymo = (T.w1 && c1 == 1)      ? (i64) (L.n[1] - 1)*L.s1 : -L.s1;
  const i64 ypo = (T.w1 && c1 == L.n[1]) ? -(i64) (L.n[1] - 1)*L.s1 : L.s1;
  double g[3], dsq;
  le_stencil_site<NPT>(plm, plc, plp, col, ymo, ypo, zmo, zpo, g, dsq);
  #pragma unroll
  for (int a = 0; a < 3; a++) bgrad[((i64) e*3 + a)*L.s0 + col] = g[a];
  bdelsq[(i64) e*L.s0 + col] = dsq;
}

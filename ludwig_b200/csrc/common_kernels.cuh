// common_kernels.cuh -- model-independent kernels: halo planes, 7-point
// gradient, fills, the last stage of the diagnostics and the host<->device
// layout conversion. Included by api.cu only.

#pragma once

#include "layout.cuh"
#include "aux_kernels.cuh"

// ---------------------------------------------------------------------------
// Halo planes of one device axis. Restates the data movement of
// halo_swap_pack_rank1 / halo_swap_unpack_rank1 (src/halo_swap.c:1074-1274)
// and, for a direction with one rank, the self-exchange of
// halo_swap_packed (src/halo_swap.c:846-860): the face spans the FULL
// halo-inclusive extent of the other two axes, so doing the axes in order
// carries edges and corners without diagonal messages.
//
// "lo list" = planes whose values travel into a LOW halo (they are read at
// coordinate n of the sender), "hi list" likewise into a HIGH halo (read at
// coordinate 1). With the full scheme both lists hold every plane; with the
// reduced scheme only the populations that stream inwards (src/model.c:1192-1219).
// Send/recv buffers are plane-major: buf[k*nface + facesite].
// ---------------------------------------------------------------------------

#define B200LB_MAXEL 54

enum { HALO_LOCAL = 0, HALO_PACK = 1, HALO_UNPACK = 2 };

struct HaloParams {
  Layout L;
  double * data;
  int axis;
  int nlo, nhi;
  unsigned char el_lo[B200LB_MAXEL];
  unsigned char el_hi[B200LB_MAXEL];
  double * to_hi;      // PACK: values for the upper neighbour's low halo
  double * to_lo;      // PACK: values for the lower neighbour's high halo
  const double * from_lo;  // UNPACK: into my low halo
  const double * from_hi;  // UNPACK: into my high halo
};

template <int MODE>
__global__ void __launch_bounds__(256)
halo_kernel(const HaloParams H) {

  const Layout & L = H.L;
  const int d = H.axis;
  // the two other axes, fa the faster one
  const int fa = (d == 2) ? 1 : 2;
  const int sl = (d == 0) ? 1 : 0;
  const int ia = blockIdx.x*blockDim.x + threadIdx.x;
  const int ib = blockIdx.y*blockDim.y + threadIdx.y;
  if (ia >= L.na[fa] || ib >= L.na[sl]) return;
  const int k = blockIdx.z;            // entry of the lo list, then of the hi list

  const i64 st_d  = (d == 0) ? L.s0 : ((d == 1) ? L.s1 : 1);
  const i64 st_fa = (fa == 1) ? L.s1 : 1;
  const i64 st_sl = (sl == 0) ? L.s0 : L.s1;
  const i64 base = L.z0 + (i64) ia*st_fa + (i64) ib*st_sl;
  const i64 nface = (i64) L.na[fa]*L.na[sl];
  const i64 fs = (i64) ib*L.na[fa] + ia;
  const int n = L.n[d];

  if (k < H.nlo) {
    double * plane = H.data + (i64) H.el_lo[k]*L.nsd + base;
    if (MODE == HALO_LOCAL)  plane[0] = plane[(i64) n*st_d];
    if (MODE == HALO_PACK)   H.to_hi[k*nface + fs] = plane[(i64) n*st_d];
    if (MODE == HALO_UNPACK) plane[0] = H.from_lo[k*nface + fs];
  }
  else {
    const int kk = k - H.nlo;
    double * plane = H.data + (i64) H.el_hi[kk]*L.nsd + base;
    if (MODE == HALO_LOCAL)  plane[(i64) (n + 1)*st_d] = plane[st_d];
    if (MODE == HALO_PACK)   H.to_lo[kk*nface + fs] = plane[st_d];
    if (MODE == HALO_UNPACK) plane[(i64) (n + 1)*st_d] = H.from_hi[kk*nface + fs];
  }
}

// ---------------------------------------------------------------------------
// 7-point gradient and Laplacian (src/gradient_3d_7pt_fluid.c:232-313, nextra
// = 0): grad = (phi(+) - phi(-))/2, delsq summed in the order
// +x -x +y -y +z -z -6 phi so that results are bit-identical.
//
// One thread per site, seven loads through the read-only path. No shared-memory staging:
// a z-y plane of phi is ~0.5 MB at 256^2, so the three planes a block's sites touch sit in
// L2 (126 MB) and the y and z neighbours in L1; DRAM sees each phi once (8 B/site) and the
// four output planes (32 B/site). Round 1 staged planes in a rolling shared-memory window
// with two block barriers per plane, which serialised one global-load latency per plane:
// 48 % of the HBM peak (0.210 ms at 256^3, profiles/r1_aux_kernels_256.log) against 87 %
// for this form (0.1175 ms, profiles/r2_aux_kernels_256.log).
// ---------------------------------------------------------------------------

#define GRAD_TZ 128
#define GRAD_TY 2
#define GRAD_NX 1      // x-planes per block (grid.z = n_x)

__global__ void __launch_bounds__(GRAD_TZ*GRAD_TY)
grad7_kernel(const Layout L, const int w0, const int w1, const int w2,
	     const double * __restrict__ phi,
	     double * __restrict__ grad, double * __restrict__ delsq) {

  const int i2 = blockIdx.x*GRAD_TZ + threadIdx.x;          // interior coordinates
  const int i1 = blockIdx.y*GRAD_TY + threadIdx.y;
  const int i0 = blockIdx.z;
  if (i2 >= L.n[2] || i1 >= L.n[1]) return;
  const i64 site = (i64)(i0 + L.h[0])*L.s0 + (i64)(i1 + L.h[1])*L.s1 + L.z0 + L.h[2] + i2;

  // Offsets to the -1 / +1 neighbours. A direction that is periodic within this rank
  // (w = 1) wraps around instead of reading the halo plane.
  const i64 xmo = (w0 && i0 == 0)           ? (i64) (L.n[0] - 1)*L.s0 : -L.s0;
  const i64 xpo = (w0 && i0 == L.n[0] - 1)  ? -(i64) (L.n[0] - 1)*L.s0 : L.s0;
  const i64 ymo = (w1 && i1 == 0)           ? (i64) (L.n[1] - 1)*L.s1 : -L.s1;
  const i64 ypo = (w1 && i1 == L.n[1] - 1)  ? -(i64) (L.n[1] - 1)*L.s1 : L.s1;
  const i64 zmo = (w2 && i2 == 0)           ? (i64) (L.n[2] - 1) : -1;
  const i64 zpo = (w2 && i2 == L.n[2] - 1)  ? -(i64) (L.n[2] - 1) : 1;

  const double * ph = phi + site;
  const double c  = __ldg(ph);
  const double xp = __ldg(ph + xpo), xm = __ldg(ph + xmo);
  const double yp = __ldg(ph + ypo), ym = __ldg(ph + ymo);
  const double zp = __ldg(ph + zpo), zm = __ldg(ph + zmo);
  grad[site]           = 0.5*(xp - xm);
  grad[L.nsd + site]   = 0.5*(yp - ym);
  grad[2*L.nsd + site] = 0.5*(zp - zm);
  // __dmul_rn keeps the product separately rounded (no FMA contraction): bit parity
  delsq[site] = xp + xm + yp + ym + zp + zm - __dmul_rn(6.0, c);
}

// The 27-point arithmetic of one site from its 3 x 3 x 3 neighbourhood, m = plane x - 1,
// c = plane x, p = plane x + 1, indexed [y][z]: the reference's order of additions.
__device__ __forceinline__ void
grad27_site(const Layout & L, const double (&m)[3][3], const double (&c)[3][3], const double (&p)[3][3],
	    const i64 site, double * __restrict__ grad, double * __restrict__ delsq) {
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
  // d/dy: planes x-1, x, x+1; z ascending.  d/dz: planes, y ascending.
  #pragma unroll
  for (int k = 1; k < 3; k++) { gy = gy + m[2][k]; gy = gy - m[0][k]; }
  #pragma unroll
  for (int k = 0; k < 3; k++) { gy = gy + c[2][k]; gy = gy - c[0][k]; }
  #pragma unroll
  for (int k = 0; k < 3; k++) { gy = gy + p[2][k]; gy = gy - p[0][k]; }
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
  // separately rounded product (no FMA contraction): bit parity
  d2 = d2 - __dmul_rn(26.0, c[1][1]);
  grad[site]           = hr9*gx;
  grad[L.nsd + site]   = hr9*gy;
  grad[2*L.nsd + site] = hr9*gz;
  delsq[site]          = r9*d2;
}

// The same stencil without staging: a thread marches G27M_NX sites along x and keeps the
// 3 x 3 x 3 neighbourhood of its site in registers, shifting it by one plane per step, so a
// site costs 9 loads through the read-only path (its row and the two next to it, in L1; the
// plane itself in L2) and no barrier. One thread per site with 27 loads is L1-bound
// (0.218 ms at 256^3); planes staged in shared memory with a barrier per plane take 0.182 ms;
// this form 0.180 ms (profiles/r2_aux_kernels_256.log). All three do the same 86 FP64
// additions per site in the reference's order, four dependent chains, no FMA: that, not
// data movement, bounds the kernel (8 TFLOP/s of plain adds).
#define G27M_NX 8

// Tried and dropped (profiles/r2b_aux_kernels_256.log): requesting the nine values of the plane
// after next before the additions start, so that a warp's own loads and arithmetic overlap.
// The nine extra live values cost registers (94, or 80 with a few spills): 0.249 and 0.201 ms
// against 0.180 ms -- the kernel lives on occupancy, not on per-warp overlap.
__global__ void __launch_bounds__(256)
grad27_direct_kernel(const Layout L, const int w0, const int w1, const int w2,
		     const double * __restrict__ phi,
		     double * __restrict__ grad, double * __restrict__ delsq) {
  const int i2 = blockIdx.x*blockDim.x + threadIdx.x;
  const int i1 = blockIdx.y*blockDim.y + threadIdx.y;
  const int x0 = blockIdx.z*G27M_NX;
  const int x1 = min(x0 + G27M_NX, L.n[0]);
  if (i2 >= L.n[2] || i1 >= L.n[1]) return;
  i64 oy[3], oz[3];
  oy[0] = (w1 && i1 == 0)          ? (i64) (L.n[1] - 1)*L.s1 : -L.s1;
  oy[1] = 0;
  oy[2] = (w1 && i1 == L.n[1] - 1) ? -(i64) (L.n[1] - 1)*L.s1 : L.s1;
  oz[0] = (w2 && i2 == 0)          ? (i64) (L.n[2] - 1) : -1;
  oz[1] = 0;
  oz[2] = (w2 && i2 == L.n[2] - 1) ? -(i64) (L.n[2] - 1) : 1;
  const i64 col = (i64)(i1 + L.h[1])*L.s1 + L.z0 + L.h[2] + i2;
  // halo-inclusive plane index xa (0 .. n_x + 1), wrapped where x is periodic within the rank
  auto plane = [&](int xa) {
    if (w0) xa = (xa == 0) ? L.n[0] : ((xa == L.n[0] + 1) ? 1 : xa);
    return phi + (i64) xa*L.s0 + col;
  };
  double m[3][3], c[3][3], p[3][3];
  {
    const double * pm = plane(x0), * pc = plane(x0 + 1);
    #pragma unroll
    for (int j = 0; j < 3; j++) {
      #pragma unroll
      for (int k = 0; k < 3; k++) { m[j][k] = __ldg(pm + oy[j] + oz[k]); c[j][k] = __ldg(pc + oy[j] + oz[k]); }
    }
  }
  for (int x = x0; x < x1; x++) {
    const double * pp = plane(x + 2);
    #pragma unroll
    for (int j = 0; j < 3; j++) {
      #pragma unroll
      for (int k = 0; k < 3; k++) p[j][k] = __ldg(pp + oy[j] + oz[k]);
    }
    const i64 site = (i64)(x + L.h[0])*L.s0 + col;
    grad27_site(L, m, c, p, site, grad, delsq);
    #pragma unroll
    for (int j = 0; j < 3; j++) {
      #pragma unroll
      for (int k = 0; k < 3; k++) { m[j][k] = c[j][k]; c[j][k] = p[j][k]; }
    }
  }
}

// ---------------------------------------------------------------------------
// 27-point gradient and Laplacian (grad_3d_27pt_fluid_d2,
// src/gradient_3d_27pt_fluid.c:75,223-352, nextra = 0):
//   d_a phi = (1/2)(1/9) sum over the 9 neighbours in the plane normal to a of
//             [phi(+a) - phi(-a)],   delsq = (1/9) [sum of 26 neighbours - 26 phi].
// The terms are added in the reference's order so the results are bit-identical:
//   d/dx (dy,dz) ascending; d/dy dx then dz ascending; d/dz dx then dy ascending,
//   each pair as +phi(+a) then -phi(-a); delsq (dx,dy,dz) ascending, centre last.
//
// One block marches a (TY x TZ) tile of columns along x. Each z-y plane (with its
// one-site apron, corners included) passes through shared memory once; a thread
// keeps the 3 x 3 x 3 neighbourhood of its site in registers and shifts it by one
// plane per step, so a site costs 9 shared loads instead of 80.
// Traffic: 8 B read + 32 B written per site (+ 2/G27_NX plane re-reads).
// ---------------------------------------------------------------------------

#define G27_TZ 64
#define G27_TY 4
#define G27_NX 32
#define G27_NE ((G27_TY + 2)*(G27_TZ + 2))

__global__ void __launch_bounds__(G27_TZ*G27_TY)
grad27_kernel(const Layout L, const int w0, const int w1, const int w2,
	      const double * __restrict__ phi,
	      double * __restrict__ grad, double * __restrict__ delsq) {

  __shared__ double sm[2][G27_TY + 2][G27_TZ + 2];

  const int tz = threadIdx.x, ty = threadIdx.y;
  const int tid = ty*G27_TZ + tz;
  const int i2 = blockIdx.x*G27_TZ + tz;           // interior coordinates, from 0
  const int i1 = blockIdx.y*G27_TY + ty;
  const int x0 = blockIdx.z*G27_NX;
  const int x1 = min(x0 + G27_NX, L.n[0]);
  const bool active = (i2 < L.n[2]) && (i1 < L.n[1]);

  // The two tile elements this thread brings in per plane: halo-inclusive (y, z),
  // clamped to the lattice, wrapped where the direction is periodic within the rank.
  i64 off[2];
  int er[2], ec[2];
  #pragma unroll
  for (int k = 0; k < 2; k++) {
    const int e = min(tid + k*G27_TZ*G27_TY, G27_NE - 1);
    er[k] = e/(G27_TZ + 2);
    ec[k] = e - er[k]*(G27_TZ + 2);
    int cy = min((int) blockIdx.y*G27_TY + er[k], L.n[1] + 1);
    int cz = min((int) blockIdx.x*G27_TZ + ec[k], L.n[2] + 1);
    if (w1) cy = (cy == 0) ? L.n[1] : ((cy == L.n[1] + 1) ? 1 : cy);
    if (w2) cz = (cz == 0) ? L.n[2] : ((cz == L.n[2] + 1) ? 1 : cz);
    off[k] = (i64) cy*L.s1 + L.z0 + cz;
  }
  const bool second = (tid + G27_TZ*G27_TY < G27_NE);

  // global -> registers -> shared, so that the loads of plane x + 3 are in flight while
  // plane x + 1 .. x - 1 is being worked on
  auto fetch = [&](int xa, double (&r)[2]) {
    if (w0) xa = (xa == 0) ? L.n[0] : ((xa == L.n[0] + 1) ? 1 : xa);
    const double * pl = phi + (i64) xa*L.s0;
    r[0] = __ldg(pl + off[0]);
    r[1] = second ? __ldg(pl + off[1]) : 0.0;
  };
  auto stash = [&](int slot, const double (&r)[2]) {
    sm[slot][er[0]][ec[0]] = r[0];
    if (second) sm[slot][er[1]][ec[1]] = r[1];
  };

  double m[3][3], c[3][3], p[3][3];
  double nxt[2];

  fetch(x0, nxt);                       // halo-inclusive plane of interior x0 - 1
  stash(0, nxt);
  fetch(x0 + 1, nxt);
  stash(1, nxt);
  fetch(x0 + 2, nxt);
  __syncthreads();
  #pragma unroll
  for (int j = 0; j < 3; j++)
    #pragma unroll
    for (int k = 0; k < 3; k++) { m[j][k] = sm[0][ty + j][tz + k]; c[j][k] = sm[1][ty + j][tz + k]; }
  __syncthreads();

  for (int x = x0; x < x1; x++) {
    const int slot = (x - x0) & 1;
    stash(slot, nxt);
    if (x + 1 < x1) fetch(x + 3, nxt);
    __syncthreads();
    #pragma unroll
    for (int j = 0; j < 3; j++)
      #pragma unroll
      for (int k = 0; k < 3; k++) p[j][k] = sm[slot][ty + j][tz + k];

    if (active) {
      const i64 site = (i64)(x + L.h[0])*L.s0 + (i64)(i1 + L.h[1])*L.s1 + L.z0 + L.h[2] + i2;
      grad27_site(L, m, c, p, site, grad, delsq);
    }
    #pragma unroll
    for (int j = 0; j < 3; j++)
      #pragma unroll
      for (int k = 0; k < 3; k++) { m[j][k] = c[j][k]; c[j][k] = p[j][k]; }
  }
}

// ---------------------------------------------------------------------------
// viscosity_model arrhenius (visc_update_kernel, src/visc_arrhenius.c:185-220):
//   eta = eta_minus^((1 - phi/phistar)/2) eta_plus^((1 + phi/phistar)/2), interior sites.
// 8 B read + 8 B written per site; the collision that follows reads eta back
// (src/collision.c:397). Device pow is within 2 ulp of the host's.
// ---------------------------------------------------------------------------

__global__ void __launch_bounds__(256)
visc_arrhenius_kernel(const Layout L, const double * __restrict__ phi, double * __restrict__ eta,
		      const double eta_plus, const double eta_minus, const double phistar) {
  const int i2 = blockIdx.x*blockDim.x + threadIdx.x;
  const int i1 = blockIdx.y*blockDim.y + threadIdx.y;
  const int i0 = blockIdx.z;
  if (i2 >= L.n[2] || i1 >= L.n[1]) return;
  const i64 site = (i64)(i0 + L.h[0])*L.s0 + (i64)(i1 + L.h[1])*L.s1 + L.z0 + L.h[2] + i2;
  const double phi0 = __ldg(phi + site)/phistar;
  const double etaminus = pow(eta_minus, 0.5*(1.0 - phi0));
  const double etaplus  = pow(eta_plus,  0.5*(1.0 + phi0));
  eta[site] = etaminus*etaplus;
}

// ---------------------------------------------------------------------------
// fills (hydro_field_set, src/hydro.c:381-398) and uniform-plane helpers
// ---------------------------------------------------------------------------

__global__ void fill_kernel(double * __restrict__ data, i64 n, i64 plane, int nf,
			    double v0, double v1, double v2) {
  const i64 i = (i64) blockIdx.x*blockDim.x + threadIdx.x;
  if (i >= n) return;
  data[i] = v0;
  if (nf > 1) data[plane + i] = v1;
  if (nf > 2) data[2*plane + i] = v2;
}

// ---------------------------------------------------------------------------
// Momentum handed to the walls by one bounce-back (the fnet tally of
// wall_bbl_kernel, src/wall.c:1073-1081,1096-1106): for every link (fluid site i on
// a wall face, velocity p pointing into a wall layer)
//   fnet += (2 f_p(i) - x_p - 2 w_p) c_p,   x_p = 2 rcs2 w_p rho0 (c_p . u_wall)
// from the post-collision populations. blockIdx.z walks the wall faces; a site on several
// faces (edges, corners) is handled by the first face in the order 0lo 0hi 1lo 1hi 2lo 2hi.
// ---------------------------------------------------------------------------

struct WallTab {
  Layout L;
  int nvel;
  int wall_lo[3], wall_hi[3];     // device axes
  signed char dcv[27][3];         // device-axis components of c_p
  signed char pcv[27][3];         // physical components of c_p
  double wv[27];
  double xout[27];                // x_p for the outgoing link p
  int slip;                       // partial slip (no wall velocity then)
  double s_lo[3], s_hi[3];        // device axes
  signed char qflip[27][8];       // velocity index with the masked device-axis components reversed
  int wrap[3];                    // device axis periodic within this rank (tangential steps wrap)
  int nface;                      // wall faces this rank holds; blockIdx.z picks one
  int face_axis[6], face_hi[6];
};

__global__ void __launch_bounds__(128)
wall_fnet_kernel(const WallTab W, const double * __restrict__ f, double * __restrict__ fnet) {
  const Layout & L = W.L;
  const int d = W.face_axis[blockIdx.z];
  const int face_hi = W.face_hi[blockIdx.z];
  const int da = (d == 0) ? 1 : 0, db = (d == 2) ? 1 : 2;      // the two in-face axes, db fastest
  const int ib = blockIdx.x*blockDim.x + threadIdx.x;
  const int ia = blockIdx.y;
  double acc[3] = {0.0, 0.0, 0.0};
  if (ib < L.n[db] && ia < L.n[da]) {
    int ic[3];
    ic[d] = face_hi ? L.n[d] - 1 : 0;
    ic[da] = ia; ic[db] = ib;
    bool lo[3], hi[3];
    bool earlier = false;
    #pragma unroll
    for (int e = 0; e < 3; e++) {
      lo[e] = W.wall_lo[e] && ic[e] == 0;
      hi[e] = W.wall_hi[e] && ic[e] == L.n[e] - 1;
      if (e < d && (lo[e] || hi[e])) earlier = true;
    }
    if (face_hi && lo[d]) earlier = true;                    // n = 1: both faces are this site
    if (!earlier) {
      const i64 site = (i64)(ic[0] + L.h[0])*L.s0 + (i64)(ic[1] + L.h[1])*L.s1 + L.z0 + L.h[2] + ic[2];
      for (int p = 1; p < W.nvel; p++) {
	bool link = false;
	#pragma unroll
	for (int e = 0; e < 3; e++) {
	  if (W.dcv[p][e] == -1 && lo[e]) link = true;
	  if (W.dcv[p][e] ==  1 && hi[e]) link = true;
	}
	if (link && !W.slip) {
	  const double fp = __ldg(f + (i64) p*L.nsd + site);
	  const double force = 2.0*fp - W.xout[p];
	  #pragma unroll
	  for (int a = 0; a < 3; a++) acc[a] += (force - 2.0*W.wv[p])*W.pcv[p][a];
	}
	if (link && W.slip) {
	  // src/wall.c:1166-1192: the bounced share hands over its full momentum, the
	  // slipping share only the component along the wall normal
	  bool nrm[3];
	  int nn = 0, mask = 0;
	  bool tang = false;
	  double sl = 0.0;
	  i64 koff = 0;
	  #pragma unroll
	  for (int e = 0; e < 3; e++) {
	    nrm[e] = (W.dcv[p][e] == -1 && lo[e]) || (W.dcv[p][e] == 1 && hi[e]);
	    if (nrm[e]) { nn++; sl += (W.dcv[p][e] == -1) ? W.s_lo[e] : W.s_hi[e]; }
	    else if (W.dcv[p][e] != 0) {
	      tang = true;
	      mask |= (1 << e);
	      // tangential step; a direction periodic within the rank wraps around
	      int c = ic[e] + W.dcv[p][e];
	      if (W.wrap[e]) c = (c + L.n[e]) % L.n[e];
	      koff += (i64)(c - ic[e])*((e == 0) ? L.s0 : ((e == 1) ? L.s1 : 1));
	    }
	  }
	  if (nn == 2) sl *= 0.5;
	  if (!tang || nn == 3) sl = 0.0;
	  const double fi = __ldg(f + (i64) p*L.nsd + site);
	  // q: c_p with its tangential components reversed
	  const int q = (sl != 0.0) ? W.qflip[p][mask] : p;
	  const double fk = (sl != 0.0) ? __ldg(f + (i64) q*L.nsd + site + koff) : 0.0;
	  #pragma unroll
	  for (int a = 0; a < 3; a++) {
	    acc[a] += 2.0*(1.0 - sl)*(fi - W.wv[p])*W.pcv[p][a];
	    const double w = -(W.pcv[p][a] + W.pcv[q][a])/2;
	    acc[a] += 2.0*w*w*sl*(fk - W.wv[q])*W.pcv[q][a];
	  }
	}
      }
    }
  }
  #pragma unroll
  for (int a = 0; a < 3; a++) {
    double v = acc[a];
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0 && v != 0.0) atomicAdd(fnet + a, v);
  }
}

// ---------------------------------------------------------------------------
// Lees-Edwards planes: lb_le_apply_boundary_conditions (src/model_le.c:57-390), which
// the reference runs on the HOST between two full-lattice copies every step. Here two
// small kernels touch only the two x-planes next to each Lees-Edwards plane:
//   le_reproject_kernel   the plane-crossing populations of a site, corrected for the
//                         velocity jump through the equilibrium projection of the change
//                         in momentum and stress (src/model_le.c:113-208), go to a buffer;
//   le_interp_kernel      they come back displaced along y by the distance the plane has
//                         slid, linearly interpolated (src/model_le.c:220-390).
// One thread per (plane side, site of the x-plane).
// ---------------------------------------------------------------------------

#define B200LB_LE_MAXPLANE 32

struct LeTab {
  Layout L;
  int nvel, ndist;
  signed char pcv[27][3];         // physical components of c_p
  double wv[27];
  int dx, dy;                     // device axes of physical x and y
  int da, db;                     // the two device axes spanning an x-plane, db fastest
  int nside;                      // 2 x local planes
  int ic[2*B200LB_LE_MAXPLANE];   // interior (0-based) device coordinate along dx of each plane side
  int cx[2*B200LB_LE_MAXPLANE];   // +1 below a plane, -1 above it: which populations cross
  double duy[2*B200LB_LE_MAXPLANE];
  int jdy[2*B200LB_LE_MAXPLANE];  // integer part of the displacement
  double fr[2*B200LB_LE_MAXPLANE];
  int nprop;                      // populations with c_x = +1 (= those with -1)
};

__device__ __forceinline__ i64 le_site(const LeTab & T, int side, int ia, int ib) {
  const Layout & L = T.L;
  int c[3];
  c[T.dx] = T.ic[side] + L.h[T.dx];
  c[T.da] = ia + L.h[T.da];
  c[T.db] = ib + L.h[T.db];
  return (i64) c[0]*L.s0 + (i64) c[1]*L.s1 + L.z0 + c[2];
}

__global__ void __launch_bounds__(128)
le_reproject_kernel(const LeTab T, const double * __restrict__ f, double * __restrict__ tmp) {
  const Layout & L = T.L;
  const int ib = blockIdx.x*blockDim.x + threadIdx.x;
  const int ia = blockIdx.y;
  const int side = blockIdx.z;
  if (ib >= L.n[T.db] || ia >= L.n[T.da]) return;
  const i64 site = le_site(T, side, ia, ib);
  const i64 nps = (i64) L.n[T.da]*L.n[T.db];
  const i64 ps = (i64) ia*L.n[T.db] + ib;
  const double cs2 = 1.0/3.0, rcs2 = 3.0;
  const double du[3] = {0.0, T.duy[side], 0.0};
  for (int n = 0; n < T.ndist; n++) {
    const double * fs = f + (i64) n*T.nvel*L.nsd + site;
    double rho = 0.0, g[3] = {0.0, 0.0, 0.0};
    for (int p = 0; p < T.nvel; p++) {
      const double fp = fs[(i64) p*L.nsd];
      rho += fp;
      for (int a = 0; a < 3; a++) g[a] += T.pcv[p][a]*fp;
    }
    double ds[3][3];
    for (int a = 0; a < 3; a++)
      for (int b = 0; b < 3; b++) ds[a][b] = g[a]*du[b] + du[a]*g[b] + rho*du[a]*du[b];
    int k = 0;
    for (int p = 1; p < T.nvel; p++) {
      if (T.pcv[p][0] != T.cx[side]) continue;
      const double udotc = du[1]*T.pcv[p][1];
      double sdotq = 0.0;
      for (int a = 0; a < 3; a++)
	for (int b = 0; b < 3; b++)
	  sdotq += ds[a][b]*(T.pcv[p][a]*T.pcv[p][b] - ((a == b) ? cs2 : 0.0));
      const double fnew = fs[(i64) p*L.nsd] + T.wv[p]*(rho*udotc*rcs2 + 0.5*sdotq*rcs2*rcs2);
      tmp[(((i64) side*T.ndist + n)*T.nprop + k)*nps + ps] = fnew;
      k++;
    }
  }
}

__global__ void __launch_bounds__(128)
le_interp_kernel(const LeTab T, const double * __restrict__ tmp, double * __restrict__ f) {
  const Layout & L = T.L;
  const int ib = blockIdx.x*blockDim.x + threadIdx.x;
  const int ia = blockIdx.y;
  const int side = blockIdx.z;
  if (ib >= L.n[T.db] || ia >= L.n[T.da]) return;
  const i64 site = le_site(T, side, ia, ib);
  const i64 nps = (i64) L.n[T.da]*L.n[T.db];
  // this site's 0-based y coordinate and the two it interpolates between
  const int ny = L.n[T.dy];
  const int jc = (T.dy == T.da) ? ia : ib;
  const int j1 = ((jc + T.jdy[side]) % ny + 2*ny) % ny;
  const int j2 = (j1 + 1) % ny;
  const i64 ps1 = (T.dy == T.da) ? (i64) j1*L.n[T.db] + ib : (i64) ia*L.n[T.db] + j1;
  const i64 ps2 = (T.dy == T.da) ? (i64) j2*L.n[T.db] + ib : (i64) ia*L.n[T.db] + j2;
  const double fr = T.fr[side];
  for (int n = 0; n < T.ndist; n++) {
    int k = 0;
    for (int p = 1; p < T.nvel; p++) {
      if (T.pcv[p][0] != T.cx[side]) continue;
      const double * t0 = tmp + (((i64) side*T.ndist + n)*T.nprop + k)*nps;
      f[(i64)(n*T.nvel + p)*L.nsd + site] = (1.0 - fr)*t0[ps1] + fr*t0[ps2];
      k++;
    }
  }
}

// sum, sum of squares, min, max of one field component over interior sites
__global__ void __launch_bounds__(256)
field_stats_kernel(const Layout L, const double * __restrict__ data, double * __restrict__ partial) {
  const int i2 = blockIdx.x*blockDim.x + threadIdx.x;
  const int i1 = blockIdx.y*blockDim.y + threadIdx.y;
  const int i0 = blockIdx.z;
  double q[8] = {0.0, 0.0, 1.0e300, -1.0e300, 0.0, 0.0, 0.0, 0.0};
  if (i2 < L.n[2] && i1 < L.n[1]) {
    const i64 site = (i64)(i0 + L.h[0])*L.s0 + (i64)(i1 + L.h[1])*L.s1 + L.z0 + L.h[2] + i2;
    const double v = __ldg(data + site);
    q[0] = v; q[1] = v*v; q[2] = v; q[3] = v;
  }
  const i64 b = ((i64) blockIdx.z*gridDim.y + blockIdx.y)*gridDim.x + blockIdx.x;
  block_reduce8(q, partial + 8*b);
}

// Folding the per-block partials of a sweep (8 values per block: two sums, min, max, four
// sums). The arithmetic is that of ONE block of 256 threads in which thread t takes partials
// t, t + 256, ... in that order and block_reduce8 folds the 256 results -- that order fixes
// the bits of the sums. Done literally, every turn of a thread waits for one memory latency:
// 119 us for the 32768 partials of a 256^3 sweep (profiles/r2_ncu_traffic.md), which was
// most of what b200lb_step_diag or a statistics step cost, and no more threads or blocks
// help a dependent chain. The 256 x 8 chains are independent of each other, though:
// stats_chain_kernel gives 64 of them to each of 32 blocks, which bring the partials in
// through shared memory 64 turns at a time (every thread loads, nothing waits on an
// addition) and then add them up in the same order as before; stats_fold_kernel is the old
// tail. Same bits, two short launches.
#define STATS_CHAIN_BLOCKS 32     // x 8 threads of the one-block form each
#define STATS_CHAIN_TURNS  64     // turns staged at a time: 64 x 64 values = 32 KB

__device__ __forceinline__ double stats_combine(const int k, const double q, const double v) {
  return (k == 2) ? fmin(q, v) : ((k == 3) ? fmax(q, v) : q + v);
}

// chain[8 t + k], t < 256: what thread t of the one-block form holds for value k
__global__ void __launch_bounds__(256)
stats_chain_kernel(const double * __restrict__ partial, const i64 nblock, double * __restrict__ chain) {
  __shared__ double tile[STATS_CHAIN_TURNS][64];
  const int tid = threadIdx.x;
  const int t0 = 8*blockIdx.x;                       // first of this block's eight threads
  const int l = (tid & 63) >> 3, k = tid & 7;         // the chain thread tid < 64 adds up
  double q = (k == 2) ? 1.0e300 : ((k == 3) ? -1.0e300 : 0.0);
  // turns of thread t0 + l: partials t0 + l + 256 i < nblock
  const i64 mine = (nblock > t0 + l) ? (nblock - (t0 + l) + 255)/256 : 0;
  const i64 nturn = (nblock + 255)/256;
  for (i64 i0 = 0; i0 < nturn; i0 += STATS_CHAIN_TURNS) {
    #pragma unroll
    for (int e = tid; e < STATS_CHAIN_TURNS*64; e += 256) {
      const int turn = e >> 6, lk = e & 63;
      const i64 b = t0 + (lk >> 3) + 256*(i0 + turn);
      tile[turn][lk] = (b < nblock) ? __ldg(partial + 8*b + (lk & 7)) : 0.0;
    }
    __syncthreads();
    if (tid < 64) {
      const i64 left = mine - i0;
      const int n = (left < STATS_CHAIN_TURNS) ? (int) ((left > 0) ? left : 0) : STATS_CHAIN_TURNS;
      for (int turn = 0; turn < n; turn++) q = stats_combine(k, q, tile[turn][tid]);
    }
    __syncthreads();
  }
  if (tid < 64) chain[8*(t0 + l) + k] = q;
}

__global__ void __launch_bounds__(256)
stats_fold_kernel(const double * __restrict__ chain, double * __restrict__ out) {
  double q[8];
  #pragma unroll
  for (int k = 0; k < 8; k++) q[k] = chain[8*threadIdx.x + k];
  block_reduce8(q, out);
}

// chain: 2048 doubles of scratch
static inline cudaError_t stats_final(const double * partial, i64 nblock, double * out, double * chain,
				      cudaStream_t st) {
  stats_chain_kernel<<<STATS_CHAIN_BLOCKS, 256, 0, st>>>(partial, nblock, chain);
  stats_fold_kernel<<<1, 256, 0, st>>>(chain, out);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// Layout conversion between a staged piece of a HOST array and the device
// planes (the job lb_memcpy / hydro_memcpy / field_memcpy do in the reference,
// src/model.c:228-270). The host site index is cs_index over the physical
// halo-inclusive box (src/coords.c:617-631); hostdim = physical nall.
//
//   AOS piece: stage[(hs - hs0)*nel + e],  hs in [hs0, hs1)
//   SOA piece: one plane e0:  stage[hs - hs0]
// TO_DEVICE = 1 host -> device, 0 device -> host.
// For a 2-d model every host z-plane maps onto the single device plane: upload
// takes kc = 1, download replicates (a periodic z halo of extent 1).
// ---------------------------------------------------------------------------

struct XferParams {
  Layout L;
  int hostdim[3];     // physical nall (x,y,z)
  int dax[3];         // device axis -> physical axis
  double * dev;       // device base (plane 0)
  double * stage;
  i64 hs0, hs1;
  int nel;            // planes per site in an AOS piece
  int e0;             // first device plane (SOA piece: the plane)
  int aos;
  int hoff[3];        // added to the decoded host coordinates: 0 = the host index space is
                      // the halo-inclusive box, 1 = interior sites only (file records)
  int vvl;            // AOS piece in blocks of vvl sites (the reference's -DADDR_AOSOA -DVVL=vvl,
                      // src/memory.h:260-284): element e of site s sits at ((s/vvl)*nel + e)*vvl + s%vvl
  int half;           // SOA piece of 32-bit words: 1 = low, 2 = high half of the 64-bit device
                      // element, stage holds unsigned ints (noise state); 0 = doubles
};

template <int TO_DEVICE>
__global__ void __launch_bounds__(256)
xfer_kernel(const XferParams X) {
  const i64 hs = X.hs0 + (i64) blockIdx.x*blockDim.x + threadIdx.x;
  if (hs >= X.hs1) return;
  const Layout & L = X.L;
  int pc[3];
  pc[2] = (int) (hs % X.hostdim[2]) + X.hoff[2];
  const i64 r = hs / X.hostdim[2];
  pc[1] = (int) (r % X.hostdim[1]) + X.hoff[1];
  pc[0] = (int) (r / X.hostdim[1]) + X.hoff[0];
  int dc[3];
  bool primary = true;     // false for replicated planes of the unit axis
  #pragma unroll
  for (int d = 0; d < 3; d++) {
    int c = pc[X.dax[d]];
    if (L.h[d] == 0) { if (c != 1) primary = false; c = 0; }
    dc[d] = c;
  }
  const i64 site = (i64) dc[0]*L.s0 + (i64) dc[1]*L.s1 + L.z0 + dc[2];
  if (TO_DEVICE && !primary) return;
  if (X.aos) {
    // pieces start on a block boundary, so the block arithmetic is local to the piece
    const i64 l = hs - X.hs0;
    const i64 es = (X.vvl > 1) ? X.vvl : 1;
    double * st = (X.vvl > 1) ? X.stage + (l/X.vvl)*X.nel*X.vvl + l%X.vvl : X.stage + l*X.nel;
    for (int e = 0; e < X.nel; e++) {
      if (TO_DEVICE) X.dev[(i64)(X.e0 + e)*L.nsd + site] = st[e*es];
      else st[e*es] = X.dev[(i64)(X.e0 + e)*L.nsd + site];
    }
  }
  else if (X.half) {
    unsigned int * dv = reinterpret_cast<unsigned int *>(X.dev + (i64) X.e0*L.nsd + site) + (X.half - 1);
    unsigned int * st = reinterpret_cast<unsigned int *>(X.stage) + (hs - X.hs0);
    if (TO_DEVICE) *dv = *st;
    else *st = *dv;
  }
  else {
    if (TO_DEVICE) X.dev[(i64) X.e0*L.nsd + site] = X.stage[hs - X.hs0];
    else X.stage[hs - X.hs0] = X.dev[(i64) X.e0*L.nsd + site];
  }
}

// ---------------------------------------------------------------------------
// noise_init (src/noise.c:136-233): decomposition-independent seeding of the per-site
// generators. A site's first state is seed0 + its GLOBAL periodic position (1-based);
// four outputs of KISS99 from there become the state it keeps. Halo sites included.
// State words are stored pairwise as 64-bit elements: plane 0 = (s0,s1), plane 1 = (s2,s3).
// ---------------------------------------------------------------------------

struct NoiseInit {
  Layout L;
  int dax[3];          // device axis -> physical axis
  int noffset[3];      // physical: global offset of this rank's interior
  int ntotal[3];       // physical: global extent
  unsigned int seed0[4];
};

__device__ __forceinline__ unsigned int noise_kiss99(unsigned int (&w)[4]) {
  w[0] = 69069u*w[0] + 1234567u;
  unsigned int x = w[1];
  x ^= x << 17; x ^= x >> 13; x ^= x << 5;
  w[1] = x;
  w[2] = 36969u*(w[2] & 65535u) + (w[2] >> 16);
  w[3] = 18000u*(w[3] & 65535u) + (w[3] >> 16);
  return w[1] + (w[0] ^ ((w[2] << 16) + w[3]));
}

__global__ void __launch_bounds__(128)
noise_init_kernel(const NoiseInit N, double * __restrict__ state) {
  const Layout & L = N.L;
  const int d2 = blockIdx.x*blockDim.x + threadIdx.x;
  const int d1 = blockIdx.y, d0 = blockIdx.z;
  if (d2 >= L.na[2]) return;
  const int dc[3] = {d0, d1, d2};
  unsigned int w[4] = {N.seed0[0], N.seed0[1], N.seed0[2], N.seed0[3]};
  #pragma unroll
  for (int d = 0; d < 3; d++) {
    const int a = N.dax[d];
    // local physical coordinate, 0 .. nlocal + 1; the unit axis of a 2-d lattice sits at 1
    const int c = (L.h[d] == 0) ? 1 : dc[d];
    int g = N.noffset[a] + c;
    if (g < 1) g += N.ntotal[a];
    if (g > N.ntotal[a]) g -= N.ntotal[a];
    w[a] += (unsigned int) g;
  }
  unsigned int o[4];
  #pragma unroll
  for (int k = 0; k < 4; k++) o[k] = noise_kiss99(w);
  const i64 site = (i64) d0*L.s0 + (i64) d1*L.s1 + L.z0 + d2;
  state[site]         = __hiloint2double((int) o[1], (int) o[0]);
  state[L.nsd + site] = __hiloint2double((int) o[3], (int) o[2]);
}

// ---------------------------------------------------------------------------
// Direct reduced halo: one message per neighbour direction (faces, edges and,
// for D3Q27, corners of the directions decomposed over ranks), each carrying only
// the populations whose velocity points into the receiver -- the message set of
// the reference's LB_HALO_OPENMP_REDUCED scheme (src/model.c:1161-1219), but
// packed by ONE kernel launch, sent in ONE NCCL group and unpacked by one launch:
// no x -> y -> z dependency chain. Directions that are periodic within the rank
// need nothing (the sweeps wrap), so regions span only interior sites there.
// ---------------------------------------------------------------------------

#define B200LB_MAXMSG 26
#define B200LB_MSGPOP 18

struct HaloMsg {
  int npop;
  int plane[B200LB_MSGPOP];   // n*nvel + p
  int slo[3];                 // first site of the send region (halo-inclusive device coords)
  int rlo[3];                 // first site of the receive region
  int dim[3];                 // region extent
  i64 nsite;                  // dim[0]*dim[1]*dim[2]
  i64 off;                    // offset of this message in the send / receive buffers
};

template <int UNPACK>
__global__ void __launch_bounds__(256)
halo_direct_kernel(const Layout L, const HaloMsg * __restrict__ msgs,
		   double * __restrict__ data, double * __restrict__ buf) {
  const HaloMsg & m = msgs[blockIdx.y];
  const i64 total = m.nsite*m.npop;
  for (i64 idx = (i64) blockIdx.x*blockDim.x + threadIdx.x; idx < total;
       idx += (i64) gridDim.x*blockDim.x) {
    const int k = (int) (idx / m.nsite);
    const i64 s = idx - (i64) k*m.nsite;
    const int a2 = (int) (s % m.dim[2]);
    const i64 r = s / m.dim[2];
    const int a1 = (int) (r % m.dim[1]);
    const int a0 = (int) (r / m.dim[1]);
    const int * o = UNPACK ? m.rlo : m.slo;
    const i64 site = (i64) (o[0] + a0)*L.s0 + (i64) (o[1] + a1)*L.s1 + L.z0 + o[2] + a2;
    double * p = data + (i64) m.plane[k]*L.nsd + site;
    if (UNPACK) *p = buf[m.off + idx];
    else buf[m.off + idx] = *p;
  }
}

// ---------------------------------------------------------------------------
// hydro_correct_momentum (src/hydro.c:829-895, kernels :904-1099): remove the net
// body force. Stage 1 sums the force over interior sites (warp shuffles, one partial
// per block, no atomics), stats_final_kernel folds the partials, the caller
// all-reduces over ranks, stage 2 adds -F_net/V to every interior site, reading the
// sum from device memory (no host round trip).
// ---------------------------------------------------------------------------

__global__ void __launch_bounds__(256)
force_sum_kernel(const Layout L, const double * __restrict__ force, double * __restrict__ partial) {
  const int i2 = blockIdx.x*blockDim.x + threadIdx.x;
  const int i1 = blockIdx.y*blockDim.y + threadIdx.y;
  const int i0 = blockIdx.z;
  double q[8] = {0.0, 0.0, 1.0e300, -1.0e300, 0.0, 0.0, 0.0, 0.0};
  if (i2 < L.n[2] && i1 < L.n[1]) {
    const i64 site = (i64)(i0 + L.h[0])*L.s0 + (i64)(i1 + L.h[1])*L.s1 + L.z0 + L.h[2] + i2;
    q[4] = force[site]; q[5] = force[L.nsd + site]; q[6] = force[2*L.nsd + site];
  }
  const i64 b = ((i64) blockIdx.z*gridDim.y + blockIdx.y)*gridDim.x + blockIdx.x;
  block_reduce8(q, partial + 8*b);
}

__global__ void __launch_bounds__(256)
force_correct_kernel(const Layout L, double * __restrict__ force,
		     const double * __restrict__ sums, const double rv) {
  const int i2 = blockIdx.x*blockDim.x + threadIdx.x;
  const int i1 = blockIdx.y*blockDim.y + threadIdx.y;
  const int i0 = blockIdx.z;
  if (i2 >= L.n[2] || i1 >= L.n[1]) return;
  const i64 site = (i64)(i0 + L.h[0])*L.s0 + (i64)(i1 + L.h[1])*L.s1 + L.z0 + L.h[2] + i2;
  #pragma unroll
  for (int a = 0; a < 3; a++) force[a*L.nsd + site] += -sums[4 + a]*rv;
}

// ---------------------------------------------------------------------------
// Peer-memory halo synchronisation. Every rank runs the same sequence
//   wait(k) -> sweep k+1 (pushes into the neighbours' halo sites) -> signal(k+1)
// where signal writes the step number into a slot of each neighbour's flag array
// (peer memory) and wait spins on the local array until every neighbour has
// reached that step. One rule covers both hazards: a neighbour's pushed values
// are complete before they are read, and a neighbour has finished reading the
// buffer this rank is about to push into. The spin has a watchdog (never hangs, and never
// lets a sweep through: see p2p_sync_kernel).
// ---------------------------------------------------------------------------

struct P2PSync {
  unsigned long long * flags;            // local: [27] per direction + [27] = error
  unsigned long long * peer_flags[27];   // neighbour's flag array per direction, or null
};

// signal(step) and wait(step) in ONE launch: between two sweeps of the time loop nothing
// else sits on the stream, so publishing this rank's step and waiting for the neighbours'
// back to back is the same protocol with one launch (and one launch gap) less. Every
// rank signals before it waits: no cycle.
// The watchdog FAILS CLOSED: a neighbour that does not arrive within timeout_ns leaves the
// step in flags[27] (read by b200lb_sync for the message) and traps, which poisons the
// context -- every later CUDA call of this process fails, nothing is ever swept on
// incomplete halos.
__global__ void p2p_sync_kernel(const P2PSync S, const unsigned long long step,
				const unsigned long long timeout_ns, const int do_signal) {
  const int t = threadIdx.x;
  // the sweep launched behind this kernel may start now (programmatic dependent launch): its
  // interior needs nothing from here, its boundary threads wait for this kernel to complete
  asm volatile("griddepcontrol.launch_dependents;");
  if (t < 27 && S.peer_flags[t] != nullptr) {
    if (do_signal) {
      __threadfence_system();
      // the neighbour in direction t sees this rank in direction -t = slot 26 - t
      *((volatile unsigned long long *) (S.peer_flags[t] + (26 - t))) = step;
    }
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    while (*((volatile unsigned long long *) (S.flags + t)) < step) {
      __nanosleep(100);
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
      if (t1 - t0 > timeout_ns) {
	*((volatile unsigned long long *) (S.flags + 27)) = step;
	__threadfence_system();
	__trap();
      }
    }
  }
  __threadfence_system();
}

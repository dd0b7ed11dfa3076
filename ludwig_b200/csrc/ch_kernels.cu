// ch_kernels.cu -- the order parameter as a finite-difference field ("free_energy
// symmetric", src/ludwig.c:727-768): one kernel for the body force on the fluid
// (phi_force_calculation, src/phi_force.c:74) and the Cahn-Hilliard update of phi
// (phi_cahn_hilliard, src/phi_cahn_hilliard.c:206).
//
// The reference makes seven passes over the lattice for this: the chemical stress into
// a 9-component array one site into the halo (pth_kernel_v, src/phi_force_stress.c:260),
// its divergence (pth_force_fluid_kernel_v, src/phi_force_colloid.c:323), the advective
// fluxes into four arrays (src/advection.c:542/790/977), the diffusive fluxes on top
// (phi_ch_flux_mu1_kernel, src/phi_cahn_hilliard.c:349; _mu_ext :1404), the no-flux mask
// (src/advection_bcs.c) and the forward step (phi_ch_ufs_kernel :1025). Here one thread
// owns one site: it reads phi, grad phi, delsq phi and u at the site and its six
// neighbours (phi two sites out for the third-order scheme), evaluates stress,
// chemical potential and the six face fluxes in registers, and writes the force (24 B)
// and the new phi (8 B): 64 B read + 32 B written per site, nothing else touches HBM.
//
// Conservation. A face flux is evaluated by both sites that share the face, by the same
// function (ch_face) from the same operands, every operation in it an explicit
// intrinsic: the two values are identical bit for bit, whatever the compiler does
// around them, and phi is conserved to round-off with no flux arrays. (The reference
// stores "west" and "east" fluxes separately and relies on the same identity.)
//
// This file is compiled twice (csrc/Makefile):
//   CH_EXACT = 1, -fmad=false   every operation rounded on its own, in the reference's
//       order: force and phi agree with the reference BIT FOR BIT (option "ch_exact").
//       348 FP64 instructions per site: the FP64 pipe, not HBM, is the floor (ncu:
//       profiles/r1_ncu_ch_kernel_exact_256.txt).
//   CH_EXACT = 0   the default: fused multiply-add, and the divergence of the stress
//       taken as -(1/2) sum_b [P_ab(r + b) - P_ab(r - b)] -- the central value, which the
//       reference adds and subtracts (src/phi_force_colloid.c:386-466), never formed.
//       Agrees with the exact build to a few ulp.

#include "ch_kernels.cuh"

#ifndef CH_EXACT
#define CH_EXACT 1
#endif

#define CH_TZ 64
#define CH_TY 4
#define CH_NX 8      // default x-planes per block (ChParams.nx): neighbours along x are re-read through L1

// Everything below has internal linkage: the two builds of this file live in one library
// and must not share a host stub (the kernel also carries CH_EXACT in its name, which is
// what a profiler shows: ch_kernel<1, ...> is the exact build).
namespace {

struct Symm { double a, b, kappa; };

// a*b + c: one rounding in the fast build, two (as the reference's C) in the exact one.
// Used with __dmul_rn / __dadd_rn, which no compiler pass fuses or reorders.
__device__ __forceinline__ double ch_mad(double a, double b, double c) {
#if CH_EXACT
  return __dadd_rn(__dmul_rn(a, b), c);
#else
  return __fma_rn(a, b, c);
#endif
}

// fe_symm_mu, src/symmetric.c:307-320: a phi + b phi phi phi - kappa delsq
__device__ __forceinline__ double ch_mu(const Symm & s, double phi, double delsq) {
  const double t1 = __dmul_rn(s.a, phi);
  const double t2 = __dmul_rn(__dmul_rn(s.b, phi), phi);
  return ch_mad(-s.kappa, delsq, ch_mad(t2, phi, t1));
}

// the isotropic part of fe_symm_str_v, src/symmetric.c:404-407
__device__ __forceinline__ double ch_p0(const Symm & s, double phi, double delsq, const double g[3]) {
  return 0.5*s.a*phi*phi + 0.75*s.b*phi*phi*phi*phi - s.kappa*phi*delsq
    - 0.5*s.kappa*(g[0]*g[0] + g[1]*g[1] + g[2]*g[2]);
}

// P_ab = p0 delta_ab + kappa d_a phi d_b phi, src/symmetric.c:409-413
__device__ __forceinline__ double ch_pab(const Symm & s, double p0, const double g[3], int ia, int ib) {
  return ((ia == ib) ? p0 : 0.0) + s.kappa*g[ia]*g[ib];
}

// The flux through the face between site A and the next site B along an axis:
// advection_x (src/advection.c: first order upwind :542, centred :790, the three-point
// upwind-biased stencil :977, four-point centred :1188), then - M (mu_B - mu_A) (phi_ch_flux_mu1_kernel,
// src/phi_cahn_hilliard.c:349) and - M grad mu^ext (:1404). Am and Bp are the sites
// beyond A and B. The reference computes this twice per face -- as the "east"/y/z flux
// of A and as the "west" flux of B, or as the y/z flux fy[index - ys] read back -- with
// the operands in these roles; sums are commutative, so the one function serves both.
template <int ORDER>
__device__ __forceinline__ double ch_face(double uA, double uB, double phAm, double phA,
					  double phB, double phBp, double muA, double muB,
					  double M, double Mgmu) {
  double adv;
  if (ORDER == 1) {
    const double u = __dmul_rn(0.5, __dadd_rn(uA, uB));
    adv = __dmul_rn(u, (u > 0.0) ? phA : phB);
  }
  else if (ORDER == 2) {
    adv = __dmul_rn(__dmul_rn(__dmul_rn(0.5, __dadd_rn(uA, uB)), 0.5), __dadd_rn(phA, phB));
  }
  else if (ORDER == 4) {
    // advection_le_4th, src/advection.c:1188-1290 (a host loop in the reference, behind a
    // copy of the fluxes to the host and back): four-point centred interpolation,
    // u (- phi_Am/16 + 9 phi_A/16 + 9 phi_B/16 - phi_Bp/16), summed in that order
    const double b1 = (1.0/16.0), b2 = (9.0/16.0);
    const double u = __dmul_rn(0.5, __dadd_rn(uA, uB));
    adv = __dmul_rn(u, ch_mad(-b1, phBp, ch_mad(b2, phB, ch_mad(b2, phA, __dmul_rn(-b1, phAm)))));
  }
  else {
    const double a1 = -0.213933, a2 = 0.927865, a3 = 0.286067;
    const double u = __dmul_rn(0.5, __dadd_rn(uA, uB));
    const bool up = (u > 0.0);
    const double fd1 = up ? phAm : phBp, fd2 = up ? phA : phB, fd3 = up ? phB : phA;
    adv = __dmul_rn(u, ch_mad(a3, fd3, ch_mad(a2, fd2, __dmul_rn(a1, fd1))));
  }
  return __dadd_rn(ch_mad(-M, __dadd_rn(muB, -muA), adv), -Mgmu);
}

// offset (in sites) from halo-inclusive coordinate c to c + o along an axis of interior
// extent n, wrapping where the axis is periodic within the rank
__device__ __forceinline__ int ch_off(int c, int n, int wrap, int o) {
  int t = c + o;
  if (wrap) { if (t < 1) t += n; else if (t > n) t -= n; }
  return t - c;
}

// METHOD: 0 = no force pass, CH_FORCE_STRESS_DIVERGENCE, CH_FORCE_PHI_GRADMU
// ORDER:  0 = no update pass, 1..4 = advection scheme of the Cahn-Hilliard update
template <int EXACT, int METHOD, int ORDER>
__global__ void __launch_bounds__(CH_TZ*CH_TY, EXACT ? 2 : 3)
ch_kernel(const ChParams P) {

  const Layout & L = P.L;
  const int i2 = blockIdx.x*CH_TZ + threadIdx.x;
  const int i1 = blockIdx.y*CH_TY + threadIdx.y;
  if (i2 >= L.n[2] || i1 >= L.n[1]) return;
  const int x0 = blockIdx.z*P.nx;
  const int x1 = min(x0 + P.nx, L.n[0]);

  const Symm S = {P.A, P.B, P.kappa};
  const int c1 = i1 + L.h[1], c2 = i2 + L.h[2];
  const i64 col = (i64) c1*L.s1 + L.z0 + c2;

  // neighbour offsets in doubles: [axis][0..3] = -2, -1, +1, +2
  i64 d[3][4];
  {
    const int o[4] = {-2, -1, 1, 2};
    #pragma unroll
    for (int k = 0; k < 4; k++) {
      d[1][k] = (i64) ch_off(c1, L.n[1], P.wrap[1], o[k])*L.s1;
      d[2][k] = (i64) ch_off(c2, L.n[2], P.wrap[2], o[k]);
    }
  }

  const double * __restrict__ phi = P.phi;
  const double * __restrict__ delsq = P.delsq;
  const double * __restrict__ grad = P.grad;
  const double * __restrict__ u = P.u;
  const i64 nsd = L.nsd;

  for (int x = x0; x < x1; x++) {
    const int c0 = x + L.h[0];
    {
      const int o[4] = {-2, -1, 1, 2};
      #pragma unroll
      for (int k = 0; k < 4; k++) d[0][k] = (i64) ch_off(c0, L.n[0], P.wrap[0], o[k])*L.s0;
    }
    const i64 site = (i64) c0*L.s0 + col;

    const double ph0 = __ldg(phi + site);
    const double dq0 = __ldg(delsq + site);
    double phm[3], php[3], dqm[3], dqp[3];
    #pragma unroll
    for (int a = 0; a < 3; a++) {
      phm[a] = __ldg(phi + site + d[a][1]);
      php[a] = __ldg(phi + site + d[a][2]);
      dqm[a] = __ldg(delsq + site + d[a][1]);
      dqp[a] = __ldg(delsq + site + d[a][2]);
    }

    if (METHOD == CH_FORCE_STRESS_DIVERGENCE) {
      // F_a = - d_b P_ab, the stress interpolated to the faces:
      // pth_force_fluid_kernel_v, src/phi_force_colloid.c:366-468
      double g1[3], f[3];
#if CH_EXACT
      double g0[3];
      #pragma unroll
      for (int a = 0; a < 3; a++) g0[a] = __ldg(grad + a*nsd + site);
      const double p00 = ch_p0(S, ph0, dq0, g0);
      #pragma unroll
      for (int b = 0; b < 3; b++) {
	#pragma unroll
	for (int a = 0; a < 3; a++) g1[a] = __ldg(grad + a*nsd + site + d[b][2]);
	const double p0p = ch_p0(S, php[b], dqp[b], g1);
	#pragma unroll
	for (int a = 0; a < 3; a++) {
	  const double t = 0.5*(ch_pab(S, p0p, g1, a, b) + ch_pab(S, p00, g0, a, b));
	  if (b == 0) f[a] = -t;
	  else f[a] -= t;
	}
	#pragma unroll
	for (int a = 0; a < 3; a++) g1[a] = __ldg(grad + a*nsd + site + d[b][1]);
	const double p0m = ch_p0(S, phm[b], dqm[b], g1);
	#pragma unroll
	for (int a = 0; a < 3; a++) {
	  f[a] += 0.5*(ch_pab(S, p0m, g1, a, b) + ch_pab(S, p00, g0, a, b));
	}
      }
#else
      // the central stress cancels between the two faces of a direction:
      // F_a = -(1/2) sum_b [P_ab(r + b) - P_ab(r - b)]
      f[0] = f[1] = f[2] = 0.0;
      #pragma unroll
      for (int b = 0; b < 3; b++) {
	#pragma unroll
	for (int a = 0; a < 3; a++) g1[a] = __ldg(grad + a*nsd + site + d[b][2]);
	const double kgp = S.kappa*g1[b];
	f[b] += ch_p0(S, php[b], dqp[b], g1);
	#pragma unroll
	for (int a = 0; a < 3; a++) f[a] += kgp*g1[a];
	#pragma unroll
	for (int a = 0; a < 3; a++) g1[a] = __ldg(grad + a*nsd + site + d[b][1]);
	const double kgm = S.kappa*g1[b];
	f[b] -= ch_p0(S, phm[b], dqm[b], g1);
	#pragma unroll
	for (int a = 0; a < 3; a++) f[a] -= kgm*g1[a];
      }
      #pragma unroll
      for (int a = 0; a < 3; a++) f[a] *= -0.5;
#endif
      #pragma unroll
      for (int a = 0; a < 3; a++) {
	const double base = P.accumulate ? P.force[a*nsd + site] : P.fconst[a];
	P.force[a*nsd + site] = base + f[a];
      }
    }

    if (METHOD == CH_FORCE_PHI_GRADMU) {
      // F_a = - phi d_a mu, then - phi grad_a mu^ext:
      // phi_grad_mu_fluid_kernel / _external_kernel, src/phi_grad_mu.c:294-380,528-560
      #pragma unroll
      for (int a = 0; a < 3; a++) {
	const double mum1 = ch_mu(S, phm[a], dqm[a]);
	const double mup1 = ch_mu(S, php[a], dqp[a]);
	double f = 0.0;
	f += -ph0*0.5*(mup1 - mum1);
	const double base = P.accumulate ? P.force[a*nsd + site] : P.fconst[a];
	P.force[a*nsd + site] = (base + f) + (-ph0*P.gmu[a]);
      }
    }

    if (ORDER > 0) {
      // the six faces of this site, then phi_ch_ufs_kernel (src/phi_cahn_hilliard.c:1025)
      const double M = P.mobility;
      const double mu0 = ch_mu(S, ph0, dq0);
      double flo[3], fhi[3];
      #pragma unroll
      for (int a = 0; a < 3; a++) {
	const double u0 = __ldg(u + a*nsd + site);
	const double um = __ldg(u + a*nsd + site + d[a][1]);
	const double up = __ldg(u + a*nsd + site + d[a][2]);
	const double phm2 = (ORDER >= 3) ? __ldg(phi + site + d[a][0]) : 0.0;
	const double php2 = (ORDER >= 3) ? __ldg(phi + site + d[a][3]) : 0.0;
	const double Mg = __dmul_rn(M, P.gmu[a]);
	flo[a] = ch_face<ORDER>(um, u0, phm2, phm[a], ph0, php[a], ch_mu(S, phm[a], dqm[a]), mu0, M, Mg);
	fhi[a] = ch_face<ORDER>(u0, up, phm[a], ph0, php[a], php2, mu0, ch_mu(S, php[a], dqp[a]), M, Mg);
      }
      // phi -= (+ fe - fw + fy - fy[index - ys] + wz*fz - wz*fz[index - 1]); wz = 0 for a
      // lattice one site thick (uniform branch)
      double dphi = __dadd_rn(fhi[0], -flo[0]);
      dphi = __dadd_rn(dphi, fhi[1]);
      dphi = __dadd_rn(dphi, -flo[1]);
      if (!P.skipz) {
	dphi = __dadd_rn(dphi, fhi[2]);
	dphi = __dadd_rn(dphi, -flo[2]);
      }
      P.phi_out[site] = __dadd_rn(ph0, -dphi);
    }
  }
}

// ---------------------------------------------------------------------------
// Lees-Edwards planes (see ch_kernels.cuh). What a site of layer l = -1 .. 2 around plane
// p (x = loc + l) sees at x = loc + r: a real plane, or -- looking across the plane --
// buffer b = r + 1 (lees_edw_ic_to_buff, src/leesedwards.c:1014-1047 with nhalo = 2).
// ---------------------------------------------------------------------------

struct LeView {
  const double * phi[5];      // x-planes of phi at r = l - 2 .. l + 2
  const double * dsq[3];      // delsq at r = l - 1 .. l + 1
  const double * g[3][3];     // [r][component] gradient planes, r = l - 1 .. l + 1
  const double * ux[3];       // u_x planes
};

__device__ __forceinline__ int le_wrapx(const ChParams & P, int x) {
  if (P.wrap[0]) { if (x < 1) x += P.L.n[0]; else if (x > P.L.n[0]) x -= P.L.n[0]; }
  return x;
}

__device__ __forceinline__ void le_view(const ChParams & P, const ChLeParams & Q, int p, int l, LeView & V) {
  const Layout & L = P.L;
  const i64 s0 = L.s0;
  #pragma unroll
  for (int o = -2; o <= 2; o++) {
    const int r = l + o;
    const bool across = (l <= 0) ? (r >= 1) : (r <= 0);
    if (across) V.phi[o + 2] = Q.bphi + ((i64) 4*p + (r + 1))*s0;
    else V.phi[o + 2] = P.phi + (i64) le_wrapx(P, Q.loc[p] + r)*s0;
    if (o >= -1 && o <= 1) {
      if (across) {
	const int e = 2*p + (r + 1) - 1;            // buffers b = 1, 2 only are ever asked for
	V.dsq[o + 1] = Q.bdelsq + (i64) e*s0;
	V.ux[o + 1] = Q.bux + (i64) e*s0;
	#pragma unroll
	for (int a = 0; a < 3; a++) V.g[o + 1][a] = Q.bgrad + ((i64) e*3 + a)*s0;
      }
      else {
	const i64 x = le_wrapx(P, Q.loc[p] + r);
	V.dsq[o + 1] = P.delsq + x*s0;
	V.ux[o + 1] = P.u + x*s0;
	#pragma unroll
	for (int a = 0; a < 3; a++) V.g[o + 1][a] = P.grad + a*L.nsd + x*s0;
      }
    }
  }
}

// P_ab e_x . at one site of a plane triple: the x column of the stress (ia = 0 .. 2)
__device__ __forceinline__ void le_stress_col(const Symm & S, const double * phi, const double * dsq,
					       const double * const g[3], i64 col, int ib, double (&out)[3]) {
  double gg[3];
  #pragma unroll
  for (int a = 0; a < 3; a++) gg[a] = __ldg(g[a] + col);
  const double p0 = ch_p0(S, __ldg(phi + col), __ldg(dsq + col), gg);
  #pragma unroll
  for (int a = 0; a < 3; a++) out[a] = ch_pab(S, p0, gg, a, ib);
}

// pass A: grid (z tiles, n_y, nplanes), block (128)
template <int EXACT, int ORDER>
__global__ void __launch_bounds__(128)
ch_le_pass_a(const ChParams P, const ChLeParams Q, const int do_force) {
  const Layout & L = P.L;
  const int p = blockIdx.z;
  const int c1 = blockIdx.y + 1;
  const int c2 = blockIdx.x*blockDim.x + threadIdx.x + 1;
  if (c2 > L.n[2]) return;
  const i64 col = (i64) c1*L.s1 + L.z0 + c2;
  const Symm S = {P.A, P.B, P.kappa};
  double * scr = Q.scr + (i64) 8*p*L.s0;
  const double M = P.mobility;
  const double Mg = __dmul_rn(M, P.gmu[0]);

  {
    // the layer below the plane: its east face, looking up into buffers b = 2, 3
    LeView V;
    le_view(P, Q, p, 0, V);
    if (do_force) {
      double p0[3], p1[3];
      le_stress_col(S, V.phi[2], V.dsq[1], V.g[1], col, 0, p0);
      le_stress_col(S, V.phi[3], V.dsq[2], V.g[2], col, 0, p1);
      #pragma unroll
      for (int a = 0; a < 3; a++) scr[(2 + a)*L.s0 + col] = 0.5*(p1[a] + p0[a]);
    }
    if (ORDER > 0) {
      const double phA = __ldg(V.phi[2] + col), phB = __ldg(V.phi[3] + col);
      scr[col] = ch_face<(ORDER > 0 ? ORDER : 1)>(__ldg(V.ux[1] + col), __ldg(V.ux[2] + col),
	  __ldg(V.phi[1] + col), phA, phB, __ldg(V.phi[4] + col),
	  ch_mu(S, phA, __ldg(V.dsq[1] + col)), ch_mu(S, phB, __ldg(V.dsq[2] + col)), M, Mg);
    }
  }
  {
    // the layer above: its west face, looking down into buffers b = 1, 0
    LeView V;
    le_view(P, Q, p, 1, V);
    if (do_force) {
      double p0[3], p1[3];
      le_stress_col(S, V.phi[2], V.dsq[1], V.g[1], col, 0, p0);
      le_stress_col(S, V.phi[1], V.dsq[0], V.g[0], col, 0, p1);
      #pragma unroll
      for (int a = 0; a < 3; a++) scr[(5 + a)*L.s0 + col] = 0.5*(p1[a] + p0[a]);
    }
    if (ORDER > 0) {
      const double phA = __ldg(V.phi[1] + col), phB = __ldg(V.phi[2] + col);
      scr[L.s0 + col] = ch_face<(ORDER > 0 ? ORDER : 1)>(__ldg(V.ux[0] + col), __ldg(V.ux[1] + col),
	  __ldg(V.phi[0] + col), phA, phB, __ldg(V.phi[3] + col),
	  ch_mu(S, phA, __ldg(V.dsq[0] + col)), ch_mu(S, phB, __ldg(V.dsq[1] + col)), M, Mg);
    }
  }
}

// plane sums of -fluxe(loc) + fluxw(loc + 1): one block of 256 threads per (plane,
// component), strided partial sums in a fixed order, then a tree (deterministic)
__global__ void __launch_bounds__(256)
ch_le_reduce(const ChParams P, const ChLeParams Q) {
  __shared__ double sm[256];
  const Layout & L = P.L;
  const int p = blockIdx.x, a = blockIdx.y;
  const double * scr = Q.scr + (i64) 8*p*L.s0;
  const int n = L.n[1]*L.n[2];
  double acc = 0.0;
  for (int i = threadIdx.x; i < n; i += 256) {
    const int jc = 1 + i/L.n[2], kc = 1 + i % L.n[2];
    const i64 col = (i64) jc*L.s1 + L.z0 + kc;
    acc += - scr[(2 + a)*L.s0 + col] + scr[(5 + a)*L.s0 + col];
  }
  sm[threadIdx.x] = acc;
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) {
    if ((int) threadIdx.x < w) sm[threadIdx.x] += sm[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x == 0) Q.fbar[3*p + a] = sm[0];
}

// pass B: grid (z tiles, n_y, nplanes), block (128); layers l = -1 .. 2
template <int EXACT, int ORDER>
__global__ void __launch_bounds__(128)
ch_le_pass_b(const ChParams P, const ChLeParams Q, const int do_force) {
  const Layout & L = P.L;
  const int p = blockIdx.z;
  const int c1 = blockIdx.y + 1;
  const int c2 = blockIdx.x*blockDim.x + threadIdx.x + 1;
  if (c2 > L.n[2]) return;
  const i64 col = (i64) c1*L.s1 + L.z0 + c2;
  const Symm S = {P.A, P.B, P.kappa};
  const double * scr = Q.scr + (i64) 8*p*L.s0;
  const double M = P.mobility;
  const i64 nsd = L.nsd;

  // offsets within a plane: [axis 1, 2][-2, -1, +1, +2]
  i64 d[3][4];
  {
    const int o[4] = {-2, -1, 1, 2};
    #pragma unroll
    for (int k = 0; k < 4; k++) {
      d[0][k] = 0;
      d[1][k] = (i64) ch_off(c1, L.n[1], P.wrap[1], o[k])*L.s1;
      d[2][k] = (i64) ch_off(c2, L.n[2], P.wrap[2], o[k]);
    }
  }

  #pragma unroll 1
  for (int l = -1; l <= 2; l++) {
    LeView V;
    le_view(P, Q, p, l, V);
    const int x = le_wrapx(P, Q.loc[p] + l);
    const i64 site = (i64) x*L.s0 + col;
    const double * phc = V.phi[2];
    const double ph0 = __ldg(phc + col);
    const double dq0 = __ldg(V.dsq[1] + col);

    if (do_force && (l == 0 || l == 1)) {
      // phi_force_flux_divergence, src/phi_force.c:431-476, with the corrected plane flux
      double pc[3], pn[3], f[3];
      double fe[3], fw[3];
      le_stress_col(S, phc, V.dsq[1], V.g[1], col, 0, pc);
      if (l == 0) {
	#pragma unroll
	for (int a = 0; a < 3; a++) fe[a] = scr[(2 + a)*L.s0 + col] + Q.ra*Q.fbar[3*p + a];
	le_stress_col(S, V.phi[1], V.dsq[0], V.g[0], col, 0, pn);
	#pragma unroll
	for (int a = 0; a < 3; a++) fw[a] = 0.5*(pn[a] + pc[a]);
      }
      else {
	le_stress_col(S, V.phi[3], V.dsq[2], V.g[2], col, 0, pn);
	#pragma unroll
	for (int a = 0; a < 3; a++) fe[a] = 0.5*(pn[a] + pc[a]);
	#pragma unroll
	for (int a = 0; a < 3; a++) fw[a] = scr[(5 + a)*L.s0 + col] - Q.ra*Q.fbar[3*p + a];
      }
      #pragma unroll
      for (int a = 0; a < 3; a++) f[a] = fe[a] - fw[a];
      #pragma unroll
      for (int b = 1; b < 3; b++) {
	double q0[3], qp[3], qm[3];
	le_stress_col(S, phc, V.dsq[1], V.g[1], col, b, q0);
	le_stress_col(S, phc, V.dsq[1], V.g[1], col + d[b][2], b, qp);
	le_stress_col(S, phc, V.dsq[1], V.g[1], col + d[b][1], b, qm);
	#pragma unroll
	for (int a = 0; a < 3; a++) {
	  f[a] = f[a] + 0.5*(qp[a] + q0[a]);
	  f[a] = f[a] - 0.5*(q0[a] + qm[a]);
	}
      }
      #pragma unroll
      for (int a = 0; a < 3; a++) P.force[a*nsd + site] = P.fconst[a] + (-f[a]);
    }

    if (ORDER > 0) {
      constexpr int O = (ORDER > 0) ? ORDER : 1;
      const double mu0 = ch_mu(S, ph0, dq0);
      double flo[3], fhi[3];
      // x faces
      {
	const double Mg = __dmul_rn(M, P.gmu[0]);
	const double phm = __ldg(V.phi[1] + col), php = __ldg(V.phi[3] + col);
	flo[0] = ch_face<O>(__ldg(V.ux[0] + col), __ldg(V.ux[1] + col), __ldg(V.phi[0] + col), phm, ph0, php,
			    ch_mu(S, phm, __ldg(V.dsq[0] + col)), mu0, M, Mg);
	fhi[0] = ch_face<O>(__ldg(V.ux[1] + col), __ldg(V.ux[2] + col), phm, ph0, php, __ldg(V.phi[4] + col),
			    mu0, ch_mu(S, php, __ldg(V.dsq[2] + col)), M, Mg);
	// phi_ch_le_fix_fluxes, src/phi_cahn_hilliard.c:663-715
	const int ny = L.n[1];
	if (l == 0) {
	  const int j1 = 1 + (c1 - Q.jdy_up - 2 + 2*ny) % ny, j2 = 1 + j1 % ny;
	  const double * fwp = scr + L.s0 + L.z0 + c2;
	  const double bw = __dadd_rn(__dmul_rn(fwp[(i64) j1*L.s1], Q.fr_up), __dmul_rn(fwp[(i64) j2*L.s1], 1.0 - Q.fr_up));
	  fhi[0] = __dmul_rn(0.5, __dadd_rn(scr[col], bw));
	}
	if (l == 1) {
	  const int j1 = 1 + (c1 - Q.jdy_dn - 2 + 2*ny) % ny, j2 = 1 + j1 % ny;
	  const double * fep = scr + L.z0 + c2;
	  const double be = __dadd_rn(__dmul_rn(fep[(i64) j1*L.s1], Q.fr_dn), __dmul_rn(fep[(i64) j2*L.s1], 1.0 - Q.fr_dn));
	  flo[0] = __dmul_rn(0.5, __dadd_rn(scr[L.s0 + col], be));
	}
      }
      // y and z faces: within the plane
      const double * dsc = V.dsq[1];
      #pragma unroll
      for (int a = 1; a < 3; a++) {
	const double * ua = P.u + a*nsd + (i64) x*L.s0;
	const double u0 = __ldg(ua + col);
	const double um = __ldg(ua + col + d[a][1]), up = __ldg(ua + col + d[a][2]);
	const double phm = __ldg(phc + col + d[a][1]), php = __ldg(phc + col + d[a][2]);
	const double phm2 = (O >= 3) ? __ldg(phc + col + d[a][0]) : 0.0;
	const double php2 = (O >= 3) ? __ldg(phc + col + d[a][3]) : 0.0;
	const double Mg = __dmul_rn(M, P.gmu[a]);
	flo[a] = ch_face<O>(um, u0, phm2, phm, ph0, php, ch_mu(S, phm, __ldg(dsc + col + d[a][1])), mu0, M, Mg);
	fhi[a] = ch_face<O>(u0, up, phm, ph0, php, php2, mu0, ch_mu(S, php, __ldg(dsc + col + d[a][2])), M, Mg);
      }
      double dphi = __dadd_rn(fhi[0], -flo[0]);
      dphi = __dadd_rn(dphi, fhi[1]);
      dphi = __dadd_rn(dphi, -flo[1]);
      if (!P.skipz) {
	dphi = __dadd_rn(dphi, fhi[2]);
	dphi = __dadd_rn(dphi, -flo[2]);
      }
      P.phi_out[site] = __dadd_rn(ph0, -dphi);
    }
  }
}

}   // namespace

#ifndef B200LB_KERNEL_SOURCE_ONLY      // tools/ch_emulate.cpp steps through the kernel source on the host

template <int METHOD, int ORDER>
static cudaError_t launch(const ChParams & P, cudaStream_t st) {
  const Layout & L = P.L;
  dim3 block(CH_TZ, CH_TY, 1);
  if (P.nx < 1) return cudaErrorInvalidValue;
  dim3 grid((L.n[2] + CH_TZ - 1)/CH_TZ, (L.n[1] + CH_TY - 1)/CH_TY, (L.n[0] + P.nx - 1)/P.nx);
  ch_kernel<CH_EXACT, METHOD, ORDER><<<grid, block, 0, st>>>(P);
  return cudaGetLastError();
}

template <int METHOD>
static cudaError_t launch_order(const ChParams & P, int order, cudaStream_t st) {
  switch (order) {
  case 0: return (METHOD == 0) ? cudaSuccess : launch<METHOD, 0>(P, st);
  case 1: return launch<METHOD, 1>(P, st);
  case 2: return launch<METHOD, 2>(P, st);
  case 3: return launch<METHOD, 3>(P, st);
  case 4: return launch<METHOD, 4>(P, st);
  }
  return cudaErrorInvalidValue;
}

#if CH_EXACT
#define CH_LAUNCH ch_launch_exact
#else
#define CH_LAUNCH ch_launch_fast
#endif

#if CH_EXACT
#define CH_LE_LAUNCH ch_le_launch_exact
#else
#define CH_LE_LAUNCH ch_le_launch_fast
#endif

template <int ORDER>
static cudaError_t le_launch(const ChParams & P, const ChLeParams & Q, bool do_force, cudaStream_t st) {
  const Layout & L = P.L;
  dim3 block(128, 1, 1);
  dim3 grid((L.n[2] + 127)/128, L.n[1], Q.nplanes);
  ch_le_pass_a<CH_EXACT, ORDER><<<grid, block, 0, st>>>(P, Q, do_force ? 1 : 0);
  if (do_force) ch_le_reduce<<<dim3(Q.nplanes, 3, 1), 256, 0, st>>>(P, Q);
  ch_le_pass_b<CH_EXACT, ORDER><<<grid, block, 0, st>>>(P, Q, do_force ? 1 : 0);
  return cudaGetLastError();
}

cudaError_t CH_LE_LAUNCH(const ChParams & P, const ChLeParams & Q, bool do_force, bool do_update, cudaStream_t st) {
  if (Q.nplanes < 1 || Q.nplanes > CH_LE_MAXPLANE) return cudaErrorInvalidValue;
  if (do_force && P.method == CH_FORCE_NONE) do_force = false;
  if (do_force && P.accumulate) return cudaErrorInvalidValue;     // pass B writes fconst + f
  switch (do_update ? P.order : 0) {
  case 0: return do_force ? le_launch<0>(P, Q, true, st) : cudaSuccess;
  case 1: return le_launch<1>(P, Q, do_force, st);
  case 2: return le_launch<2>(P, Q, do_force, st);
  case 3: return le_launch<3>(P, Q, do_force, st);
  case 4: return le_launch<4>(P, Q, do_force, st);
  }
  return cudaErrorInvalidValue;
}

cudaError_t CH_LAUNCH(const ChParams & P, bool do_force, bool do_update, cudaStream_t st) {
  const int order = do_update ? P.order : 0;
  int method = do_force ? P.method : 0;
  if (method == CH_FORCE_NONE) method = 0;
  if (do_update && (order < 1 || order > 4)) return cudaErrorInvalidValue;
  switch (method) {
  case 0: return launch_order<0>(P, order, st);
  case CH_FORCE_STRESS_DIVERGENCE: return launch_order<CH_FORCE_STRESS_DIVERGENCE>(P, order, st);
  case CH_FORCE_PHI_GRADMU: return launch_order<CH_FORCE_PHI_GRADMU>(P, order, st);
  }
  return cudaErrorInvalidValue;
}

#endif

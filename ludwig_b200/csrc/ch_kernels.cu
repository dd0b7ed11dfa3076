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
      // phi -= (+ fe - fw + fy - fy[index - ys] + wz*fz - wz*fz[index - 1]), wz = 1
      double dphi = __dadd_rn(fhi[0], -flo[0]);
      dphi = __dadd_rn(dphi, fhi[1]);
      dphi = __dadd_rn(dphi, -flo[1]);
      dphi = __dadd_rn(dphi, fhi[2]);
      dphi = __dadd_rn(dphi, -flo[2]);
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

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
// A face flux is evaluated by both sites that share the face from the same operands in
// the same order, so the two values are identical and phi is conserved exactly as in
// the reference (which also stores fw and fe separately).
//
// Compiled with -fmad=false: every operation is rounded on its own and the expressions
// keep the reference's order, so force and phi agree with it bit for bit.

#include "ch_kernels.cuh"

#define CH_TZ 64
#define CH_TY 4
#define CH_NX 8      // default x-planes per block (ChParams.nx): neighbours along x are re-read through L1

struct Symm { double a, b, kappa; };

// fe_symm_mu, src/symmetric.c:307-320
__device__ __forceinline__ double ch_mu(const Symm & s, double phi, double delsq) {
  return s.a*phi + s.b*phi*phi*phi - s.kappa*delsq;
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

// One face of advection_x (src/advection.c). The flux is stored at the site whose
// order parameter is phi0; the face lies between it and the site of phi1; phim and phi2
// are the sites beyond (behind phi0, beyond phi1). west: the face is the lower x face
// (upwind of it is phi1's side when u > 0), otherwise an upper face (when u < 0).
template <int ORDER>
__device__ __forceinline__ double ch_adv(bool west, double u0, double u1, double phim,
					 double phi0, double phi1, double phi2) {
  if (ORDER == 1) {
    const double u = 0.5*(u0 + u1);
    double ph = phi0;
    if (west ? (u > 0.0) : (u < 0.0)) ph = phi1;
    return u*ph;
  }
  if (ORDER == 2) {
    if (west) return 0.5*(u0 + u1)*0.5*(phi1 + phi0);
    return 0.5*(u0 + u1)*0.5*(phi0 + phi1);
  }
  {
    const double a1 = -0.213933, a2 = 0.927865, a3 = 0.286067;
    const double u = 0.5*(u0 + u1);
    double fd1 = phim, fd2 = phi0, fd3 = phi1;
    if (west ? (u > 0.0) : (u < 0.0)) { fd1 = phi2; fd2 = phi1; fd3 = phi0; }
    return u*(a1*fd1 + a2*fd2 + a3*fd3);
  }
}

// offset (in sites) from halo-inclusive coordinate c to c + o along an axis of interior
// extent n, wrapping where the axis is periodic within the rank
__device__ __forceinline__ int ch_off(int c, int n, int wrap, int o) {
  int t = c + o;
  if (wrap) { if (t < 1) t += n; else if (t > n) t -= n; }
  return t - c;
}

// METHOD: 0 = no force pass, CH_FORCE_STRESS_DIVERGENCE, CH_FORCE_PHI_GRADMU
// ORDER:  0 = no update pass, 1..3 = advection scheme of the Cahn-Hilliard update
template <int METHOD, int ORDER>
__global__ void __launch_bounds__(CH_TZ*CH_TY)
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
      double g0[3], g1[3], f[3];
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
      // the six face fluxes of this site: advection_x, then phi_ch_flux_mu1_kernel and
      // phi_ch_flux_mu_ext_kernel accumulated on top, then phi_ch_ufs_kernel
      const double M = P.mobility;
      const double mu0 = ch_mu(S, ph0, dq0);
      double u0[3], um[3], up[3], phm2[3], php2[3];
      #pragma unroll
      for (int a = 0; a < 3; a++) {
	u0[a] = __ldg(u + a*nsd + site);
	um[a] = __ldg(u + a*nsd + site + d[a][1]);
	up[a] = __ldg(u + a*nsd + site + d[a][2]);
	phm2[a] = (ORDER == 3) ? __ldg(phi + site + d[a][0]) : 0.0;
	php2[a] = (ORDER == 3) ? __ldg(phi + site + d[a][3]) : 0.0;
      }

      double fw = ch_adv<ORDER>(true, u0[0], um[0], php[0], ph0, phm[0], phm2[0]);
      fw -= M*(mu0 - ch_mu(S, phm[0], dqm[0]));
      fw -= M*P.gmu[0];
      double fe = ch_adv<ORDER>(false, u0[0], up[0], phm[0], ph0, php[0], php2[0]);
      fe -= M*(ch_mu(S, php[0], dqp[0]) - mu0);
      fe -= M*P.gmu[0];

      // upper faces in y and z belong to this site, the lower ones to the site below:
      // the same expressions with that site in the role of "this"
      double fhi[3], flo[3];
      #pragma unroll
      for (int a = 1; a < 3; a++) {
	double fl = ch_adv<ORDER>(false, u0[a], up[a], phm[a], ph0, php[a], php2[a]);
	fl -= M*(ch_mu(S, php[a], dqp[a]) - mu0);
	fl -= M*P.gmu[a];
	fhi[a] = fl;
	fl = ch_adv<ORDER>(false, um[a], u0[a], phm2[a], phm[a], ph0, php[a]);
	fl -= M*(mu0 - ch_mu(S, phm[a], dqm[a]));
	fl -= M*P.gmu[a];
	flo[a] = fl;
      }

      const double wz = 1.0;
      double ph = ph0;
      ph -= (+ fe - fw + fhi[1] - flo[1] + wz*fhi[2] - wz*flo[2]);
      P.phi_out[site] = ph;
    }
  }
}

#ifndef B200LB_KERNEL_SOURCE_ONLY      // tools/ch_emulate.cpp steps through the kernel source on the host

template <int METHOD, int ORDER>
static cudaError_t launch(const ChParams & P, cudaStream_t st) {
  const Layout & L = P.L;
  dim3 block(CH_TZ, CH_TY, 1);
  if (P.nx < 1) return cudaErrorInvalidValue;
  dim3 grid((L.n[2] + CH_TZ - 1)/CH_TZ, (L.n[1] + CH_TY - 1)/CH_TY, (L.n[0] + P.nx - 1)/P.nx);
  ch_kernel<METHOD, ORDER><<<grid, block, 0, st>>>(P);
  return cudaGetLastError();
}

template <int METHOD>
static cudaError_t launch_order(const ChParams & P, int order, cudaStream_t st) {
  switch (order) {
  case 0: return (METHOD == 0) ? cudaSuccess : launch<METHOD, 0>(P, st);
  case 1: return launch<METHOD, 1>(P, st);
  case 2: return launch<METHOD, 2>(P, st);
  case 3: return launch<METHOD, 3>(P, st);
  }
  return cudaErrorInvalidValue;
}

cudaError_t ch_launch(const ChParams & P, bool do_force, bool do_update, cudaStream_t st) {
  const int order = do_update ? P.order : 0;
  int method = do_force ? P.method : 0;
  if (method == CH_FORCE_NONE) method = 0;
  if (do_update && (order < 1 || order > 3)) return cudaErrorInvalidValue;
  switch (method) {
  case 0: return launch_order<0>(P, order, st);
  case CH_FORCE_STRESS_DIVERGENCE: return launch_order<CH_FORCE_STRESS_DIVERGENCE>(P, order, st);
  case CH_FORCE_PHI_GRADMU: return launch_order<CH_FORCE_PHI_GRADMU>(P, order, st);
  }
  return cudaErrorInvalidValue;
}

#endif

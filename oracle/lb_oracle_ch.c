/*
 * lb_oracle_ch.c -- TEST INFRASTRUCTURE ONLY. Not part of the product.
 *
 * CPU restatement of the finite-difference order-parameter route of Ludwig
 * (zazu29/ludwig v0.20.1) that "free_energy symmetric" selects: the order
 * parameter phi is a field (not a second distribution), the fluid feels it as
 * a body force, phi itself follows Cahn-Hilliard with the fluid velocity of the
 * previous collision (src/ludwig.c:558-575, 727-768):
 *
 *   field_halo(phi); field_grad_compute(phi_grad)
 *   phi_force_calculation      src/phi_force.c:74
 *     pth_stress_compute       src/phi_force_stress.c:171 (fe_symm_str_v src/symmetric.c:371-420)
 *     pth_force_fluid_kernel_v src/phi_force_colloid.c:323-480
 *     or phi_grad_mu_fluid + phi_grad_mu_external  src/phi_grad_mu.c:294-380, 528-560
 *   phi_cahn_hilliard          src/phi_cahn_hilliard.c:206
 *     hydro_u_halo             src/hydro.c
 *     advection_x              src/advection.c:433 (orders 1, 2, 3: :542, :790, :977)
 *     phi_ch_flux_mu1          src/phi_cahn_hilliard.c:349 (fe_symm_mu src/symmetric.c:307)
 *     phi_ch_flux_mu_ext       src/phi_cahn_hilliard.c:1404
 *     phi_ch_ufs_kernel        src/phi_cahn_hilliard.c:1025
 *
 * Parity status: PINNED (tests/test_oracle_vs_ref.py::test_cahn_hilliard_* against the
 * unmodified reference built in place with nhalo = 2, tests/golden/ref_ch.npz, and the
 * reference's regression logs serial-spin-fd1, serial-symm-dr1, serial-symm-pat,
 * serial-muex-st1).
 *
 * Geometry. The reference runs this route with a halo two sites wide: the stress is
 * evaluated one site into the halo (so the gradient is too), the third-order advection
 * reads phi two sites away. Here every array keeps the one-site halo of lb_oracle.h and the
 * box is periodic in all three directions on one rank: a halo site is the periodic
 * image of an interior site, so the same arithmetic on the same operands (gradient and
 * stress of the image = gradient and stress of the site) gives bit-identical values,
 * and a read two sites out wraps around. All arithmetic keeps the reference's order of
 * operations.
 */

#include <stdlib.h>
#include <string.h>

#include "lb_oracle.h"

static orc_ch_t * ch_attached_ = NULL;

int orc_ch_attach(orc_ch_t * ch) { ch_attached_ = ch; return 0; }
orc_ch_t * orc_ch_attached(void) { return ch_attached_; }

/* 0-based halo-inclusive coordinate (0 .. n+1 stored) of a coordinate up to two sites
 * outside the interior, periodic image where it is not stored */
static int wrapc(int c, int n) {
  if (c < 0) return c + n;
  if (c > n + 1) return c - n;
  return c;
}

static long site(const orc_box_t * box, int ic, int jc, int kc) {
  ic = wrapc(ic, box->nlocal[0]);
  jc = wrapc(jc, box->nlocal[1]);
  kc = wrapc(kc, box->nlocal[2]);
  return ((long) ic*box->nall[1] + jc)*box->nall[2] + kc;
}

/* fe_symm_mu, src/symmetric.c:307-320 */
static double mu_symm(const orc_param_t * prm, double phi, double delsq) {
  return prm->a*phi + prm->b*phi*phi*phi - prm->kappa*delsq;
}

/* pth_stress_compute: the chemical stress at every site of the box, halo included
 * (the reference's region is nextra = 1). phi, grad, delsq need their halos.
 * pth[index*9 + ia*3 + ib]; fe_symm_str_v src/symmetric.c:371-420. */
int orc_ch_stress(const orc_box_t * box, const orc_param_t * prm,
		  const double * phi, const double * grad, const double * delsq,
		  double * pth) {
  const double a = prm->a, b = prm->b, kappa = prm->kappa;
  long index;

  #pragma omp parallel for
  for (index = 0; index < box->nsite; index++) {
    const double * g = grad + 3*index;
    const double ph = phi[index];
    const double d2 = delsq[index];
    int ia, ib;
    double p0 = 0.5*a*ph*ph + 0.75*b*ph*ph*ph*ph - kappa*ph*d2
      - 0.5*kappa*(g[0]*g[0] + g[1]*g[1] + g[2]*g[2]);
    for (ia = 0; ia < 3; ia++) {
      for (ib = 0; ib < 3; ib++) {
	const double d = (ia == ib);
	pth[9*index + 3*ia + ib] = p0*d + kappa*g[ia]*g[ib];
      }
    }
  }
  return 0;
}

/* pth_force_fluid_kernel_v, src/phi_force_colloid.c:323-480: F_a = - d_b P_ab with the
 * stress interpolated to the faces; added to the force field of interior sites. */
int orc_ch_force_divergence(const orc_box_t * box, const double * pth, double * force) {
  int ic;
  const long xs = (long) box->nall[1]*box->nall[2], ys = box->nall[2];

  #pragma omp parallel for
  for (ic = 1; ic <= box->nlocal[0]; ic++) {
    int jc, kc, ia;
    for (jc = 1; jc <= box->nlocal[1]; jc++) {
      for (kc = 1; kc <= box->nlocal[2]; kc++) {
	const long index = ((long) ic*box->nall[1] + jc)*box->nall[2] + kc;
	const double * p0 = pth + 9*index;
	double f[3];
	const double * p1;
	p1 = pth + 9*(index + xs);
	for (ia = 0; ia < 3; ia++) f[ia] = -0.5*(p1[3*ia + 0] + p0[3*ia + 0]);
	p1 = pth + 9*(index - xs);
	for (ia = 0; ia < 3; ia++) f[ia] += 0.5*(p1[3*ia + 0] + p0[3*ia + 0]);
	p1 = pth + 9*(index + ys);
	for (ia = 0; ia < 3; ia++) f[ia] -= 0.5*(p1[3*ia + 1] + p0[3*ia + 1]);
	p1 = pth + 9*(index - ys);
	for (ia = 0; ia < 3; ia++) f[ia] += 0.5*(p1[3*ia + 1] + p0[3*ia + 1]);
	p1 = pth + 9*(index + 1);
	for (ia = 0; ia < 3; ia++) f[ia] -= 0.5*(p1[3*ia + 2] + p0[3*ia + 2]);
	p1 = pth + 9*(index - 1);
	for (ia = 0; ia < 3; ia++) f[ia] += 0.5*(p1[3*ia + 2] + p0[3*ia + 2]);
	for (ia = 0; ia < 3; ia++) force[3*index + ia] += f[ia];
      }
    }
  }
  return 0;
}

/* phi_grad_mu_fluid_kernel + phi_grad_mu_external_kernel, src/phi_grad_mu.c:294-380,
 * 528-560: F_a = - phi d_a mu (central difference), then F_a -= phi grad_mu^ext_a. */
int orc_ch_force_gradmu(const orc_box_t * box, const orc_param_t * prm,
			const double grad_mu[3], const double * phi,
			const double * delsq, double * force) {
  int ic;
  const long xs = (long) box->nall[1]*box->nall[2], ys = box->nall[2];
  const long str[3] = {xs, ys, 1};

  #pragma omp parallel for
  for (ic = 1; ic <= box->nlocal[0]; ic++) {
    int jc, kc, ia;
    for (jc = 1; jc <= box->nlocal[1]; jc++) {
      for (kc = 1; kc <= box->nlocal[2]; kc++) {
	const long index = ((long) ic*box->nall[1] + jc)*box->nall[2] + kc;
	const double phi0 = phi[index];
	for (ia = 0; ia < 3; ia++) {
	  const double mum1 = mu_symm(prm, phi[index - str[ia]], delsq[index - str[ia]]);
	  const double mup1 = mu_symm(prm, phi[index + str[ia]], delsq[index + str[ia]]);
	  double f = 0.0;
	  f += -phi0*0.5*(mup1 - mum1);
	  force[3*index + ia] += f;
	}
	for (ia = 0; ia < 3; ia++) {
	  const double f = -phi0*grad_mu[ia];
	  force[3*index + ia] += f;
	}
      }
    }
  }
  return 0;
}

/* one face of advection_x: the velocity at the face is the mean of the two sites, the
 * order parameter is taken upwind (order 1), centred (2) or from the three-point
 * upwind-biased stencil of src/advection.c:977-1180 (3).
 *   up = +1: the second site of the face is the one at +1 (east, y, z faces:
 *            upwind of it when u < 0); up = -1: west face (upwind when u > 0).
 * phim, phi0, phi1, phi2: the order parameter at -1, 0, +1, +2 along the face normal
 * relative to the site the flux is stored at (for the west face mirrored: +1, 0, -1, -2). */
static double adv_face(int order, int west, double u0, double u1,
		       double phim, double phi0, double phi1, double phi2) {
  const double a1 = -0.213933, a2 = 0.927865, a3 = 0.286067;
  if (order == 1) {
    const double u = 0.5*(u0 + u1);
    double ph = phi0;
    if (west) { if (u > 0.0) ph = phi1; }
    else      { if (u < 0.0) ph = phi1; }
    return u*ph;
  }
  if (order == 2) {
    if (west) return 0.5*(u0 + u1)*1*0.5*(phi1 + phi0);
    return 0.5*(u0 + u1)*1*0.5*(phi0 + phi1);
  }
  {
    const double u = 0.5*1*(u0 + u1);
    double fd1, fd2, fd3;
    if (west ? (u > 0.0) : (u < 0.0)) {
      /* the far site is upwind: stencil phi2, phi1, phi0 */
      fd1 = phi2; fd2 = phi1; fd3 = phi0;
    }
    else {
      fd1 = phim; fd2 = phi0; fd3 = phi1;
    }
    return u*(a1*fd1 + a2*fd2 + a3*fd3);
  }
}

/* advection_x (src/advection.c:433) for orders 1-3 and no Lees-Edwards planes:
 * flux[0] = fw, flux[1] = fe, flux[2] = fy, flux[3] = fz, each [nsite], on the region
 * ic = 1..n, jc = 0..n, kc = 0..n as in the reference. u needs its halo. */
int orc_ch_advection(const orc_box_t * box, int order, const double * u,
		     const double * phi, double * flux) {
  int ic;
  double * fw = flux, * fe = flux + box->nsite, * fy = flux + 2*box->nsite,
    * fz = flux + 3*box->nsite;

  if (order < 1 || order > 3) return -1;

  #pragma omp parallel for
  for (ic = 1; ic <= box->nlocal[0]; ic++) {
    int jc, kc;
    for (jc = 0; jc <= box->nlocal[1]; jc++) {
      for (kc = 0; kc <= box->nlocal[2]; kc++) {
	const long i0 = site(box, ic, jc, kc);
	const double * u0 = u + 3*i0;
	const double p0 = phi[i0];
	fw[i0] = adv_face(order, 1, u0[0], u[3*site(box, ic - 1, jc, kc) + 0],
			  phi[site(box, ic + 1, jc, kc)], p0,
			  phi[site(box, ic - 1, jc, kc)], phi[site(box, ic - 2, jc, kc)]);
	fe[i0] = adv_face(order, 0, u0[0], u[3*site(box, ic + 1, jc, kc) + 0],
			  phi[site(box, ic - 1, jc, kc)], p0,
			  phi[site(box, ic + 1, jc, kc)], phi[site(box, ic + 2, jc, kc)]);
	fy[i0] = adv_face(order, 0, u0[1], u[3*site(box, ic, jc + 1, kc) + 1],
			  phi[site(box, ic, jc - 1, kc)], p0,
			  phi[site(box, ic, jc + 1, kc)], phi[site(box, ic, jc + 2, kc)]);
	fz[i0] = adv_face(order, 0, u0[2], u[3*site(box, ic, jc, kc + 1) + 2],
			  phi[site(box, ic, jc, kc - 1)], p0,
			  phi[site(box, ic, jc, kc + 1)], phi[site(box, ic, jc, kc + 2)]);
      }
    }
  }
  return 0;
}

/* phi_ch_flux_mu1_kernel + phi_ch_flux_mu_ext_kernel (src/phi_cahn_hilliard.c:349-418,
 * 1404-1436): the diffusive fluxes - M (mu(+) - mu(-)) and - M grad_mu^ext are
 * accumulated onto the advective ones, same region. phi and delsq need their halos. */
int orc_ch_flux_mu(const orc_box_t * box, const orc_param_t * prm,
		   const double grad_mu[3], const double * phi, const double * delsq,
		   double * flux) {
  int ic;
  const double mobility = prm->mobility;
  double * fw = flux, * fe = flux + box->nsite, * fy = flux + 2*box->nsite,
    * fz = flux + 3*box->nsite;

  #pragma omp parallel for
  for (ic = 1; ic <= box->nlocal[0]; ic++) {
    int jc, kc;
    for (jc = 0; jc <= box->nlocal[1]; jc++) {
      for (kc = 0; kc <= box->nlocal[2]; kc++) {
	const long i0 = site(box, ic, jc, kc);
	long i1;
	double mu0, mu1;
	mu0 = mu_symm(prm, phi[i0], delsq[i0]);
	i1 = site(box, ic - 1, jc, kc);
	mu1 = mu_symm(prm, phi[i1], delsq[i1]);
	fw[i0] -= mobility*(mu0 - mu1);
	i1 = site(box, ic + 1, jc, kc);
	mu1 = mu_symm(prm, phi[i1], delsq[i1]);
	fe[i0] -= mobility*(mu1 - mu0);
	i1 = site(box, ic, jc + 1, kc);
	mu1 = mu_symm(prm, phi[i1], delsq[i1]);
	fy[i0] -= mobility*(mu1 - mu0);
	i1 = site(box, ic, jc, kc + 1);
	mu1 = mu_symm(prm, phi[i1], delsq[i1]);
	fz[i0] -= mobility*(mu1 - mu0);
      }
    }
  }
  #pragma omp parallel for
  for (ic = 1; ic <= box->nlocal[0]; ic++) {
    int jc, kc;
    for (jc = 0; jc <= box->nlocal[1]; jc++) {
      for (kc = 0; kc <= box->nlocal[2]; kc++) {
	const long i0 = site(box, ic, jc, kc);
	fw[i0] -= mobility*grad_mu[0];
	fe[i0] -= mobility*grad_mu[0];
	fy[i0] -= mobility*grad_mu[1];
	fz[i0] -= mobility*grad_mu[2];
      }
    }
  }
  return 0;
}

/* phi_ch_ufs_kernel, src/phi_cahn_hilliard.c:1025-1062 (three-dimensional: wz = 1) */
int orc_ch_update(const orc_box_t * box, const double * flux, double * phi) {
  int ic;
  const double wz = 1.0;
  const long ys = box->nall[2];
  const double * fw = flux, * fe = flux + box->nsite, * fy = flux + 2*box->nsite,
    * fz = flux + 3*box->nsite;

  #pragma omp parallel for
  for (ic = 1; ic <= box->nlocal[0]; ic++) {
    int jc, kc;
    for (jc = 1; jc <= box->nlocal[1]; jc++) {
      for (kc = 1; kc <= box->nlocal[2]; kc++) {
	const long index = ((long) ic*box->nall[1] + jc)*box->nall[2] + kc;
	double ph = phi[index];
	ph -= (+ fe[index] - fw[index] + fy[index] - fy[index - ys]
	       + wz*fz[index] - wz*fz[index - 1]);
	phi[index] = ph;
      }
    }
  }
  return 0;
}

/* phi_force_calculation (src/phi_force.c:74-139) for a fluid-only periodic box.
 * Expects phi with its halo; fills the halos of grad and delsq (see "Geometry"). */
int orc_phi_force(const orc_box_t * box, const orc_param_t * prm, const orc_ch_t * ch,
		  const double * phi, double * grad, double * delsq, double * force) {
  if (ch->method == ORC_FORCE_NONE) return 0;
  orc_halo_periodic(box, 3, grad);
  orc_halo_periodic(box, 1, delsq);
  if (ch->method == ORC_FORCE_STRESS_DIVERGENCE) {
    double * pth = (double *) malloc(sizeof(double)*9*box->nsite);
    if (pth == NULL) return -1;
    orc_ch_stress(box, prm, phi, grad, delsq, pth);
    orc_ch_force_divergence(box, pth, force);
    free(pth);
    return 0;
  }
  if (ch->method == ORC_FORCE_PHI_GRADMU) {
    return orc_ch_force_gradmu(box, prm, ch->grad_mu, phi, delsq, force);
  }
  return -1;
}

/* phi_cahn_hilliard (src/phi_cahn_hilliard.c:206-283): hydro_u_halo, advective then
 * diffusive fluxes, forward step. phi and delsq with halos on entry; the halo of phi is
 * stale on return (the next field_halo refreshes it), as in the reference. */
int orc_cahn_hilliard(const orc_box_t * box, const orc_param_t * prm, const orc_ch_t * ch,
		      double * phi, double * delsq, double * u) {
  int rc;
  double * flux = (double *) calloc(4*box->nsite, sizeof(double));
  if (flux == NULL) return -1;
  orc_halo_periodic(box, 3, u);
  orc_halo_periodic(box, 1, delsq);
  rc = orc_ch_advection(box, ch->order, u, phi, flux);
  if (rc == 0) rc = orc_ch_flux_mu(box, prm, ch->grad_mu, phi, delsq, flux);
  if (rc == 0) rc = orc_ch_update(box, flux, phi);
  free(flux);
  return rc;
}

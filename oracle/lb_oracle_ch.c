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

/* a site whose coordinate along a wall direction lies outside 1 .. n is solid */
static int ch_solid(const orc_box_t * box, const orc_wall_t * wall, int ic, int jc, int kc) {
  const int c[3] = {ic, jc, kc};
  if (wall == NULL) return 0;
  for (int d = 0; d < 3; d++) {
    if (wall->isboundary[d] && (c[d] < 1 || c[d] > box->nlocal[d])) return 1;
  }
  return 0;
}

/* pth_force_fluid_wall_driver (src/phi_force_colloid.c:227-268): pth_force_map_kernel
 * (:490-676) takes the stress of the site itself at a face shared with solid, and
 * pth_force_wall_kernel (:677-790) books what crosses those faces to the wall
 * (wall_momentum_add). Same face order as without walls. */
static int ch_force_divergence_walls(const orc_box_t * box, orc_wall_t * wall,
				     const double * pth, double * force) {
  const long xs = (long) box->nall[1]*box->nall[2], ys = box->nall[2];
  const long str[3] = {xs, ys, 1};
  double fwall[3] = {0.0, 0.0, 0.0};
  for (int ic = 1; ic <= box->nlocal[0]; ic++) {
    for (int jc = 1; jc <= box->nlocal[1]; jc++) {
      for (int kc = 1; kc <= box->nlocal[2]; kc++) {
	const long index = ((long) ic*box->nall[1] + jc)*box->nall[2] + kc;
	const double * p0 = pth + 9*index;
	double f[3] = {0.0, 0.0, 0.0};
	for (int b = 0; b < 3; b++) {
	  for (int side = +1; side >= -1; side -= 2) {
	    const int n[3] = {ic + side*(b == 0), jc + side*(b == 1), kc + side*(b == 2)};
	    const int solid = ch_solid(box, wall, n[0], n[1], n[2]);
	    const double * p1 = pth + 9*(index + side*str[b]);
	    for (int ia = 0; ia < 3; ia++) {
	      const double t = solid ? p0[3*ia + b] : 0.5*(p1[3*ia + b] + p0[3*ia + b]);
	      if (b == 0 && side == +1) f[ia] = -t;
	      else if (side == +1) f[ia] -= t;
	      else f[ia] += t;
	      if (solid) fwall[ia] += (side == +1) ? -p0[3*ia + b] : p0[3*ia + b];
	    }
	  }
	}
	for (int ia = 0; ia < 3; ia++) force[3*index + ia] += f[ia];
      }
    }
  }
  for (int ia = 0; ia < 3; ia++) wall->fnet[ia] += -fwall[ia];
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
  if (order == 4) {
    /* advection_le_4th, src/advection.c:1188-1290: four-point centred interpolation */
    const double b1 = (1.0/16.0), b2 = (9.0/16.0);
    const double u = 0.5*(u0 + u1);
    if (west) return u*(- b1*phi2 + b2*phi1 + b2*phi0 - b1*phim);
    return u*(- b1*phim + b2*phi0 + b2*phi1 - b1*phi2);
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

  if (order < 1 || order > 4) return -1;

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

/* phi_ch_ufs_kernel, src/phi_cahn_hilliard.c:1025-1062; wz = 0 when the lattice is one
 * site thick (phi_ch_update_forward_step :999) */
int orc_ch_update(const orc_box_t * box, const double * flux, double * phi) {
  int ic;
  const double wz = (box->nlocal[2] == 1) ? 0.0 : 1.0;
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
    if (orc_wall_attached()) ch_force_divergence_walls(box, orc_wall_attached(), pth, force);
    else orc_ch_force_divergence(box, pth, force);
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
  if (rc == 0 && orc_wall_attached()) {
    /* advection_bcs_no_normal_flux (src/advection_bcs.c:38-140): no flux through a face
     * shared with solid */
    const orc_wall_t * wall = orc_wall_attached();
    double * fw = flux, * fe = flux + box->nsite, * fy = flux + 2*box->nsite, * fz = flux + 3*box->nsite;
    for (int ic = 1; ic <= box->nlocal[0]; ic++) {
      for (int jc = 0; jc <= box->nlocal[1]; jc++) {
	for (int kc = 0; kc <= box->nlocal[2]; kc++) {
	  const long i0 = site(box, ic, jc, kc);
	  const double mask = !ch_solid(box, wall, ic, jc, kc);
	  fw[i0] *= mask*(!ch_solid(box, wall, ic - 1, jc, kc));
	  fe[i0] *= mask*(!ch_solid(box, wall, ic + 1, jc, kc));
	  fy[i0] *= mask*(!ch_solid(box, wall, ic, jc + 1, kc));
	  fz[i0] *= mask*(!ch_solid(box, wall, ic, jc, kc + 1));
	}
      }
    }
  }
  if (rc == 0) rc = orc_ch_update(box, flux, phi);
  free(flux);
  return rc;
}

/* ===========================================================================
 * The finite-difference order parameter across Lees-Edwards planes
 * ("free_energy symmetric" with N_LE_plane > 0; regression logs serial-le3d-st*).
 *
 * With planes the reference cannot use periodic images: a site next to a plane sees,
 * on the far side, values displaced along y by what the plane has slid. It keeps them
 * in buffer x-planes behind the real lattice (halo two sites wide, 2 x 2 buffer planes
 * per plane) and reaches them through lees_edw_ic_to_buff. This part of the oracle
 * therefore works on arrays laid out like the reference's (x-planes -1 .. n_x + 2 then
 * the buffers; y, z from -1 to n + 2) and restates, in the reference's order:
 *
 *   field_halo (periodic, width 2) ............. le_fill
 *   field_leesedwards   src/field.c:500-600 .... le_field_buffers
 *   grad_3d_{7pt,27pt}_fluid_operator + _le
 *        src/gradient_3d_7pt_fluid.c:185-460, _27pt_fluid.c:216-560 ... le_gradient
 *   phi_force_flux      src/phi_force.c:289-340 (compute_fluxes :350, fix_local :595,
 *                       divergence :431): with planes the force is ALWAYS the
 *                       divergence of the face-interpolated stress, whatever the method
 *   hydro_u_halo, hydro_lees_edwards  src/hydro.c:190, 412-500 (time t + 1, linear
 *                       interpolation, velocity jump added) ........ le_u_buffers
 *   advection_le_{1st,2nd,3rd,4th}  src/advection.c:542-1290 ....... le_advection
 *   phi_ch_flux_mu1, phi_ch_flux_mu_ext  src/phi_cahn_hilliard.c:349, 1404
 *   phi_ch_le_fix_fluxes  src/phi_cahn_hilliard.c:618-735: the east flux of the sites
 *                       below a plane and the west flux of those above it are
 *                       averaged with their (interpolated) counterparts
 *   phi_ch_ufs_kernel   src/phi_cahn_hilliard.c:1025 (wz = 0 when n_z = 1)
 *
 * One rank; steady or oscillatory planes; advection orders 1-4.
 * Parity status: PINNED, tests/test_oracle_vs_ref.py::test_cahn_hilliard_lees_edwards.
 * =========================================================================== */

#include <math.h>

typedef struct {
  int n[3];            /* nlocal */
  int np;              /* planes */
  int nxb;             /* buffer planes: 2 nhalo np = 4 np */
  long sx, sy;         /* strides of the extended arrays */
  long nreal, nall;    /* sites without / with the buffer planes */
  const orc_box_t * box;
  const orc_le_t * le;
} lebox_t;

static void lebox_init(lebox_t * E, const orc_box_t * box, const orc_le_t * le) {
  for (int a = 0; a < 3; a++) E->n[a] = box->nlocal[a];
  E->np = le->nplanes;
  E->nxb = 4*le->nplanes;
  E->sy = E->n[2] + 4;
  E->sx = (long) (E->n[1] + 4)*E->sy;
  E->nreal = (long) (E->n[0] + 4)*E->sx;
  E->nall = (long) (E->n[0] + 4 + E->nxb)*E->sx;
  E->box = box; E->le = le;
}

/* lees_edw_index, src/leesedwards.c:767-781, nhalo = 2 */
static long le_idx(const lebox_t * E, int ic, int jc, int kc) {
  return (long) (ic + 1)*E->sx + (long) (jc + 1)*E->sy + (kc + 1);
}

/* lees_edw_ic_to_buff, src/leesedwards.c:1014-1047, nhalo = 2 */
static int le_ic_to_buff(const lebox_t * E, int ic, int di) {
  const int nh = 2;
  int p = ic/(E->n[0]/E->np);
  int ip;
  if (p > E->np - 1) p = E->np - 1;
  if (p < 0) p = 0;
  ip = orc_le_plane_location(E->box, E->le, p) - (nh - 1);
  if (di > 0 && (ic >= ip && ic < ip + nh) && (ic + di >= ip + nh))
    return E->n[0] + (1 + 2*p)*nh + (ic - ip + 1) + di;
  ip = orc_le_plane_location(E->box, E->le, p) + 1;
  if (di < 0 && (ic >= ip && ic < ip + nh) && (ic + di < ip))
    return E->n[0] + (2 + 2*p)*nh + (ic - ip + 1) + di;
  return ic + di;
}

/* displacement of buffer plane ib at time t: lees_edw_buffer_displacement
 * (src/leesedwards.c:633-655) with the sign of lees_edw_buffer_duy (:1065-1078) */
static int le_buffer_duy(int ib) { return (ib % 4 < 2) ? -1 : +1; }

static double le_buffer_dy(const lebox_t * E, int ib, double t) {
  const orc_le_t * le = E->le;
  double tle;
  if (t < 0.0) t = 0.0;
  tle = t - 1.0*le->nt0;
  if (le->oscillatory) {
    const double omega = 2.0*4.0*atan(1.0)/le->period;
    return le->uy*sin(omega*tle)/omega;
  }
  return tle*le->uy*le_buffer_duy(ib);
}

/* interior values of a halo-1 array into the extended layout, then every real site out
 * to two sites from the interior takes its periodic image (field_halo, width 2) */
static void le_fill(const lebox_t * E, int nf, const double * src, double * dst) {
  const orc_box_t * box = E->box;
  for (int ic = -1; ic <= E->n[0] + 2; ic++) {
    const int iw = 1 + (ic - 1 + 2*E->n[0]) % E->n[0];
    for (int jc = -1; jc <= E->n[1] + 2; jc++) {
      const int jw = 1 + (jc - 1 + 2*E->n[1]) % E->n[1];
      for (int kc = -1; kc <= E->n[2] + 2; kc++) {
	const int kw = 1 + (kc - 1 + 2*E->n[2]) % E->n[2];
	const long is = orc_index(box, iw, jw, kw), id = le_idx(E, ic, jc, kc);
	for (int a = 0; a < nf; a++) dst[nf*id + a] = src[nf*is + a];
      }
    }
  }
}

/* field_leesedwards, src/field.c:546-585 (nhalo = 2): buffer ib holds the real plane
 * lees_edw_ibuff_to_real(ib) (src/leesedwards.c:991-1005) displaced, four-point Lagrange */
static void le_field_buffers(const lebox_t * E, double t, double * phi) {
  const int ny = E->n[1];
  const int ib0 = E->n[0] + 2 + 1;
  const double r6 = (1.0/6.0);
  for (int ib = 0; ib < E->nxb; ib++) {
    const int ic = orc_le_plane_location(E->box, E->le, ib/4) - 1 + ib % 4;
    double dy = fmod(le_buffer_dy(E, ib, t), 1.0*ny);
    const int jdy = (int) floor(dy);
    const double fr = 1.0 - (dy - jdy);
    for (int jc = -1; jc <= ny + 2; jc++) {
      const int j0 = 1 + (jc - jdy - 3 + 2*ny) % ny;
      const int j1 = 1 + j0 % ny;
      const int j2 = 1 + j1 % ny;
      const int j3 = 1 + j2 % ny;
      for (int kc = -1; kc <= E->n[2] + 2; kc++) {
	phi[le_idx(E, ib0 + ib, jc, kc)] =
	  -  r6*fr*(fr-1.0)*(fr-2.0)*phi[le_idx(E, ic, j0, kc)]
	  + 0.5*(fr*fr-1.0)*(fr-2.0)*phi[le_idx(E, ic, j1, kc)]
	  - 0.5*fr*(fr+1.0)*(fr-2.0)*phi[le_idx(E, ic, j2, kc)]
	  +        r6*fr*(fr*fr-1.0)*phi[le_idx(E, ic, j3, kc)];
      }
    }
  }
}

/* one site of either stencil from the three x-planes it spans (im1, i0, ip1: indices of
 * the sites at x - 1, x, x + 1 with the same jc, kc), terms in the reference's order */
static void le_stencil(const lebox_t * E, int npt, const double * f, long im1, long i0, long ip1,
		       double * grad, double * delsq) {
  const long ys = E->sy;
  if (npt == 7) {
    grad[3*i0 + 0] = 0.5*(f[ip1] - f[im1]);
    grad[3*i0 + 1] = 0.5*(f[i0 + ys] - f[i0 - ys]);
    grad[3*i0 + 2] = 0.5*(f[i0 + 1] - f[i0 - 1]);
    delsq[i0] = f[ip1] + f[im1] + f[i0 + ys] + f[i0 - ys] + f[i0 + 1] + f[i0 - 1] - 6.0*f[i0];
  }
  else {
    const double r9 = (1.0/9.0);
    const long q[3] = {im1, i0, ip1};
    double gx = 0.0, gy = 0.0, gz = 0.0, d2 = 0.0;
    int first = 1;
    for (int b = -1; b <= 1; b++) {
      for (int c = -1; c <= 1; c++) {
	gx = first ? f[ip1 + b*ys + c] : gx + f[ip1 + b*ys + c];
	gx = gx - f[im1 + b*ys + c];
	first = 0;
      }
    }
    first = 1;
    for (int a = 0; a < 3; a++) {
      for (int c = -1; c <= 1; c++) {
	gy = first ? f[q[a] + ys + c] : gy + f[q[a] + ys + c];
	gy = gy - f[q[a] - ys + c];
	first = 0;
      }
    }
    first = 1;
    for (int a = 0; a < 3; a++) {
      for (int b = -1; b <= 1; b++) {
	gz = first ? f[q[a] + b*ys + 1] : gz + f[q[a] + b*ys + 1];
	gz = gz - f[q[a] + b*ys - 1];
	first = 0;
      }
    }
    first = 1;
    for (int a = 0; a < 3; a++) {
      for (int b = -1; b <= 1; b++) {
	for (int c = -1; c <= 1; c++) {
	  if (a == 1 && b == 0 && c == 0) continue;
	  d2 = first ? f[q[a] + b*ys + c] : d2 + f[q[a] + b*ys + c];
	  first = 0;
	}
      }
    }
    d2 = d2 - 26.0*f[i0];
    grad[3*i0 + 0] = 0.5*r9*gx;
    grad[3*i0 + 1] = 0.5*r9*gy;
    grad[3*i0 + 2] = 0.5*r9*gz;
    delsq[i0] = r9*d2;
  }
}

/* the operator over 0 .. n + 1 (nextra = 1) with plane-aware x neighbours, then the
 * first buffer plane on either side of every plane (grad_3d_*_fluid_le, nextra = 1) */
static void le_gradient(const lebox_t * E, int npt, const double * phi, double * grad, double * delsq) {
  for (int ic = 0; ic <= E->n[0] + 1; ic++) {
    const int icm1 = le_ic_to_buff(E, ic, -1), icp1 = le_ic_to_buff(E, ic, +1);
    for (int jc = 0; jc <= E->n[1] + 1; jc++) {
      for (int kc = 0; kc <= E->n[2] + 1; kc++) {
	le_stencil(E, npt, phi, le_idx(E, icm1, jc, kc), le_idx(E, ic, jc, kc), le_idx(E, icp1, jc, kc),
		   grad, delsq);
      }
    }
  }
  for (int p = 0; p < E->np; p++) {
    int ic = orc_le_plane_location(E->box, E->le, p);
    int ic0 = le_ic_to_buff(E, ic, 0), ic1 = le_ic_to_buff(E, ic, 1), ic2 = le_ic_to_buff(E, ic, 2);
    for (int jc = 0; jc <= E->n[1] + 1; jc++) {
      for (int kc = 0; kc <= E->n[2] + 1; kc++) {
	le_stencil(E, npt, phi, le_idx(E, ic0, jc, kc), le_idx(E, ic1, jc, kc), le_idx(E, ic2, jc, kc),
		   grad, delsq);
      }
    }
    ic += 1;
    ic2 = le_ic_to_buff(E, ic, 0); ic1 = le_ic_to_buff(E, ic, -1); ic0 = le_ic_to_buff(E, ic, -2);
    for (int jc = 0; jc <= E->n[1] + 1; jc++) {
      for (int kc = 0; kc <= E->n[2] + 1; kc++) {
	le_stencil(E, npt, phi, le_idx(E, ic0, jc, kc), le_idx(E, ic1, jc, kc), le_idx(E, ic2, jc, kc),
		   grad, delsq);
      }
    }
  }
}

/* fe_symm_str (src/symmetric.c:371-420) at one site of the extended arrays */
static void le_stress(const orc_param_t * prm, const double * phi, const double * grad,
		      const double * delsq, long index, double s[3][3]) {
  const double a = prm->a, b = prm->b, kappa = prm->kappa;
  const double * g = grad + 3*index;
  const double ph = phi[index], d2 = delsq[index];
  const double p0 = 0.5*a*ph*ph + 0.75*b*ph*ph*ph*ph - kappa*ph*d2
    - 0.5*kappa*(g[0]*g[0] + g[1]*g[1] + g[2]*g[2]);
  for (int ia = 0; ia < 3; ia++) {
    for (int ib = 0; ib < 3; ib++) {
      const double d = (ia == ib);
      s[ia][ib] = p0*d + kappa*g[ia]*g[ib];
    }
  }
}

/* phi_force_flux, src/phi_force.c:289-340; force [halo-1 layout] += */
static int le_force(const lebox_t * E, const orc_param_t * prm, const double * phi,
		    const double * grad, const double * delsq, double * force) {
  const long n = E->nreal;
  double * fluxe = (double *) calloc(12*n, sizeof(double));
  double * fluxw = fluxe + 3*n, * fluxy = fluxe + 6*n, * fluxz = fluxe + 9*n;
  if (fluxe == NULL) return -1;
  for (int ic = 1; ic <= E->n[0]; ic++) {
    const int icm1 = le_ic_to_buff(E, ic, -1), icp1 = le_ic_to_buff(E, ic, +1);
    for (int jc = 0; jc <= E->n[1]; jc++) {
      for (int kc = 0; kc <= E->n[2]; kc++) {
	const long index = le_idx(E, ic, jc, kc);
	double p0[3][3], p1[3][3];
	le_stress(prm, phi, grad, delsq, index, p0);
	le_stress(prm, phi, grad, delsq, le_idx(E, icm1, jc, kc), p1);
	for (int ia = 0; ia < 3; ia++) fluxw[3*index + ia] = 0.5*(p1[ia][0] + p0[ia][0]);
	le_stress(prm, phi, grad, delsq, le_idx(E, icp1, jc, kc), p1);
	for (int ia = 0; ia < 3; ia++) fluxe[3*index + ia] = 0.5*(p1[ia][0] + p0[ia][0]);
	le_stress(prm, phi, grad, delsq, le_idx(E, ic, jc + 1, kc), p1);
	for (int ia = 0; ia < 3; ia++) fluxy[3*index + ia] = 0.5*(p1[ia][1] + p0[ia][1]);
	le_stress(prm, phi, grad, delsq, le_idx(E, ic, jc, kc + 1), p1);
	for (int ia = 0; ia < 3; ia++) fluxz[3*index + ia] = 0.5*(p1[ia][2] + p0[ia][2]);
      }
    }
  }
  /* phi_force_flux_fix_local: the mismatch of the two fluxes through a plane, averaged
   * over the plane, is shared between them */
  {
    const double ra = 0.5/(1.0*E->n[1]*E->n[2]);
    for (int p = 0; p < E->np; p++) {
      const int ic = orc_le_plane_location(E->box, E->le, p);
      double fbar[3] = {0.0, 0.0, 0.0};
      for (int jc = 1; jc <= E->n[1]; jc++) {
	for (int kc = 1; kc <= E->n[2]; kc++) {
	  const long index = le_idx(E, ic, jc, kc), index1 = le_idx(E, ic + 1, jc, kc);
	  for (int ia = 0; ia < 3; ia++) fbar[ia] += - fluxe[3*index + ia] + fluxw[3*index1 + ia];
	}
      }
      for (int jc = 1; jc <= E->n[1]; jc++) {
	for (int kc = 1; kc <= E->n[2]; kc++) {
	  const long index = le_idx(E, ic, jc, kc), index1 = le_idx(E, ic + 1, jc, kc);
	  for (int ia = 0; ia < 3; ia++) {
	    fluxe[3*index + ia] += ra*fbar[ia];
	    fluxw[3*index1 + ia] -= ra*fbar[ia];
	  }
	}
      }
    }
  }
  for (int ic = 1; ic <= E->n[0]; ic++) {
    for (int jc = 1; jc <= E->n[1]; jc++) {
      for (int kc = 1; kc <= E->n[2]; kc++) {
	const long index = le_idx(E, ic, jc, kc);
	const long indexj = le_idx(E, ic, jc - 1, kc), indexk = le_idx(E, ic, jc, kc - 1);
	const long io = orc_index(E->box, ic, jc, kc);
	for (int ia = 0; ia < 3; ia++) {
	  const double f = -(+ fluxe[3*index + ia] - fluxw[3*index + ia]
			     + fluxy[3*index + ia] - fluxy[3*indexj + ia]
			     + fluxz[3*index + ia] - fluxz[3*indexk + ia]);
	  force[3*io + ia] += f;
	}
      }
    }
  }
  free(fluxe);
  return 0;
}

/* hydro_lees_edwards, src/hydro.c:449-492: the buffer planes of u at time t + 1, linear
 * interpolation plus the velocity jump (lees_edw_buffer_du, src/leesedwards.c:968-983) */
static void le_u_buffers(const lebox_t * E, double t, double * u) {
  const int ny = E->n[1];
  const int ib0 = E->n[0] + 2 + 1;
  for (int ib = 0; ib < E->nxb; ib++) {
    const int ic = orc_le_plane_location(E->box, E->le, ib/4) - 1 + ib % 4;
    const double ule[3] = {0.0, E->le->uy*le_buffer_duy(ib), 0.0};
    double dy = fmod(le_buffer_dy(E, ib, t + 1.0), 1.0*ny);
    const int jdy = (int) floor(dy);
    const double fr = dy - jdy;
    for (int jc = -1; jc <= ny + 2; jc++) {
      const int j1 = 1 + (jc - jdy - 2 + 2*ny) % ny;
      const int j2 = 1 + j1 % ny;
      for (int kc = 0; kc <= E->n[2] + 1; kc++) {
	const long i0 = le_idx(E, ib0 + ib, jc, kc);
	const long i1 = le_idx(E, ic, j1, kc), i2 = le_idx(E, ic, j2, kc);
	for (int ia = 0; ia < 3; ia++) u[3*i0 + ia] = ule[ia] + u[3*i1 + ia]*fr + u[3*i2 + ia]*(1.0 - fr);
      }
    }
  }
}

/* advection_le_{1st,2nd,3rd}: the faces of adv_face with plane-aware x neighbours;
 * fw, fe, fy, fz [nreal] */
static void le_advection(const lebox_t * E, int order, const double * u, const double * phi,
			 double * fw, double * fe, double * fy, double * fz) {
  for (int ic = 1; ic <= E->n[0]; ic++) {
    const int m2 = le_ic_to_buff(E, ic, -2), m1 = le_ic_to_buff(E, ic, -1);
    const int p1 = le_ic_to_buff(E, ic, +1), p2 = le_ic_to_buff(E, ic, +2);
    for (int jc = 0; jc <= E->n[1]; jc++) {
      for (int kc = 0; kc <= E->n[2]; kc++) {
	const long i0 = le_idx(E, ic, jc, kc);
	const double * u0 = u + 3*i0;
	const double ph = phi[i0];
	fw[i0] = adv_face(order, 1, u0[0], u[3*le_idx(E, m1, jc, kc) + 0],
			  phi[le_idx(E, p1, jc, kc)], ph, phi[le_idx(E, m1, jc, kc)], phi[le_idx(E, m2, jc, kc)]);
	fe[i0] = adv_face(order, 0, u0[0], u[3*le_idx(E, p1, jc, kc) + 0],
			  phi[le_idx(E, m1, jc, kc)], ph, phi[le_idx(E, p1, jc, kc)], phi[le_idx(E, p2, jc, kc)]);
	fy[i0] = adv_face(order, 0, u0[1], u[3*le_idx(E, ic, jc + 1, kc) + 1],
			  phi[le_idx(E, ic, jc - 1, kc)], ph, phi[le_idx(E, ic, jc + 1, kc)], phi[le_idx(E, ic, jc + 2, kc)]);
	fz[i0] = adv_face(order, 0, u0[2], u[3*le_idx(E, ic, jc, kc + 1) + 2],
			  phi[le_idx(E, ic, jc, kc - 1)], ph, phi[le_idx(E, ic, jc, kc + 1)], phi[le_idx(E, ic, jc, kc + 2)]);
      }
    }
  }
}

/* phi_ch_le_fix_fluxes, src/phi_cahn_hilliard.c:618-735; the displacement is
 * lees_edw_plane_dy = t u_y (src/leesedwards.c:923-935: steady planes, no time offset) */
static int le_fix_fluxes(const lebox_t * E, double t, double * fw, double * fe) {
  const int ny = E->n[1], nz = E->n[2];
  double * bufw = (double *) malloc(sizeof(double)*2*ny*nz);
  double * bufe = bufw + (long) ny*nz;
  if (bufw == NULL) return -1;
  for (int p = 0; p < E->np; p++) {
    const int ic = orc_le_plane_location(E->box, E->le, p);
    double dy, fr;
    int jdy;
    dy = fmod(+(t*E->le->uy), 1.0*ny);
    jdy = (int) floor(dy);
    fr = dy - jdy;
    for (int jc = 1; jc <= ny; jc++) {
      const int j1 = 1 + (jc - jdy - 2 + 2*ny) % ny;
      const int j2 = 1 + j1 % ny;
      for (int kc = 1; kc <= nz; kc++) {
	bufw[nz*(jc - 1) + (kc - 1)] = fw[le_idx(E, ic + 1, j1, kc)]*fr + fw[le_idx(E, ic + 1, j2, kc)]*(1.0 - fr);
      }
    }
    dy = fmod(-(t*E->le->uy), 1.0*ny);
    jdy = (int) floor(dy);
    fr = dy - jdy;
    for (int jc = 1; jc <= ny; jc++) {
      const int j1 = 1 + (jc - jdy - 2 + 2*ny) % ny;
      const int j2 = 1 + j1 % ny;
      for (int kc = 1; kc <= nz; kc++) {
	bufe[nz*(jc - 1) + (kc - 1)] = fe[le_idx(E, ic, j1, kc)]*fr + fe[le_idx(E, ic, j2, kc)]*(1.0 - fr);
      }
    }
    for (int jc = 1; jc <= ny; jc++) {
      for (int kc = 1; kc <= nz; kc++) {
	const long i0 = le_idx(E, ic, jc, kc), i1 = le_idx(E, ic + 1, jc, kc);
	fe[i0] = 0.5*(fe[i0] + bufw[nz*(jc - 1) + (kc - 1)]);
	fw[i1] = 0.5*(fw[i1] + bufe[nz*(jc - 1) + (kc - 1)]);
      }
    }
  }
  free(bufw);
  return 0;
}

/* One time step's worth of the order parameter with planes (src/ludwig.c:558-575,
 * 727-768): field_halo, field_grad_compute, phi_force_calculation, phi_cahn_hilliard.
 * t = physics_control_time. phi, u, force in the halo-1 layout of lb_oracle.h (interior
 * values are read and written); grad and delsq receive the interior of the gradient.
 * do_force / do_update select the stages (the gradient is always computed). */
int orc_ch_le_step(const orc_box_t * box, const orc_param_t * prm, const orc_ch_t * ch,
		   const orc_le_t * le, double t, int npt, int do_force, int do_update,
		   double * phi, const double * u, double * force, double * grad, double * delsq) {
  lebox_t E;
  double * pe, * ge, * de, * ue, * flux;
  int rc = 0;

  if (le->nplanes <= 0 || box->nlocal[0] % le->nplanes) return -1;
  if (ch->order < 1 || ch->order > 4) return -1;
  if (npt != 7 && npt != 27) return -1;
  lebox_init(&E, box, le);
  pe = (double *) calloc(8*E.nall + 4*E.nreal, sizeof(double));
  if (pe == NULL) return -1;
  ge = pe + E.nall; de = ge + 3*E.nall; ue = de + E.nall; flux = ue + 3*E.nall;

  le_fill(&E, 1, phi, pe);
  le_field_buffers(&E, t, pe);
  le_gradient(&E, npt, pe, ge, de);
  for (int ic = 1; ic <= E.n[0]; ic++) {
    for (int jc = 1; jc <= E.n[1]; jc++) {
      for (int kc = 1; kc <= E.n[2]; kc++) {
	const long io = orc_index(box, ic, jc, kc), ie = le_idx(&E, ic, jc, kc);
	for (int a = 0; a < 3; a++) grad[3*io + a] = ge[3*ie + a];
	delsq[io] = de[ie];
      }
    }
  }
  if (do_force && ch->method != ORC_FORCE_NONE) rc = le_force(&E, prm, pe, ge, de, force);
  if (rc == 0 && do_update) {
    double * fw = flux, * fe = flux + E.nreal, * fy = flux + 2*E.nreal, * fz = flux + 3*E.nreal;
    const double mobility = prm->mobility;
    const double wz = (E.n[2] == 1) ? 0.0 : 1.0;
    le_fill(&E, 3, u, ue);
    le_u_buffers(&E, t, ue);
    le_advection(&E, ch->order, ue, pe, fw, fe, fy, fz);
    for (int ic = 1; ic <= E.n[0]; ic++) {
      const int icm1 = le_ic_to_buff(&E, ic, -1), icp1 = le_ic_to_buff(&E, ic, +1);
      for (int jc = 0; jc <= E.n[1]; jc++) {
	for (int kc = 0; kc <= E.n[2]; kc++) {
	  const long i0 = le_idx(&E, ic, jc, kc);
	  long i1;
	  double mu0, mu1;
	  mu0 = mu_symm(prm, pe[i0], de[i0]);
	  i1 = le_idx(&E, icm1, jc, kc);
	  mu1 = mu_symm(prm, pe[i1], de[i1]);
	  fw[i0] -= mobility*(mu0 - mu1);
	  i1 = le_idx(&E, icp1, jc, kc);
	  mu1 = mu_symm(prm, pe[i1], de[i1]);
	  fe[i0] -= mobility*(mu1 - mu0);
	  i1 = le_idx(&E, ic, jc + 1, kc);
	  mu1 = mu_symm(prm, pe[i1], de[i1]);
	  fy[i0] -= mobility*(mu1 - mu0);
	  i1 = le_idx(&E, ic, jc, kc + 1);
	  mu1 = mu_symm(prm, pe[i1], de[i1]);
	  fz[i0] -= mobility*(mu1 - mu0);
	}
      }
    }
    for (int ic = 1; ic <= E.n[0]; ic++) {
      for (int jc = 0; jc <= E.n[1]; jc++) {
	for (int kc = 0; kc <= E.n[2]; kc++) {
	  const long i0 = le_idx(&E, ic, jc, kc);
	  fw[i0] -= mobility*ch->grad_mu[0];
	  fe[i0] -= mobility*ch->grad_mu[0];
	  fy[i0] -= mobility*ch->grad_mu[1];
	  fz[i0] -= mobility*ch->grad_mu[2];
	}
      }
    }
    rc = le_fix_fluxes(&E, t, fw, fe);
    for (int ic = 1; ic <= E.n[0] && rc == 0; ic++) {
      for (int jc = 1; jc <= E.n[1]; jc++) {
	for (int kc = 1; kc <= E.n[2]; kc++) {
	  const long index = le_idx(&E, ic, jc, kc);
	  const long io = orc_index(box, ic, jc, kc);
	  double ph = pe[index];
	  ph -= (+ fe[index] - fw[index] + fy[index] - fy[index - E.sy]
		 + wz*fz[index] - wz*fz[index - 1]);
	  phi[io] = ph;
	}
      }
    }
  }
  free(pe);
  return rc;
}

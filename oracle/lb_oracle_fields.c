/*
 * lb_oracle_fields.c -- TEST INFRASTRUCTURE ONLY. See lb_oracle.h.
 *
 * Binary (two-distribution) collision, halo, propagation, order parameter,
 * gradient, fills, diagnostics and the time loop of the CPU restatement.
 */

#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "lb_oracle.h"

void orc_relax_modes(const orc_model_t * model, const double * rtau,
		     const double force[3], const double sextra[3][3],
		     const double shat[3][3], const double * ghat,
		     double * m, double * rho_out, double u[3]);
void orc_f_to_modes(const orc_model_t * model, const double * fs, double * m);
void orc_modes_to_f(const orc_model_t * model, const double * m, double * fs);

/* ------------------------------------------------------------------------
 * orc_collide_binary: src/collision.c:720-1027 (lb_collision_mrt2_site),
 * with the symmetric free energy stress and chemical potential of
 * src/symmetric.c:371-420 and :307-320. ndist = 2, three dimensions only
 * (src/collision.c:618). Same extent as the single-fluid collision.
 * u is written; rho is NOT (src/collision.c:824-829).
 * ---------------------------------------------------------------------- */

int orc_collide_binary(const orc_model_t * model, const orc_box_t * box,
		       const orc_param_t * prm, double * f,
		       const double * force, double * u,
		       const double * phi, const double * grad,
		       const double * delsq) {

  const int nvel = model->nvel;
  const int ndist = 2;
  const double cs2 = 1.0/3.0;
  const long first = orc_index(box, 1, 0, 0);
  const long last  = orc_index(box, box->nlocal[0] + 1, 0, 0);
  const double rtau2 = 2.0/(1.0 + 2.0*prm->mobility);
  double rtau[ORC_NVELMAX];
  double kt = 0.0;
  unsigned int * nstate = orc_noise_attached(&kt);

  if (model->ndim != 3) return -1;
  if (orc_rtau(model, prm, rtau) != 0) return -1;
  if (nstate && nvel - model->nhydro > 10) return -1;
  if (prm->ghost_off) {
    /* lb_collision_noise_var_set, src/collision.c:1607-1614: lb->param->rtau, which
     * only this collision reads for the ghosts (src/collision.c:918) */
    for (int k = model->nhydro; k < nvel; k++) rtau[k] = 1.0;
  }

  #pragma omp parallel for schedule(static)
  for (long index = first; index < last; index++) {

    double m[ORC_NVELMAX];
    double fl[3], ul[3], rl;
    double sth[3][3];
    double * fs = f + (index*ndist + 0)*nvel;
    double * gs = f + (index*ndist + 1)*nvel;
    const double ph = phi[index];
    const double dsq = delsq[index];
    const double * gr = grad + 3*index;

    for (int a = 0; a < 3; a++) {
      fl[a] = prm->force_global[a] + force[3*index + a];
    }

    {
      double p0 = 0.5*prm->a*ph*ph + 0.75*prm->b*ph*ph*ph*ph
	- prm->kappa*ph*dsq
	- 0.5*prm->kappa*(gr[0]*gr[0] + gr[1]*gr[1] + gr[2]*gr[2]);
      for (int a = 0; a < 3; a++) {
	for (int b = 0; b < 3; b++) {
	  double dab = (a == b);
	  sth[a][b] = p0*dab + prm->kappa*gr[a]*gr[b];
	}
      }
    }

    orc_f_to_modes(model, fs, m);
    if (nstate) {
      /* src/collision.c:884-900 */
      double shat[3][3], ghat[ORC_NVELMAX];
      orc_fluctuations(model, rtau, kt, prm->ghost_off, nstate, index, shat, ghat);
      orc_relax_modes(model, rtau, fl, (const double (*)[3]) sth,
		      (const double (*)[3]) shat, ghat, m, &rl, ul);
    }
    else {
      orc_relax_modes(model, rtau, fl, (const double (*)[3]) sth, NULL, NULL, m, &rl, ul);
    }
    orc_modes_to_f(model, m, fs);
    for (int a = 0; a < 3; a++) u[3*index + a] = ul[a];

    /* Order parameter distribution */
    {
      double mu = prm->a*ph + prm->b*ph*ph*ph - prm->kappa*dsq;
      double jphi[3] = {0.0, 0.0, 0.0};
      double sphi[3][3];

      for (int p = 1; p < nvel; p++) {
	for (int a = 0; a < 3; a++) jphi[a] += model->cv[p][a]*gs[p];
      }

      for (int a = 0; a < 3; a++) {
	for (int b = 0; b < 3; b++) {
	  double dab = (a == b);
	  sphi[a][b] = ph*ul[a]*ul[b] + mu*dab;
	}
	jphi[a] = jphi[a] - rtau2*(jphi[a] - ph*ul[a]);
      }

      for (int p = 0; p < nvel; p++) {
	double jdotc = 0.0;
	double sphidotq = 0.0;
	for (int a = 0; a < 3; a++) {
	  jdotc += jphi[a]*model->cv[p][a];
	  for (int b = 0; b < 3; b++) {
	    double dab = (a == b);
	    sphidotq += sphi[a][b]*(model->cv[p][a]*model->cv[p][b] - cs2*dab);
	  }
	}
	gs[p] = model->wv[p]*(jdotc*3.0 + sphidotq*4.5) + ph*(p == 0);
      }
    }
  }

  return 0;
}

/* ------------------------------------------------------------------------
 * orc_halo_periodic: one-rank periodic halo of width 1 for a field with
 * nfel doubles per site. Restates the net effect of halo_swap_packed for
 * cartsz = 1 in every direction (src/halo_swap.c:709-1063): x faces over the
 * full y,z extent, then y over full x,z, then z over full x,y, so edges and
 * corners come for free. Pure copies: bit-exact.
 * ---------------------------------------------------------------------- */

int orc_halo_periodic(const orc_box_t * box, int nfel, double * data) {

  const int nx = box->nlocal[0], ny = box->nlocal[1], nz = box->nlocal[2];
  const size_t sz = (size_t) nfel*sizeof(double);

  for (int j = 0; j < box->nall[1]; j++) {
    for (int k = 0; k < box->nall[2]; k++) {
      memcpy(data + nfel*orc_index(box, 0, j, k),
	     data + nfel*orc_index(box, nx, j, k), sz);
      memcpy(data + nfel*orc_index(box, nx + 1, j, k),
	     data + nfel*orc_index(box, 1, j, k), sz);
    }
  }
  for (int i = 0; i < box->nall[0]; i++) {
    for (int k = 0; k < box->nall[2]; k++) {
      memcpy(data + nfel*orc_index(box, i, 0, k),
	     data + nfel*orc_index(box, i, ny, k), sz);
      memcpy(data + nfel*orc_index(box, i, ny + 1, k),
	     data + nfel*orc_index(box, i, 1, k), sz);
    }
  }
  for (int i = 0; i < box->nall[0]; i++) {
    for (int j = 0; j < box->nall[1]; j++) {
      memcpy(data + nfel*orc_index(box, i, j, 0),
	     data + nfel*orc_index(box, i, j, nz), sz);
      memcpy(data + nfel*orc_index(box, i, j, nz + 1),
	     data + nfel*orc_index(box, i, j, 1), sz);
    }
  }

  return 0;
}

/* ------------------------------------------------------------------------
 * orc_propagate: src/propagation.c:162-212. Pull: interior sites take
 * f(r - c_p); y,z halo sites of planes 1..nlocal[X] copy themselves; the
 * two x halo planes of fprime are not written. Bit-exact.
 * ---------------------------------------------------------------------- */

int orc_propagate(const orc_model_t * model, const orc_box_t * box,
		  int ndist, const double * f, double * fprime) {

  const int nvel = model->nvel;
  const long sx = (long) box->nall[1]*box->nall[2];
  const long sy = box->nall[2];

  #pragma omp parallel for schedule(static)
  for (int ic = 1; ic <= box->nlocal[0]; ic++) {
    for (int jc = 0; jc < box->nall[1]; jc++) {
      for (int kc = 0; kc < box->nall[2]; kc++) {
	long index = orc_index(box, ic, jc, kc);
	int inside = (jc >= 1 && jc <= box->nlocal[1] &&
		      kc >= 1 && kc <= box->nlocal[2]);
	for (int n = 0; n < ndist; n++) {
	  for (int p = 0; p < nvel; p++) {
	    long from = index;
	    if (inside) {
	      from -= model->cv[p][0]*sx + model->cv[p][1]*sy + model->cv[p][2];
	    }
	    fprime[(index*ndist + n)*nvel + p] = f[(from*ndist + n)*nvel + p];
	  }
	}
      }
    }
  }

  return 0;
}

/* src/phi_lb_coupler.c:72-102: phi = sum_p g_p, p in index order, interior */

int orc_phi_from_g(const orc_model_t * model, const orc_box_t * box,
		   const double * f, double * phi) {

  const int nvel = model->nvel;

  #pragma omp parallel for schedule(static)
  for (int ic = 1; ic <= box->nlocal[0]; ic++) {
    for (int jc = 1; jc <= box->nlocal[1]; jc++) {
      for (int kc = 1; kc <= box->nlocal[2]; kc++) {
	long index = orc_index(box, ic, jc, kc);
	double phi0 = 0.0;
	for (int p = 0; p < nvel; p++) phi0 += f[(index*2 + 1)*nvel + p];
	phi[index] = phi0;
      }
    }
  }
  return 0;
}

/* src/gradient_3d_7pt_fluid.c:232-313 with nextra = 0: central differences
 * and the 7-point Laplacian summed in the order +x -x +y -y +z -z -6 phi. */

int orc_grad_7pt(const orc_box_t * box, const double * phi,
		 double * grad, double * delsq) {

  const long sx = (long) box->nall[1]*box->nall[2];
  const long sy = box->nall[2];

  #pragma omp parallel for schedule(static)
  for (int ic = 1; ic <= box->nlocal[0]; ic++) {
    for (int jc = 1; jc <= box->nlocal[1]; jc++) {
      for (int kc = 1; kc <= box->nlocal[2]; kc++) {
	long i = orc_index(box, ic, jc, kc);
	grad[3*i + 0] = 0.5*(phi[i + sx] - phi[i - sx]);
	grad[3*i + 1] = 0.5*(phi[i + sy] - phi[i - sy]);
	grad[3*i + 2] = 0.5*(phi[i + 1] - phi[i - 1]);
	delsq[i] = phi[i + sx] + phi[i - sx] + phi[i + sy] + phi[i - sy]
	  + phi[i + 1] + phi[i - 1] - 6.0*phi[i];
      }
    }
  }
  return 0;
}

/* src/gradient_3d_27pt_fluid.c:223-352 (GRAD_DEL2, nextra = 0). The 27-point
 * box stencil: each first derivative is the sum of nine central differences
 * scaled by (1/2)(1/9); the Laplacian is (1/9)(sum of 26 neighbours - 26 phi).
 * Floating-point sums are not associative, so the loops below visit the
 * terms in the order the reference writes them:
 *   d/dx : (dy,dz) ascending, +phi(x+1) then -phi(x-1) for each pair
 *   d/dy : dx ascending then dz ascending, +phi(y+1) then -phi(y-1)
 *   d/dz : dx ascending then dy ascending, +phi(z+1) then -phi(z-1)
 *   delsq: (dx,dy,dz) ascending without the centre, centre term last. */

static int grad_select_npt = 7;

int orc_gradient_select(int npt) {
  if (npt != 7 && npt != 27 && npt != 28) return -1;      /* 28: the 27-point stencil next to walls */
  grad_select_npt = npt;
  return 0;
}

int orc_grad_27pt(const orc_box_t * box, const double * phi,
		  double * grad, double * delsq) {

  const long sx = (long) box->nall[1]*box->nall[2];
  const long sy = box->nall[2];
  const double r9 = (1.0/9.0);

  #pragma omp parallel for schedule(static)
  for (int ic = 1; ic <= box->nlocal[0]; ic++) {
    for (int jc = 1; jc <= box->nlocal[1]; jc++) {
      for (int kc = 1; kc <= box->nlocal[2]; kc++) {
	const long i = orc_index(box, ic, jc, kc);
	double gx = 0.0, gy = 0.0, gz = 0.0, d2 = 0.0;
	int first;

	first = 1;
	for (int b = -1; b <= 1; b++) {
	  for (int c = -1; c <= 1; c++) {
	    const long o = b*sy + c;
	    gx = first ? phi[i + sx + o] : gx + phi[i + sx + o];
	    gx = gx - phi[i - sx + o];
	    first = 0;
	  }
	}
	first = 1;
	for (int a = -1; a <= 1; a++) {
	  for (int c = -1; c <= 1; c++) {
	    const long o = a*sx + c;
	    gy = first ? phi[i + sy + o] : gy + phi[i + sy + o];
	    gy = gy - phi[i - sy + o];
	    first = 0;
	  }
	}
	first = 1;
	for (int a = -1; a <= 1; a++) {
	  for (int b = -1; b <= 1; b++) {
	    const long o = a*sx + b*sy;
	    gz = first ? phi[i + 1 + o] : gz + phi[i + 1 + o];
	    gz = gz - phi[i - 1 + o];
	    first = 0;
	  }
	}
	first = 1;
	for (int a = -1; a <= 1; a++) {
	  for (int b = -1; b <= 1; b++) {
	    for (int c = -1; c <= 1; c++) {
	      if (a == 0 && b == 0 && c == 0) continue;
	      d2 = first ? phi[i + a*sx + b*sy + c] : d2 + phi[i + a*sx + b*sy + c];
	      first = 0;
	    }
	  }
	}
	d2 = d2 - 26.0*phi[i];

	grad[3*i + 0] = 0.5*r9*gx;
	grad[3*i + 1] = 0.5*r9*gy;
	grad[3*i + 2] = 0.5*r9*gz;
	delsq[i] = r9*d2;
      }
    }
  }
  return 0;
}

/* src/hydro.c:279-398: hydro_u_zero / hydro_f_zero fill ALL sites */

int orc_fill(const orc_box_t * box, int nf, const double * v, double * data) {
  #pragma omp parallel for schedule(static)
  for (long i = 0; i < box->nsite; i++) {
    for (int a = 0; a < nf; a++) data[nf*i + a] = v[a];
  }
  return 0;
}

/* src/stats_distribution.c:55-118 and :201-265 (plain sums here) */

int orc_stats(const orc_model_t * model, const orc_box_t * box, int ndist,
	      const double * f, double stats[8]) {

  const int nvel = model->nvel;
  double sr = 0.0, sr2 = 0.0, rmin = 1.0e300, rmax = -1.0e300;
  double g[3] = {0.0, 0.0, 0.0}, sphi = 0.0;

  for (int ic = 1; ic <= box->nlocal[0]; ic++) {
    for (int jc = 1; jc <= box->nlocal[1]; jc++) {
      for (int kc = 1; kc <= box->nlocal[2]; kc++) {
	long index = orc_index(box, ic, jc, kc);
	double rho = 0.0;
	for (int p = 0; p < nvel; p++) {
	  double fp = f[(index*ndist + 0)*nvel + p];
	  rho += fp;
	  for (int a = 0; a < 3; a++) g[a] += model->cv[p][a]*fp;
	}
	sr += rho; sr2 += rho*rho;
	if (rho < rmin) rmin = rho;
	if (rho > rmax) rmax = rho;
	if (ndist == 2) {
	  for (int p = 0; p < nvel; p++) sphi += f[(index*ndist + 1)*nvel + p];
	}
      }
    }
  }
  stats[0] = sr; stats[1] = sr2; stats[2] = rmin; stats[3] = rmax;
  stats[4] = g[0]; stats[5] = g[1]; stats[6] = g[2]; stats[7] = sphi;
  return 0;
}

/* ------------------------------------------------------------------------
 * orc_step: the time loop of src/ludwig.c:530-862 restricted to the hot
 * path (periodic, no colloids/walls/LE planes):
 *   hydro_f_zero; [phi_lb_to_field; field_halo; grad]; hydro_u_zero;
 *   lb_collide; lb_halo; lb_propagation (+ pointer swap).
 * ---------------------------------------------------------------------- */

int orc_step(const orc_model_t * model, const orc_box_t * box,
	     const orc_param_t * prm, int ndist, int nsteps,
	     double * f, double * fprime, double * rho, double * u,
	     double * force, double * phi, double * grad, double * delsq) {

  const double zero[3] = {0.0, 0.0, 0.0};
  const size_t nbytes = (size_t) box->nsite*ndist*model->nvel*sizeof(double);
  double * a = f;
  double * b = fprime;

  for (int n = 0; n < nsteps; n++) {
    double * t;
    if (orc_le_attached()) orc_le_attached()->t_current += 1;
    orc_fill(box, 3, zero, force);
    if (ndist == 2) {
      orc_phi_from_g(model, box, a, phi);
      orc_halo_periodic(box, 1, phi);
      if (orc_le_attached()) {
	/* field_grad_compute: field_leesedwards, then the stencil across the planes */
	const orc_le_t * le = orc_le_attached();
	const double t = 1.0*(le->t_start + le->t_current - 1.0);
	if (orc_le_field_grad(box, le, t, grad_select_npt, phi, grad, delsq)) return -1;
      }
      else if (grad_select_npt == 28) orc_grad_27pt_solid(box, orc_wall_attached(), phi, grad, delsq);
      else if (grad_select_npt == 27) orc_grad_27pt(box, phi, grad, delsq);
      else orc_grad_7pt(box, phi, grad, delsq);
    }
    if (ndist == 1 && orc_ch_attached()) {
      /* finite-difference order parameter ("free_energy symmetric"), src/ludwig.c:558-575,727-768 */
      const orc_ch_t * ch = orc_ch_attached();
      if (orc_le_attached()) {
	const orc_le_t * le = orc_le_attached();
	const double t = 1.0*(le->t_start + le->t_current - 1.0);
	if (orc_ch_le_step(box, prm, ch, le, t, grad_select_npt, 1, 1, phi, u, force, grad, delsq)) return -1;
      }
      else {
      orc_halo_periodic(box, 1, phi);
      if (grad_select_npt == 28) orc_grad_27pt_solid(box, orc_wall_attached(), phi, grad, delsq);
      else if (grad_select_npt == 27) orc_grad_27pt(box, phi, grad, delsq);
      else orc_grad_7pt(box, phi, grad, delsq);
      if (orc_phi_force(box, prm, ch, phi, grad, delsq, force)) return -1;
      if (orc_cahn_hilliard(box, prm, ch, phi, delsq, u)) return -1;
      }
    }
    orc_fill(box, 3, zero, u);
    {
      const double * eta = NULL;
      const orc_visc_t * visc = orc_visc_attached(&eta);
      /* src/ludwig.c:794: the model's update ahead of the collision */
      if (visc && eta && phi) orc_visc_arrhenius_update(box, visc, phi, (double *) eta);
    }
    if (ndist == 1) {
      if (orc_collide_mrt(model, box, prm, 1, a, force, rho, u)) return -1;
    }
    else {
      if (orc_collide_binary(model, box, prm, a, force, u, phi, grad, delsq)) {
	return -1;
      }
    }
    if (orc_le_attached()) {
      orc_le_t * le = orc_le_attached();
      if (orc_le_apply(model, box, ndist, le, 1.0*le->t_current, a)) return -1;
    }
    orc_halo_periodic(box, ndist*model->nvel, a);
    if (orc_open_attached()) {
      if (orc_open_apply(model, box, ndist, orc_open_attached(), rho, u, a)) return -1;
    }
    if (orc_wall_attached()) {
      /* src/ludwig.c:837-842. The periodic copies above also filled the wall layers,
       * which the reference leaves alone; every population a fluid site pulls out of
       * a wall layer is overwritten here, so it makes no difference. */
      orc_wall_bbl(model, box, prm, ndist, orc_wall_attached(), a);
    }
    orc_propagate(model, box, ndist, a, b);
    t = a; a = b; b = t;
  }

  if (a != f) {
    /* Odd number of steps: result is in the caller's fprime. Swap contents
     * so that f holds the current state and fprime the previous one. */
    double * tmp = (double *) malloc(nbytes);
    if (tmp == NULL) return -2;
    memcpy(tmp, f, nbytes);
    memcpy(f, fprime, nbytes);
    memcpy(fprime, tmp, nbytes);
    free(tmp);
  }

  return 0;
}

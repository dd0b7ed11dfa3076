/*
 * lb_oracle_le.c -- TEST INFRASTRUCTURE ONLY. See lb_oracle.h.
 *
 * Lees-Edwards sliding periodic boundaries for the distributions:
 * lb_le_apply_boundary_conditions, src/model_le.c:57-96, applied to the
 * post-collision populations between lb_collide and lb_halo (src/ludwig.c:810).
 *
 * nplanes planes normal to x, half a spacing from the ends and a spacing
 * ntotal_x/nplanes apart (src/leesedwards.c:240-241,599-617), move with velocity
 * jump uy along y. A population about to cross a plane is
 *  1. reprojected (src/model_le.c:113-208): the momentum and stress of its site
 *     seen from the other side of the plane differ by the jump; that difference is
 *     added through the equilibrium projection, for the crossing velocities only;
 *  2. displaced along y by the distance the plane has slid so far, with linear
 *     interpolation between the two sites it falls between (src/model_le.c:220-390).
 * One rank, as the rest of the oracle (the reference's own one-rank branch).
 */

#include <math.h>
#include <stdlib.h>

#include "lb_oracle.h"

static orc_le_t * attached_le = NULL;

int orc_le_attach(orc_le_t * le) {
  attached_le = le;
  return 0;
}

orc_le_t * orc_le_attached(void) { return attached_le; }

/* local x coordinate (the plane sits between it and the next), src/leesedwards.c:599-617 */
int orc_le_plane_location(const orc_box_t * box, const orc_le_t * le, int np) {
  const double dx_sep = 1.0*box->nlocal[0]/le->nplanes;
  const double dx_min = 0.5*dx_sep;
  return (int) (dx_min + np*dx_sep);
}

/* velocity jump and accumulated displacement at time t (src/leesedwards.c:572-587,633-655) */
static void le_now(const orc_le_t * le, double t, double * uy, double * dy) {
  const double tle = ((t < 0.0) ? 0.0 : t) - 1.0*le->nt0;
  *uy = le->uy;
  *dy = tle*le->uy;
  if (le->oscillatory) {
    const double omega = 2.0*4.0*atan(1.0)/le->period;
    *uy = le->uy*cos(omega*(t - 1.0*le->nt0));
    *dy = le->uy*sin(omega*tle)/omega;
  }
}

int orc_le_apply(const orc_model_t * model, const orc_box_t * box, int ndist,
		 const orc_le_t * le, double t, double * f) {

  const int nvel = model->nvel;
  const int ny = box->nlocal[1], nz = box->nlocal[2];
  const double cs2 = 1.0/3.0, rcs2 = 3.0;
  double uy_now, dy_now;
  double * buf;
  int nprop = 0;

  if (le->nplanes <= 0) return 0;
  if (box->nlocal[0] % le->nplanes) return -1;
  le_now(le, t, &uy_now, &dy_now);

  /* 1. reprojection */
  for (int plane = 0; plane < le->nplanes; plane++) {
    for (int side = 0; side < 2; side++) {
      const int ic = orc_le_plane_location(box, le, plane) + side;
      const int cx = (side == 0) ? +1 : -1;
      const double du[3] = {0.0, (side == 0) ? -1.0*uy_now : uy_now, 0.0};
      for (int jc = 1; jc <= ny; jc++) {
	for (int kc = 1; kc <= nz; kc++) {
	  const long index = orc_index(box, ic, jc, kc);
	  for (int n = 0; n < ndist; n++) {
	    double * fs = f + (index*ndist + n)*nvel;
	    double rho = 0.0, g[3] = {0.0, 0.0, 0.0}, ds[3][3];
	    for (int p = 0; p < nvel; p++) {
	      rho += fs[p];
	      for (int a = 0; a < 3; a++) g[a] += model->cv[p][a]*fs[p];
	    }
	    for (int a = 0; a < 3; a++) {
	      for (int b = 0; b < 3; b++) {
		ds[a][b] = g[a]*du[b] + du[a]*g[b] + rho*du[a]*du[b];
	      }
	    }
	    for (int p = 1; p < nvel; p++) {
	      double udotc, sdotq = 0.0;
	      if (model->cv[p][0] != cx) continue;
	      udotc = du[1]*model->cv[p][1];
	      for (int a = 0; a < 3; a++) {
		for (int b = 0; b < 3; b++) {
		  double dab = cs2*(a == b);
		  sdotq += ds[a][b]*(model->cv[p][a]*model->cv[p][b] - dab);
		}
	      }
	      fs[p] += model->wv[p]*(rho*udotc*rcs2 + 0.5*sdotq*rcs2*rcs2);
	    }
	  }
	}
      }
    }
  }

  /* 2. displacement along y with linear interpolation, through a buffer */
  for (int p = 1; p < nvel; p++) nprop += (model->cv[p][0] == +1);
  buf = (double *) malloc(sizeof(double)*ndist*nprop*ny*nz);
  if (buf == NULL) return -1;

  for (int plane = 0; plane < le->nplanes; plane++) {
    for (int side = 0; side < 2; side++) {
      const int ic = orc_le_plane_location(box, le, plane) + side;
      const int cx = (side == 0) ? +1 : -1;
      const double dy = fmod((side == 0) ? dy_now : -dy_now, 1.0*ny);
      const int jdy = (int) floor(dy);
      const double fr = dy - jdy;
      long nd = 0;
      for (int jc = 1; jc <= ny; jc++) {
	const int j1 = 1 + (jc + jdy - 1 + 2*ny) % ny;
	const int j2 = 1 + (j1 % ny);
	for (int kc = 1; kc <= nz; kc++) {
	  const long i0 = orc_index(box, ic, j1, kc), i1 = orc_index(box, ic, j2, kc);
	  for (int n = 0; n < ndist; n++) {
	    for (int p = 1; p < nvel; p++) {
	      if (model->cv[p][0] != cx) continue;
	      buf[nd++] = (1.0 - fr)*f[(i0*ndist + n)*nvel + p] + fr*f[(i1*ndist + n)*nvel + p];
	    }
	  }
	}
      }
      nd = 0;
      for (int jc = 1; jc <= ny; jc++) {
	for (int kc = 1; kc <= nz; kc++) {
	  const long i0 = orc_index(box, ic, jc, kc);
	  for (int n = 0; n < ndist; n++) {
	    for (int p = 1; p < nvel; p++) {
	      if (model->cv[p][0] != cx) continue;
	      f[(i0*ndist + n)*nvel + p] = buf[nd++];
	    }
	  }
	}
      }
    }
  }

  free(buf);
  return 0;
}

/* ---------------------------------------------------------------------------
 * The order parameter across the planes (binary fluid, one-site halo).
 *
 * field_grad_compute first calls field_leesedwards (src/field_grad.c:310,
 * src/field.c:500-600): for every plane two buffer planes of phi, one per side,
 *   buffer 2p     the plane of sites x = loc(p),     seen from x = loc(p) + 1
 *   buffer 2p + 1 the plane of sites x = loc(p) + 1, seen from x = loc(p)
 * (lees_edw_ibuff_to_real, src/leesedwards.c:991-1005), displaced along y by what the
 * plane has slid so far, -/+ u_y t (lees_edw_buffer_duy :1065-1078,
 * lees_edw_buffer_displacement :633-655), with four-point Lagrange interpolation.
 * The gradient stencils then take the x - 1 / x + 1 neighbours of the sites next to
 * a plane from those buffers (lees_edw_ic_to_buff :1014-1047 as called from
 * src/gradient_3d_7pt_fluid.c:264-265 and src/gradient_3d_27pt_fluid.c:262-265);
 * with a one-site halo nothing else changes (the loops of grad_3d_*_fluid_le run over
 * nextra = nhalo - 1 = 0 planes).
 *
 * buf[2 nplanes][nall_y][nall_z]; one rank (the reference's own branch for a
 * y-extent of one rank, src/field.c:540-585). phi needs current y and z halos.
 * --------------------------------------------------------------------------- */

int orc_le_field_buffers(const orc_box_t * box, const orc_le_t * le, double t,
			 const double * phi, double * buf) {

  const int ny = box->nlocal[1], nz = box->nlocal[2];
  const long sx = (long) box->nall[1]*box->nall[2];
  const long sy = box->nall[2];
  const double r6 = (1.0/6.0);

  if (le->nplanes <= 0) return 0;
  if (box->nlocal[0] % le->nplanes) return -1;

  for (int ib = 0; ib < 2*le->nplanes; ib++) {
    const int ic = orc_le_plane_location(box, le, ib/2) + ib % 2;
    const double tle = ((t < 0.0) ? 0.0 : t) - 1.0*le->nt0;
    double dy = tle*le->uy*((ib % 2 < 1) ? -1 : +1);
    int jdy;
    double fr;
    if (le->oscillatory) {
      const double omega = 2.0*4.0*atan(1.0)/le->period;
      dy = le->uy*sin(omega*tle)/omega;
    }
    dy = fmod(dy, 1.0*ny);
    jdy = (int) floor(dy);
    fr = 1.0 - (dy - jdy);
    for (int jc = 0; jc <= ny + 1; jc++) {
      const int j0 = 1 + (jc - jdy - 3 + 2*ny) % ny;
      const int j1 = 1 + j0 % ny;
      const int j2 = 1 + j1 % ny;
      const int j3 = 1 + j2 % ny;
      for (int kc = 0; kc <= nz + 1; kc++) {
	buf[ib*sx + jc*sy + kc] =
	  -  r6*fr*(fr-1.0)*(fr-2.0)*phi[orc_index(box, ic, j0, kc)]
	  + 0.5*(fr*fr-1.0)*(fr-2.0)*phi[orc_index(box, ic, j1, kc)]
	  - 0.5*fr*(fr+1.0)*(fr-2.0)*phi[orc_index(box, ic, j2, kc)]
	  +        r6*fr*(fr*fr-1.0)*phi[orc_index(box, ic, j3, kc)];
      }
    }
  }
  return 0;
}

/* the x-plane (its element jc = 0, kc = 0) that a site at ic sees at ic + di */
static const double * le_xplane(const orc_box_t * box, const orc_le_t * le,
				const double * phi, const double * buf, int ic, int di) {
  const long sx = (long) box->nall[1]*box->nall[2];
  /* one candidate plane per x, as lees_edw_ic_to_buff finds it (src/leesedwards.c:1028-1029) */
  int p = ic/(box->nlocal[0]/le->nplanes);
  if (p > le->nplanes - 1) p = le->nplanes - 1;
  if (p < 0) p = 0;
  {
    const int loc = orc_le_plane_location(box, le, p);
    if (di == +1 && ic == loc) return buf + (2*p + 1)*sx;
    if (di == -1 && ic == loc + 1) return buf + (2*p)*sx;
  }
  return phi + (ic + di)*sx;
}

/* field_grad_compute with planes: npt = 7 (src/gradient_3d_7pt_fluid.c:232-313) or
 * 27 (src/gradient_3d_27pt_fluid.c:216-365), terms in the reference's order */
int orc_le_field_grad(const orc_box_t * box, const orc_le_t * le, double t, int npt,
		      const double * phi, double * grad, double * delsq) {

  const long sx = (long) box->nall[1]*box->nall[2];
  const long sy = box->nall[2];
  const double r9 = (1.0/9.0);
  double * buf;

  if (npt != 7 && npt != 27) return -1;
  if (le->nplanes <= 0 || box->nlocal[0] % le->nplanes) return -1;
  buf = (double *) malloc(sizeof(double)*2*le->nplanes*sx);
  if (buf == NULL) return -1;
  if (orc_le_field_buffers(box, le, t, phi, buf)) { free(buf); return -1; }

  for (int ic = 1; ic <= box->nlocal[0]; ic++) {
    const double * pl[3];
    pl[0] = le_xplane(box, le, phi, buf, ic, -1);
    pl[1] = phi + ic*sx;
    pl[2] = le_xplane(box, le, phi, buf, ic, +1);
    for (int jc = 1; jc <= box->nlocal[1]; jc++) {
      for (int kc = 1; kc <= box->nlocal[2]; kc++) {
	const long i = orc_index(box, ic, jc, kc);
	const long o = jc*sy + kc;
	const double * xm = pl[0] + o, * x0 = pl[1] + o, * xp = pl[2] + o;
	if (npt == 7) {
	  grad[3*i + 0] = 0.5*(xp[0] - xm[0]);
	  grad[3*i + 1] = 0.5*(x0[sy] - x0[-sy]);
	  grad[3*i + 2] = 0.5*(x0[1] - x0[-1]);
	  delsq[i] = xp[0] + xm[0] + x0[sy] + x0[-sy] + x0[1] + x0[-1] - 6.0*x0[0];
	}
	else {
	  const double * q[3] = {xm, x0, xp};
	  double gx = 0.0, gy = 0.0, gz = 0.0, d2 = 0.0;
	  int first;
	  first = 1;
	  for (int b = -1; b <= 1; b++) {
	    for (int c = -1; c <= 1; c++) {
	      gx = first ? xp[b*sy + c] : gx + xp[b*sy + c];
	      gx = gx - xm[b*sy + c];
	      first = 0;
	    }
	  }
	  first = 1;
	  for (int a = 0; a < 3; a++) {
	    for (int c = -1; c <= 1; c++) {
	      gy = first ? q[a][sy + c] : gy + q[a][sy + c];
	      gy = gy - q[a][-sy + c];
	      first = 0;
	    }
	  }
	  first = 1;
	  for (int a = 0; a < 3; a++) {
	    for (int b = -1; b <= 1; b++) {
	      gz = first ? q[a][b*sy + 1] : gz + q[a][b*sy + 1];
	      gz = gz - q[a][b*sy - 1];
	      first = 0;
	    }
	  }
	  first = 1;
	  for (int a = 0; a < 3; a++) {
	    for (int b = -1; b <= 1; b++) {
	      for (int c = -1; c <= 1; c++) {
		if (a == 1 && b == 0 && c == 0) continue;
		d2 = first ? q[a][b*sy + c] : d2 + q[a][b*sy + c];
		first = 0;
	      }
	    }
	  }
	  d2 = d2 - 26.0*x0[0];
	  grad[3*i + 0] = 0.5*r9*gx;
	  grad[3*i + 1] = 0.5*r9*gy;
	  grad[3*i + 2] = 0.5*r9*gz;
	  delsq[i] = r9*d2;
	}
      }
    }
  }
  free(buf);
  return 0;
}

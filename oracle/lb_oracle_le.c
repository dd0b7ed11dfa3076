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

/*
 * lb_oracle_wall.c -- TEST INFRASTRUCTURE ONLY. See lb_oracle.h.
 *
 * Flat no-slip walls: bounce-back on links, src/wall.c.
 *
 * A wall in direction a occupies the halo layer at global coordinate 0 and
 * ntotal[a] + 1 (wall_init_map, src/wall.c:1219-1268). A link is a fluid site
 * i and a velocity p with i + c_p inside a wall (wall_init_boundaries,
 * src/wall.c:381-462). Between lb_halo and lb_propagation (src/ludwig.c:837-842)
 * wall_bbl (src/wall.c:997-1107) writes into the solid site j = i + c_p, for the
 * reversed velocity, the population that left i along p -- less the momentum a
 * moving wall feeds in -- so that the propagation brings it back to i; and
 * accumulates the momentum handed to the wall.
 *
 * One rank only (local = global coordinates), like the rest of the oracle.
 */

#include <stddef.h>

#include "lb_oracle.h"

static orc_wall_t * attached_wall = NULL;

int orc_wall_attach(orc_wall_t * wall) {
  attached_wall = wall;
  return 0;
}

orc_wall_t * orc_wall_attached(void) { return attached_wall; }

static int in_wall(const orc_box_t * box, const orc_wall_t * wall, const int r[3]) {
  for (int a = 0; a < 3; a++) {
    if (wall->isboundary[a] && (r[a] == 0 || r[a] == box->nlocal[a] + 1)) return 1;
  }
  return 0;
}

int orc_wall_nlink(const orc_model_t * model, const orc_box_t * box,
		   const orc_wall_t * wall) {
  int nlink = 0;
  for (int ic = 1; ic <= box->nlocal[0]; ic++) {
    for (int jc = 1; jc <= box->nlocal[1]; jc++) {
      for (int kc = 1; kc <= box->nlocal[2]; kc++) {
	for (int p = 1; p < model->nvel; p++) {
	  const int r[3] = {ic + model->cv[p][0], jc + model->cv[p][1], kc + model->cv[p][2]};
	  nlink += in_wall(box, wall, r);
	}
      }
    }
  }
  return nlink;
}

/* wall_bbl_slip_kernel, src/wall.c:1118-1210, with the link geometry of
 * wall_init_boundaries_slip (src/wall.c:489-600): for a link (i, p) the wall normal
 * collects the axes along which the step c_p alone already lands in a wall; the
 * remaining (tangential) part of c_p leads to the fluid site k whose population q --
 * c_p with its tangential components reversed -- slips along the wall and arrives at i
 * together with the bounced one:  f(j, -p) = (1 - s) f(i, p) + s f(k, q). */

static int wall_bbl_slip(const orc_model_t * model, const orc_box_t * box,
			 int ndist, orc_wall_t * wall, double * f) {

  const int nvel = model->nvel;

  if (ndist != 1) return -1;

  for (int ic = 1; ic <= box->nlocal[0]; ic++) {
    for (int jc = 1; jc <= box->nlocal[1]; jc++) {
      for (int kc = 1; kc <= box->nlocal[2]; kc++) {
	const int r0[3] = {ic, jc, kc};
	const long i = orc_index(box, ic, jc, kc);
	for (int p = 1; p < nvel; p++) {
	  const int * c = model->cv[p];
	  const int r[3] = {ic + c[0], jc + c[1], kc + c[2]};
	  int wn[3] = {0, 0, 0}, wt[3], cq[3], rk[3];
	  int modwn = 0, modwt = 0, q = -1;
	  double s = 0.0, fi, fk;
	  long j, k;
	  if (!in_wall(box, wall, r)) continue;
	  j = orc_index(box, r[0], r[1], r[2]);
	  for (int a = 0; a < 3; a++) {
	    int ra[3] = {r0[0], r0[1], r0[2]};
	    ra[a] += c[a];
	    if (c[a] != 0 && in_wall(box, wall, ra)) wn[a] = -c[a];
	    modwn += wn[a]*wn[a];
	  }
	  for (int a = 0; a < 3; a++) {
	    wt[a] = wn[a] ? 0 : c[a];
	    modwt += wt[a]*wt[a];
	    cq[a] = -2*wn[a] - c[a];
	    rk[a] = r0[a] + wt[a];
	  }
	  k = i; q = p;
	  if (modwt != 0) {
	    k = orc_index(box, rk[0], rk[1], rk[2]);
	    for (int pp = 0; pp < nvel; pp++) {
	      if (model->cv[pp][0] == cq[0] && model->cv[pp][1] == cq[1] && model->cv[pp][2] == cq[2]) q = pp;
	    }
	    if (modwn == 1) {
	      for (int a = 0; a < 3; a++) {
		if (wn[a] == +1) s = wall->sbot[a];
		if (wn[a] == -1) s = wall->stop[a];
	      }
	    }
	    if (modwn == 2) {
	      for (int a = 0; a < 3; a++) {
		if (wn[a] == +1) s += 0.5*wall->sbot[a];
		if (wn[a] == -1) s += 0.5*wall->stop[a];
	      }
	    }
	  }
	  fi = f[i*nvel + p];
	  fk = f[k*nvel + q];
	  f[j*nvel + (nvel - p)] = (1.0 - s)*fi + s*fk;
	  for (int a = 0; a < 3; a++) {
	    double w = -(c[a] + model->cv[q][a])/2;
	    wall->fnet[a] += 2.0*(1.0 - s)*(fi - model->wv[p])*c[a];
	    wall->fnet[a] += 2.0*w*w*s*(fk - model->wv[q])*model->cv[q][a];
	  }
	}
      }
    }
  }
  return 0;
}

int orc_wall_bbl(const orc_model_t * model, const orc_box_t * box,
		 const orc_param_t * prm, int ndist, orc_wall_t * wall,
		 double * f) {

  const int nvel = model->nvel;
  const double rcs2 = 3.0;
  const int nwall = wall->isboundary[0] + wall->isboundary[1] + wall->isboundary[2];
  int iw = -1;

  if (wall->slip_active) return wall_bbl_slip(model, box, ndist, wall, f);

  /* wall_init_uw, src/wall.c:864-893: wall velocities only for a single pair */
  if (nwall == 1) {
    for (int a = 0; a < 3; a++) if (wall->isboundary[a]) iw = a;
  }

  for (int ic = 1; ic <= box->nlocal[0]; ic++) {
    for (int jc = 1; jc <= box->nlocal[1]; jc++) {
      for (int kc = 1; kc <= box->nlocal[2]; kc++) {
	const long i = orc_index(box, ic, jc, kc);
	for (int p = 1; p < nvel; p++) {
	  const int r[3] = {ic + model->cv[p][0], jc + model->cv[p][1], kc + model->cv[p][2]};
	  const double * uw = NULL;
	  double cdotu = 0.0, fp, force;
	  long j;
	  if (!in_wall(box, wall, r)) continue;
	  j = orc_index(box, r[0], r[1], r[2]);
	  if (iw >= 0 && model->cv[p][iw] == -1) uw = wall->ubot;
	  if (iw >= 0 && model->cv[p][iw] == +1) uw = wall->utop;
	  if (uw) {
	    cdotu = model->cv[p][0]*uw[0] + model->cv[p][1]*uw[1] + model->cv[p][2]*uw[2];
	  }
	  fp = f[(i*ndist + 0)*nvel + p];
	  force = 2.0*fp - 2.0*rcs2*model->wv[p]*prm->rho0*cdotu;
	  for (int a = 0; a < 3; a++) {
	    wall->fnet[a] += (force - 2.0*model->wv[p])*model->cv[p][a];
	  }
	  f[(j*ndist + 0)*nvel + (nvel - p)] = fp - 2.0*rcs2*model->wv[p]*prm->rho0*cdotu;
	  if (ndist > 1) {
	    double gp = f[(i*ndist + 1)*nvel + p];
	    f[(j*ndist + 1)*nvel + (nvel - p)] = gp - 2.0*rcs2*model->wv[p]*prm->rho0*cdotu;
	  }
	}
      }
    }
  }

  return 0;
}

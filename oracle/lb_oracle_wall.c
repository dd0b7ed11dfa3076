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

/* ---------------------------------------------------------------------------
 * Open boundaries for the fluid ("lb_bc_open yes", lb_bc_inflow_type / _outflow_type
 * rhou: src/lb_bc_inflow_rhou.c, src/lb_bc_outflow_rhou.c, selected in
 * src/lb_bc_open_rt.c:48-170): the one direction without walls carries a flow. After
 * lb_halo and ahead of the bounce-back and the propagation (src/ludwig.c:820-831)
 *   update: the layer of halo sites in front of the inlet takes the density of the
 *           fluid sites behind it and the prescribed velocity u0; the layer behind the
 *           outlet takes the prescribed density rho0 and the velocity of the fluid
 *           sites in front of it (:235-281 of either file);
 *   impose: every population that the propagation will carry from such a halo site
 *           into the fluid is set to the equilibrium of the halo site's (rho, u)
 *           (:478-531; links whose halo end lies behind a wall are left to the walls).
 * One rank. rho, u: the hydrodynamic fields the collision has just written.
 * --------------------------------------------------------------------------- */

static orc_open_t * attached_open = NULL;

int orc_open_attach(orc_open_t * open) { attached_open = open; return 0; }
orc_open_t * orc_open_attached(void) { return attached_open; }

int orc_open_apply(const orc_model_t * model, const orc_box_t * box, int ndist,
		   const orc_open_t * open, double * rho, double * u, double * f) {
  const int id = open->id;
  const int id1 = (id + 1) % 3, id2 = (id + 2) % 3;
  const int nvel = model->nvel;
  const double cs2 = 1.0/3.0, rcs2 = 3.0;

  if (id < 0 || id > 2) return -1;

  for (int side = 0; side < 2; side++) {          /* 0: inflow (low end), 1: outflow */
    const int dir = (side == 0) ? -1 : +1;        /* where the halo layer is */
    int lo[3] = {1, 1, 1}, hi[3] = {box->nlocal[0], box->nlocal[1], box->nlocal[2]};
    if (side == 0) hi[id] = 1; else lo[id] = box->nlocal[id];

    /* update */
    for (int ic = lo[0]; ic <= hi[0]; ic++) {
      for (int jc = lo[1]; jc <= hi[1]; jc++) {
	for (int kc = lo[2]; kc <= hi[2]; kc++) {
	  int c[3] = {ic, jc, kc};
	  long index0, index1 = orc_index(box, ic, jc, kc);
	  c[id] += dir;
	  index0 = orc_index(box, c[0], c[1], c[2]);
	  if (side == 0) {
	    rho[index0] = rho[index1];
	    for (int a = 0; a < 3; a++) u[3*index0 + a] = open->u0[a];
	  }
	  else {
	    rho[index0] = open->rho0;
	    for (int a = 0; a < 3; a++) u[3*index0 + a] = u[3*index1 + a];
	  }
	}
      }
    }
    /* impose */
    for (int ic = lo[0]; ic <= hi[0]; ic++) {
      for (int jc = lo[1]; jc <= hi[1]; jc++) {
	for (int kc = lo[2]; kc <= hi[2]; kc++) {
	  for (int p = 1; p < nvel; p++) {
	    int c[3] = {ic + model->cv[p][0], jc + model->cv[p][1], kc + model->cv[p][2]};
	    const int q = nvel - p;                 /* the reverse: into the fluid */
	    long index;
	    double udotc = 0.0, sdotq = 0.0;
	    if (model->cv[p][id] != dir) continue;
	    if (c[id1] < 1 || c[id1] > box->nlocal[id1]) continue;
	    if (c[id2] < 1 || c[id2] > box->nlocal[id2]) continue;
	    index = orc_index(box, c[0], c[1], c[2]);
	    for (int ia = 0; ia < 3; ia++) {
	      udotc += u[3*index + ia]*model->cv[q][ia];
	      for (int ib = 0; ib < 3; ib++) {
		const double d_ab = (ia == ib);
		const double s_ab = model->cv[q][ia]*model->cv[q][ib] - cs2*d_ab;
		sdotq += s_ab*u[3*index + ia]*u[3*index + ib];
	      }
	    }
	    f[(index*ndist + 0)*nvel + q] = rho[index]*model->wv[q]*(1.0 + rcs2*udotc + 0.5*rcs2*rcs2*sdotq);
	  }
	}
      }
    }
  }
  return 0;
}

/* ---------------------------------------------------------------------------
 * grad_3d_27pt_solid_d2 (src/gradient_3d_27pt_solid.c:131-325) next to flat walls, neutral
 * wetting (C = H = 0: no "wetting" keys): differences to the 26 neighbours, those inside a
 * wall contributing the boundary gradient -(C phi_b + H)/kappa = -0; delsq = (1/9) sum,
 * grad_a = (1/18) sum c_a, both over the links in the order of the reference's table.
 * A site is solid when a coordinate along a wall direction lies outside 1 .. n.
 * --------------------------------------------------------------------------- */

int orc_grad_27pt_solid(const orc_box_t * box, const orc_wall_t * wall, const double * phi,
			double * grad, double * delsq) {
  const double r9 = (1.0/9.0), r18 = (1.0/18.0);
  const double c = 0.0, h = 0.0, rkappa = 1.0;     /* neutral: the value of 1/kappa drops out */

  for (int ic = 1; ic <= box->nlocal[0]; ic++) {
    for (int jc = 1; jc <= box->nlocal[1]; jc++) {
      for (int kc = 1; kc <= box->nlocal[2]; kc++) {
	const long index = orc_index(box, ic, jc, kc);
	double gradt[27], gradn[3] = {0.0, 0.0, 0.0}, count[3] = {0.0, 0.0, 0.0};
	int solid[27], cv[27][3];
	double dphi;
	int p = 1;
	for (int a = -1; a <= 1; a++) for (int b = -1; b <= 1; b++) for (int cc = -1; cc <= 1; cc++) {
	  if (a == 0 && b == 0 && cc == 0) continue;
	  cv[p][0] = a; cv[p][1] = b; cv[p][2] = cc;
	  p++;
	}
	for (p = 1; p < 27; p++) {
	  const int n[3] = {ic + cv[p][0], jc + cv[p][1], kc + cv[p][2]};
	  solid[p] = 0;
	  for (int d = 0; d < 3; d++) {
	    if (wall && wall->isboundary[d] && (n[d] < 1 || n[d] > box->nlocal[d])) solid[p] = 1;
	  }
	  if (solid[p]) continue;
	  dphi = phi[orc_index(box, n[0], n[1], n[2])] - phi[index];
	  gradt[p] = dphi;
	  for (int ia = 0; ia < 3; ia++) {
	    gradn[ia] += cv[p][ia]*dphi;
	    count[ia] += cv[p][ia]*cv[p][ia];
	  }
	}
	for (int ia = 0; ia < 3; ia++) if (count[ia] > 0.0) gradn[ia] /= count[ia];
	for (p = 1; p < 27; p++) {
	  if (solid[p]) {
	    const double phi_b = phi[index] + 0.5*(cv[p][0]*gradn[0] + cv[p][1]*gradn[1] + cv[p][2]*gradn[2]);
	    gradt[p] = -(c*phi_b + h)*rkappa;
	  }
	}
	dphi = 0.0;
	for (int ia = 0; ia < 3; ia++) gradn[ia] = 0.0;
	for (p = 1; p < 27; p++) {
	  dphi += gradt[p];
	  for (int ia = 0; ia < 3; ia++) gradn[ia] += gradt[p]*cv[p][ia];
	}
	delsq[index] = r9*dphi;
	for (int ia = 0; ia < 3; ia++) grad[3*index + ia] = r18*gradn[ia];
      }
    }
  }
  return 0;
}

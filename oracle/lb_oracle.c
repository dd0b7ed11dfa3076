/*
 * lb_oracle.c -- TEST INFRASTRUCTURE ONLY. See lb_oracle.h.
 *
 * CPU restatement (not a copy) of the reference algorithm. Each function
 * cites the reference file:line whose behaviour it restates. The arithmetic
 * is plain IEEE double, compiled with -ffp-contract=off.
 */

#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "lb_oracle.h"

/* ------------------------------------------------------------------------
 * Velocity sets. Ordering and weights are interface facts of the reference:
 *   D3Q19 src/lb_d3q19.h:27-40, D3Q27 src/lb_d3q27.h:27-43,
 *   D2Q9 src/lb_d2q9.h:26-35.
 * ---------------------------------------------------------------------- */

static const signed char cv_d3q19[19][3] = {
  { 0, 0, 0},
  { 1, 1, 0}, { 1, 0, 1}, { 1, 0, 0}, { 1, 0,-1}, { 1,-1, 0}, { 0, 1, 1},
  { 0, 1, 0}, { 0, 1,-1}, { 0, 0, 1}, { 0, 0,-1}, { 0,-1, 1}, { 0,-1, 0},
  { 0,-1,-1}, {-1, 1, 0}, {-1, 0, 1}, {-1, 0, 0}, {-1, 0,-1}, {-1,-1, 0}};

static const signed char cv_d2q9[9][3] = {
  { 0, 0, 0},
  { 1, 1, 0}, { 1, 0, 0}, { 1,-1, 0}, { 0, 1, 0},
  { 0,-1, 0}, {-1, 1, 0}, {-1, 0, 0}, {-1,-1, 0}};

/* D3Q27: p = 0 is rest; then lexicographic in (x,y,z) over {-1,0,1}^3
 * skipping the origin. */

static void cv_d3q27_fill(int cv[ORC_NVELMAX][3]) {
  int p = 1;
  cv[0][0] = cv[0][1] = cv[0][2] = 0;
  for (int x = -1; x <= 1; x++) {
    for (int y = -1; y <= 1; y++) {
      for (int z = -1; z <= 1; z++) {
	if (x == 0 && y == 0 && z == 0) continue;
	cv[p][0] = x; cv[p][1] = y; cv[p][2] = z;
	p++;
      }
    }
  }
}

/* Weights depend only on |c|^2. */

static double weight(int nvel, int c2) {
  if (nvel == 19) {
    const double w[3] = {12.0/36.0, 2.0/36.0, 1.0/36.0};
    return w[c2];
  }
  if (nvel == 27) {
    const double w[4] = {64.0/216.0, 16.0/216.0, 4.0/216.0, 1.0/216.0};
    return w[c2];
  }
  {
    const double w[3] = {16.0/36.0, 4.0/36.0, 1.0/36.0};
    return w[c2];
  }
}

/* Rows of the mode matrix: src/lb_d3q19.c:112-150, src/lb_d3q27.c:152-195,
 * src/lb_d2q9.c:137-168. */

static void ma_column(int nvel, const int c[3], double cs2, double * col) {

  double cx = c[0], cy = c[1], cz = c[2];
  double qxx = cx*cx - cs2, qyy = cy*cy - cs2, qzz = cz*cz - cs2;

  if (nvel == 9) {
    col[0] = 1.0; col[1] = cx; col[2] = cy;
    col[3] = qxx; col[4] = cx*cy; col[5] = qyy;
    col[6] = 9.0*qxx*qyy;
    col[7] = 6.0*qyy*cx;
    col[8] = 6.0*qxx*cy;
    return;
  }

  col[0] = 1.0; col[1] = cx; col[2] = cy; col[3] = cz;
  col[4] = qxx; col[5] = cx*cy; col[6] = cx*cz;
  col[7] = qyy; col[8] = cy*cz; col[9] = qzz;

  if (nvel == 19) {
    double c2 = cx*cx + cy*cy + cz*cz;
    double chi1 = (2.0*c2 - 3.0)*(3.0*cz*cz - c2);
    double chi2 = (2.0*c2 - 3.0)*(cy*cy - cx*cx);
    double chi3 = 3.0*c2*c2 - 6.0*c2 + 1;
    col[10] = chi1; col[11] = chi1*cx; col[12] = chi1*cy; col[13] = chi1*cz;
    col[14] = chi2; col[15] = chi2*cx; col[16] = chi2*cy; col[17] = chi2*cz;
    col[18] = chi3;
  }
  else {
    /* nvel == 27: complete Hermite basis */
    col[10] = 3.0*qxx*cy; col[11] = 3.0*qxx*cz;
    col[12] = 3.0*qyy*cz; col[13] = 3.0*qyy*cx;
    col[14] = 3.0*qzz*cx; col[15] = 3.0*qzz*cy;
    col[16] = cx*cy*cz;
    col[17] = 9.0*qxx*qyy; col[18] = 9.0*qyy*qzz; col[19] = 9.0*qzz*qxx;
    col[20] = 9.0*qxx*cy*cz; col[21] = 9.0*qyy*cz*cx; col[22] = 9.0*qzz*cx*cy;
    col[23] = 9.0*qxx*qyy*cz; col[24] = 9.0*qyy*qzz*cx;
    col[25] = 9.0*qzz*qxx*cy;
    col[26] = 27.0*qxx*qyy*qzz;
  }
}

/* orc_model_init: src/lb_model.c:69, src/lb_d3q19.c:30-88 (normalisers
 * na[m] = 1/sum_p w_p ma[m][p]^2) and src/model.c:358-390
 * (mi[p][m] = w_p na[m] ma[m][p]). */

int orc_model_init(int nvel, orc_model_t * model) {

  const double cs2 = 1.0/3.0;

  if (nvel != 9 && nvel != 19 && nvel != 27) return -1;

  memset(model, 0, sizeof(orc_model_t));
  model->nvel = nvel;
  model->ndim = (nvel == 9) ? 2 : 3;
  model->nhydro = 1 + model->ndim + model->ndim*(model->ndim + 1)/2;

  if (nvel == 27) cv_d3q27_fill(model->cv);
  for (int p = 0; p < nvel; p++) {
    int c2 = 0;
    for (int a = 0; a < 3; a++) {
      if (nvel == 19) model->cv[p][a] = cv_d3q19[p][a];
      if (nvel == 9)  model->cv[p][a] = cv_d2q9[p][a];
      c2 += model->cv[p][a]*model->cv[p][a];
    }
    model->wv[p] = weight(nvel, c2);
  }

  for (int p = 0; p < nvel; p++) {
    double col[ORC_NVELMAX];
    ma_column(nvel, model->cv[p], cs2, col);
    for (int m = 0; m < nvel; m++) model->ma[m][p] = col[m];
  }

  for (int m = 0; m < nvel; m++) {
    double wip = 0.0;
    for (int p = 0; p < nvel; p++) {
      wip += model->wv[p]*model->ma[m][p]*model->ma[m][p];
    }
    model->na[m] = 1.0/wip;
  }

  for (int p = 0; p < nvel; p++) {
    for (int m = 0; m < nvel; m++) {
      model->mi[p][m] = model->wv[p]*model->na[m]*model->ma[m][p];
      model->maf[m][p] = model->ma[m][p];
    }
  }
  if (nvel == 19) model->maf[13][4] = 0.0;  /* src/collision.c:2300 */

  return 0;
}

int orc_box_init(const int nlocal[3], orc_box_t * box) {
  box->nsite = 1;
  for (int a = 0; a < 3; a++) {
    box->nlocal[a] = nlocal[a];
    box->nall[a] = nlocal[a] + 2;
    box->nsite *= box->nall[a];
  }
  return 0;
}

/* src/coords.c:617-631 with nhalo = 1: 1-based local coordinates */

long orc_index(const orc_box_t * box, int ic, int jc, int kc) {
  return ((long) ic*box->nall[1] + jc)*box->nall[2] + kc;
}

/* ------------------------------------------------------------------------
 * orc_rtau: inverse relaxation times for every mode.
 * src/collision.c:1181-1264 (lb_collision_relaxation_times_set) and the
 * per-site versions src/collision.c:1287-1538, which give the same numbers
 * for a fixed viscosity. Conserved modes get 0 here (never relaxed).
 * ---------------------------------------------------------------------- */

int orc_rtau(const orc_model_t * model, const orc_param_t * prm,
	     double rtau[ORC_NVELMAX]) {

  const double cs2 = 1.0/3.0;
  const int nd = model->ndim;
  const int ibulk = 1 + nd;          /* LB_TAU_BULK  = 1 + NDIM + XX */
  const int ishear = 1 + nd + 1;     /* LB_TAU_SHEAR = 1 + NDIM + XY */
  double rtau_shear = 1.0/(0.5 + prm->eta_shear/(prm->rho0*cs2));
  double rtau_bulk  = 1.0/(0.5 + prm->eta_bulk /(prm->rho0*cs2));

  for (int m = 0; m < ORC_NVELMAX; m++) rtau[m] = 0.0;

  if (prm->nrelax == ORC_RELAX_M10) {
    for (int m = model->nhydro; m < model->nvel; m++) rtau[m] = 1.0;
  }
  else if (prm->nrelax == ORC_RELAX_BGK) {
    for (int m = 0; m < model->nvel; m++) rtau[m] = rtau_shear;
    rtau_bulk = rtau_shear;
  }
  else if (prm->nrelax == ORC_RELAX_TRT) {
    double tau = prm->eta_shear/(prm->rho0*cs2);
    double rt = 0.5 + 2.0*tau/(tau + 3.0/8.0);
    if (rt > 2.0) rt = 2.0;
    if (model->nvel != 19) return -1;  /* reference: d3q15/d3q19 only */
    rtau[10] = rtau[14] = rtau[18] = rtau_shear;
    rtau[11] = rtau[12] = rtau[13] = rt;
    rtau[15] = rtau[16] = rtau[17] = rt;
  }
  else {
    return -1;
  }

  rtau[ishear] = rtau_shear;
  rtau[ibulk]  = rtau_bulk;

  return 0;
}

/* src/model.c:916-941 lb_1st_moment_equilib_set */

int orc_equilib_set(const orc_model_t * model, int ndist, double * f,
		    long index, double rho, const double u[3]) {

  const double cs2 = 1.0/3.0;
  const double rcs2 = 1.0/cs2;

  for (int p = 0; p < model->nvel; p++) {
    double udotc = 0.0;
    double sdotq = 0.0;
    for (int a = 0; a < 3; a++) {
      udotc += u[a]*model->cv[p][a];
      for (int b = 0; b < 3; b++) {
	double dab = (a == b);
	sdotq += (model->cv[p][a]*model->cv[p][b] - cs2*dab)*u[a]*u[b];
      }
    }
    f[(index*ndist + 0)*model->nvel + p]
      = rho*model->wv[p]*(1.0 + rcs2*udotc + 0.5*rcs2*rcs2*sdotq);
  }
  return 0;
}

/* ------------------------------------------------------------------------
 * Shared part of both collisions: from modes m[] (all nvel of them) and the
 * local force, compute velocity, relax the stress towards
 * rho u u + extra, add the force correction, relax the ghosts, and leave
 * post-collision modes in m[]. src/collision.c:350-544 and :800-925.
 * shat / ghat: random stress and ghost terms (lb_oracle_noise.c) or NULL.
 * ---------------------------------------------------------------------- */

void orc_relax_modes(const orc_model_t * model, const double * rtau,
			const double force[3], const double sextra[3][3],
			const double shat[3][3], const double * ghat,
			double * m, double * rho_out, double u[3]) {

  const int nd = model->ndim;
  const int ibulk = 1 + nd, ishear = 1 + nd + 1;
  const double rdim = 1.0/nd;
  double s[3][3], seq[3][3];
  double rho = m[0];
  double rrho = 1.0/rho;
  double tr_s = 0.0, tr_seq = 0.0;
  int k;

  u[0] = u[1] = u[2] = 0.0;
  for (int a = 0; a < nd; a++) u[a] = rrho*(m[1 + a] + 0.5*force[a]);

  k = 0;
  for (int a = 0; a < nd; a++) {
    for (int b = a; b < nd; b++) {
      s[a][b] = m[1 + nd + k];
      s[b][a] = s[a][b];
      k++;
    }
  }

  for (int a = 0; a < nd; a++) {
    for (int b = 0; b < nd; b++) {
      seq[a][b] = rho*u[a]*u[b] + sextra[a][b];
    }
    tr_s   += s[a][a];
    tr_seq += seq[a][a];
  }

  for (int a = 0; a < nd; a++) {
    s[a][a]   -= rdim*tr_s;
    seq[a][a] -= rdim*tr_seq;
  }

  tr_s = tr_s - rtau[ibulk]*(tr_s - tr_seq);

  for (int a = 0; a < nd; a++) {
    for (int b = 0; b < nd; b++) {
      double dab = (a == b);
      s[a][b] -= rtau[ishear]*(s[a][b] - seq[a][b]);
      s[a][b] += dab*rdim*tr_s;
      s[a][b] += (2.0 - rtau[ishear])*(u[a]*force[b] + force[a]*u[b]);
    }
  }

  for (int a = 0; a < nd; a++) m[1 + a] += force[a];

  k = 0;
  for (int a = 0; a < nd; a++) {
    for (int b = a; b < nd; b++) {
      m[1 + nd + k] = s[a][b];
      if (shat) m[1 + nd + k] += shat[a][b];      /* src/collision.c:529 */
      k++;
    }
  }

  for (k = model->nhydro; k < model->nvel; k++) {
    m[k] = m[k] - rtau[k]*(m[k] - 0.0);
    if (ghat) m[k] += ghat[k];                    /* src/collision.c:541 */
  }

  *rho_out = rho;
}

void orc_f_to_modes(const orc_model_t * model, const double * fs,
		       double * m) {
  for (int k = 0; k < model->nvel; k++) {
    double sum = 0.0;
    for (int p = 0; p < model->nvel; p++) sum += fs[p]*model->maf[k][p];
    m[k] = sum;
  }
}

void orc_modes_to_f(const orc_model_t * model, const double * m,
		       double * fs) {
  for (int p = 0; p < model->nvel; p++) {
    double sum = 0.0;
    for (int k = 0; k < model->nvel; k++) sum += model->mi[p][k]*m[k];
    fs[p] = sum;
  }
}

/* ------------------------------------------------------------------------
 * orc_collide_mrt: single-fluid MRT collision, in place, over x-planes
 * ic = 1..nlocal[X] and the FULL y,z extent including halo sites, which is
 * the reference's kernel extent (src/kernel.c:198-209, src/collision.c:224).
 * All sites are fluid. Writes rho and u (src/collision.c:570-579).
 * ---------------------------------------------------------------------- */

/* Viscosity model (src/collision.c:386-404 with hydro->eta; visc_arrhenius.c): while an
 * eta array is attached the single-fluid collision takes the shear viscosity of each
 * site from it, the bulk viscosity in the Newtonian ratio, and every relaxation time
 * that depends on them (lb_relaxation_time_shear_v/_bulk_v/_ghosts_v). */
static const double * visc_eta_ = NULL;
static const orc_visc_t * visc_ = NULL;

int orc_visc_attach(const orc_visc_t * visc, const double * eta) {
  visc_ = visc;
  visc_eta_ = eta;
  return 0;
}
const orc_visc_t * orc_visc_attached(const double ** eta) {
  if (eta) *eta = visc_eta_;
  return visc_;
}

/* visc_update_kernel, src/visc_arrhenius.c:185-220 (interior sites) */
int orc_visc_arrhenius_update(const orc_box_t * box, const orc_visc_t * visc,
			      const double * phi, double * eta) {
  for (int ic = 1; ic <= box->nlocal[0]; ic++) {
    for (int jc = 1; jc <= box->nlocal[1]; jc++) {
      for (int kc = 1; kc <= box->nlocal[2]; kc++) {
	const long index = orc_index(box, ic, jc, kc);
	double phi0 = phi[index];
	double etaplus, etaminus;
	phi0 = phi0/visc->phistar;
	etaminus = pow(visc->eta_minus, 0.5*(1.0 - phi0));
	etaplus  = pow(visc->eta_plus,  0.5*(1.0 + phi0));
	eta[index] = etaminus*etaplus;
      }
    }
  }
  return 0;
}

int orc_collide_mrt(const orc_model_t * model, const orc_box_t * box,
		    const orc_param_t * prm, int ndist, double * f,
		    const double * force, double * rho, double * u) {

  const int nvel = model->nvel;
  const long first = orc_index(box, 1, 0, 0);
  const long last  = orc_index(box, box->nlocal[0] + 1, 0, 0);
  double rtau[ORC_NVELMAX];
  const double zero33[3][3] = {{0.0}};
  double kt = 0.0;
  unsigned int * nstate = orc_noise_attached(&kt);

  if (orc_rtau(model, prm, rtau) != 0) return -1;
  if (nstate && nvel - model->nhydro > 10) return -1;
  if (nstate && visc_eta_) return -1;     /* noise variances follow one viscosity */

  #pragma omp parallel for schedule(static) firstprivate(rtau)
  for (long index = first; index < last; index++) {
    double m[ORC_NVELMAX];
    double fl[3], ul[3], rl;
    double * fs = f + (index*ndist + 0)*nvel;
    for (int a = 0; a < 3; a++) {
      fl[a] = prm->force_global[a] + force[3*index + a];
    }
    if (visc_eta_) {
      orc_param_t local = *prm;
      local.eta_shear = visc_eta_[index];
      local.eta_bulk = (prm->eta_bulk/prm->eta_shear)*visc_eta_[index];
      orc_rtau(model, &local, rtau);
    }
    orc_f_to_modes(model, fs, m);
    if (nstate) {
      double shat[3][3], ghat[ORC_NVELMAX];
      orc_fluctuations(model, rtau, kt, prm->ghost_off, nstate, index, shat, ghat);
      orc_relax_modes(model, rtau, fl, zero33, (const double (*)[3]) shat, ghat, m, &rl, ul);
    }
    else {
      orc_relax_modes(model, rtau, fl, zero33, NULL, NULL, m, &rl, ul);
    }
    orc_modes_to_f(model, m, fs);
    rho[index] = rl;
    for (int a = 0; a < 3; a++) u[3*index + a] = ul[a];
  }

  return 0;
}

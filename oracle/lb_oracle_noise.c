/*
 * lb_oracle_noise.c -- TEST INFRASTRUCTURE ONLY. See lb_oracle.h.
 *
 * Isothermal fluctuations ("noise") of the collision: the per-site random
 * number generator of src/noise.c and the random stress / ghost-mode terms of
 * src/collision.c:476-515 (single fluid) and :1663-1745 (binary fluid).
 *
 * The uniform generator is Marsaglia's KISS99 (four 32-bit words per site,
 * src/noise.c:456-472); the variates are Ladd's discrete eight-level
 * approximation to a unit Gaussian (src/noise.c:70-77, 398-425): one 32-bit
 * uniform yields up to ten variates, three bits each.
 */

#include <math.h>
#include <stddef.h>

#include "lb_oracle.h"

/* state attached to the collisions (NULL = fluctuations off) */
static unsigned int * noise_state = NULL;
static double noise_kt = 0.0;

int orc_noise_attach(unsigned int * state, double kt) {
  noise_state = state;
  noise_kt = kt;
  return 0;
}

unsigned int * orc_noise_attached(double * kt) {
  if (kt) *kt = noise_kt;
  return noise_state;
}

/* src/noise.c:456-472: one step of KISS99 = a linear congruential word, a
 * three-shift xorshift word and two 16-bit multiply-with-carry words. */

unsigned int orc_noise_uniform(unsigned int s[4]) {

  unsigned int x, mwc;

  s[0] = 69069u*s[0] + 1234567u;

  x = s[1];
  x ^= x << 17;
  x ^= x >> 13;
  x ^= x << 5;
  s[1] = x;

  s[2] = 36969u*(s[2] & 65535u) + (s[2] >> 16);
  s[3] = 18000u*(s[3] & 65535u) + (s[3] >> 16);
  mwc = (s[2] << 16) + s[3];

  return s[1] + (s[0] ^ mwc);
}

/* src/noise.c:136-233: decomposition-independent seeding. Every site of the
 * halo-inclusive local box starts from a seed built from its GLOBAL periodic
 * position (1-based), then takes four outputs of the generator as its state.
 * state[index*4 + ia] (host AOS layout, src/memory.h:124-130). */

int orc_noise_init(const orc_box_t * box, int master_seed,
		   const int noffset[3], const int ntotal[3],
		   unsigned int * state) {

  unsigned int seed0[4] = {13u, 12953u, 712357u, 22383979u};

  if (master_seed > 0) seed0[0] = (unsigned int) master_seed;

  for (int ic = 0; ic <= box->nlocal[0] + 1; ic++) {
    for (int jc = 0; jc <= box->nlocal[1] + 1; jc++) {
      for (int kc = 0; kc <= box->nlocal[2] + 1; kc++) {
	int g[3] = {noffset[0] + ic, noffset[1] + jc, noffset[2] + kc};
	unsigned int local[4];
	long index = orc_index(box, ic, jc, kc);
	for (int a = 0; a < 3; a++) {
	  if (g[a] < 1) g[a] += ntotal[a];
	  if (g[a] > ntotal[a]) g[a] -= ntotal[a];
	  local[a] = seed0[a] + (unsigned int) g[a];
	}
	local[3] = seed0[3];
	for (int ia = 0; ia < 4; ia++) {
	  state[4*index + ia] = orc_noise_uniform(local);
	}
      }
    }
  }

  return 0;
}

/* src/noise.c:398-425: advance the site's generator once; drop the two low
 * bits, then each following group of three bits picks one of the eight levels
 * -sqrt(2+sqrt2), -sqrt(2-sqrt2), 0, 0, 0, 0, +sqrt(2-sqrt2), +sqrt(2+sqrt2). */

int orc_noise_reap_n(unsigned int * state, long index, int nmax, double * reap) {

  const double big = sqrt(2.0 + sqrt(2.0));
  const double small = sqrt(2.0 - sqrt(2.0));
  const double level[8] = {-big, -small, 0.0, 0.0, 0.0, 0.0, small, big};
  unsigned int bits;

  if (nmax > 10) return -1;        /* NNOISE_MAX, src/noise.h:18 */

  bits = orc_noise_uniform(state + 4*index) >> 2;
  for (int n = 0; n < nmax; n++) {
    reap[n] = level[bits & 7u];
    bits >>= 3;
  }

  return 0;
}

/* Random stress and ghost-mode terms for one site.
 * src/collision.c:1826-1883 (lb_fluctuations_stress) with the variances of
 * :1756-1780, and :1900-1920 (lb_fluctuations_ghosts) with :1797-1815; the
 * binary collision's lb_collision_fluctuations (:1663-1745) evaluates the same
 * expressions from lb->param->var_shear / var_bulk (:1601-1604).
 * Six variates are always drawn for the stress (also in two dimensions), then
 * nvel - nhydro for the ghosts from a second draw unless ghost modes are off. */

int orc_fluctuations(const orc_model_t * model, const double * rtau, double kt,
		     int ghost_off, unsigned int * state, long index,
		     double shat[3][3], double * ghat) {

  const int nd = model->ndim;
  const int ishear = 1 + nd + 1, ibulk = 1 + nd;
  const double rcs2 = 3.0;
  const double ktn = kt*rcs2;             /* "without normalisation kT = cs^2" */
  const double tau_s = 1.0/rtau[ishear];
  const double tau_b = 1.0/rtau[ibulk];
  const double var_eta  = sqrt(ktn)*sqrt(1.0/9.0)*sqrt((tau_s + tau_s - 1.0)/(tau_s*tau_s));
  const double var_bulk = sqrt(ktn)*sqrt(2.0/9.0)*sqrt((tau_b + tau_b - 1.0)/(tau_b*tau_b));
  double r[10];
  double tr;

  if (model->nvel - model->nhydro > 10) return -1;

  orc_noise_reap_n(state, index, 6, r);

  shat[0][0] = r[0]; shat[0][1] = r[1]; shat[0][2] = r[2];
  shat[1][0] = r[1]; shat[1][1] = r[3]; shat[1][2] = r[4];
  shat[2][0] = r[2]; shat[2][1] = r[4]; shat[2][2] = r[5];

  tr = (1.0/nd)*(shat[0][0] + shat[1][1] + (nd - 2.0)*shat[2][2]);

  for (int a = 0; a < 3; a++) {
    shat[a][a] -= tr;
    for (int b = 0; b < 3; b++) {
      shat[a][b] *= (a == b) ? var_eta*sqrt(2.0) : var_eta;
    }
  }

  tr *= var_bulk;
  for (int a = 0; a < 3; a++) shat[a][a] += tr;

  for (int k = 0; k < model->nvel; k++) ghat[k] = 0.0;

  /* ghost modes off: no second draw (src/collision.c:503, :1730) */
  if (ghost_off) return 0;

  orc_noise_reap_n(state, index, model->nvel - model->nhydro, r);

  for (int k = model->nhydro; k < model->nvel; k++) {
    double tau_g = 1.0/rtau[k];
    double rna = 1.0/model->na[k];
    double var = sqrt(ktn*rna)*sqrt((tau_g + tau_g - 1.0)/(tau_g*tau_g));
    ghat[k] = var*r[k - model->nhydro];
  }

  return 0;
}

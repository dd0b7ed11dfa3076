/*
 * lb_b200.c -- see lb_b200.h. Plain C; every operation forwards to the C ABI of
 * libb200lb.so. No arithmetic on lattice data happens on the host.
 */

#include <float.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "lb_b200.h"

static char host_err[512] = "";

static int host_fail(const char * what) {
  snprintf(host_err, sizeof(host_err), "%s: %s", what, b200lb_last_error());
  return -1;
}

const char * lb_b200_error(void) { return host_err; }

#define TRY(call) do { if ((call) != 0) return host_fail(#call); } while (0)

lb_data_options_t lb_data_options_default(void) {
  /* src/lb_data_options.c: D3Q19, one distribution, M10, LB_HALO_TARGET */
  lb_data_options_t opts = {3, 19, 1, LB_RELAXATION_M10, LB_HALO_TARGET};
  return opts;
}

int lb_data_create(cs_t * cs, physics_t * phys, const lb_data_options_t * opts, lb_t ** plb) {

  b200lb_desc_t d;
  lb_t * lb = NULL;

  if (cs == NULL || phys == NULL || opts == NULL || plb == NULL) {
    snprintf(host_err, sizeof(host_err), "lb_data_create: null argument");
    return -1;
  }
  lb = (lb_t *) calloc(1, sizeof(lb_t));
  if (lb == NULL) return -1;

  memset(&d, 0, sizeof(d));
  d.ndim = opts->ndim; d.nvel = opts->nvel; d.ndist = opts->ndist;
  d.nhalo = cs->nhalo;
  for (int a = 0; a < 3; a++) {
    d.nlocal[a] = cs->nlocal[a];
    d.cartsz[a] = cs->mpi_cartsz[a];
    d.cartcoords[a] = cs->mpi_cartcoords[a];
    d.periodic[a] = cs->periodic[a];
    d.nbr_lo[a] = cs->mpi_cart_neighbours[0][a];
    d.nbr_hi[a] = cs->mpi_cart_neighbours[1][a];
  }
  d.rank = cs->mpi_cartrank; d.nranks = cs->nranks;
  d.device = -1;

  if (b200lb_create(&d, &lb->target) != 0) { free(lb); return host_fail("lb_data_create"); }

  lb->ndim = opts->ndim; lb->nvel = opts->nvel; lb->ndist = opts->ndist;
  lb->nsite = (int) b200lb_nsite(lb->target);
  lb->cs = cs; lb->phys = phys;
  lb->nrelax = opts->nrelax; lb->haloscheme = opts->halo;
  lb->f = (double *) calloc((size_t) lb->nsite*lb->ndist*lb->nvel, sizeof(double));
  if (lb->f == NULL) { b200lb_destroy(lb->target); free(lb); return -1; }
  *plb = lb;
  return 0;
}

int lb_free(lb_t * lb) {
  if (lb == NULL) return 0;
  b200lb_destroy(lb->target);
  free(lb->f);
  free(lb);
  return 0;
}

int lb_memcpy(lb_t * lb, tdpMemcpyKind flag) {
  if (flag == tdpMemcpyHostToDevice) TRY(b200lb_f_upload(lb->target, lb->f, B200LB_HOST_AOS));
  else TRY(b200lb_f_download(lb->target, lb->f, B200LB_HOST_AOS));
  return 0;
}

int lb_collision_relaxation_set(lb_t * lb, lb_relaxation_enum_t nrelax) {
  lb->nrelax = nrelax;
  return 0;
}

/* lb_collide: gathers the parameters the reference commits every step
 * (src/collision.c:144-163, :1181-1264, :1928-1980) and runs the collision. */

/* what the reference commits from physics_ref() and the free energy */
static int gather_params(lb_t * lb, fe_t * fe, b200lb_param_t * p) {
  memset(p, 0, sizeof(*p));
  p->rho0 = lb->phys->rho0;
  p->eta_shear = lb->phys->eta_shear;
  p->eta_bulk = lb->phys->eta_bulk;
  for (int a = 0; a < 3; a++) p->force_global[a] = lb->phys->fbody[a];
  p->nrelax = (int) lb->nrelax;
  p->mobility = lb->phys->mobility;
  if (lb->ndist == 2 && fe == NULL) {
    snprintf(host_err, sizeof(host_err), "lb_collide: ndist = 2 needs the symmetric free energy");
    return -1;
  }
  if (fe != NULL) {
    fe_symm_t * fs = (fe_symm_t *) fe;     /* src/collision.c:160 */
    p->a = fs->param.a; p->b = fs->param.b; p->kappa = fs->param.kappa;
  }
  /* lb_collision_noise_var_set (src/collision.c:1572): variances follow kt */
  p->kt = lb->phys->kt;
  p->ghost_off = (lb->isghost == LB_GHOST_OFF);
  return 0;
}

int lb_collide(lb_t * lb, hydro_t * hydro, map_t * map, noise_t * noise, fe_t * fe, visc_t * visc) {

  b200lb_param_t p;

  (void) map;
  if (hydro == NULL) return 0;               /* src/collision.c:149 */

  if (gather_params(lb, fe, &p)) return -1;
  TRY(b200lb_visc_model_set(lb->target, visc != NULL));       /* p.have_visc_model, src/collision.c:1954 */
  TRY(b200lb_noise_present_set(lb->target, (noise != NULL) && noise->on[NOISE_RHO]));
  TRY(b200lb_set_param(lb->target, &p));
  TRY(b200lb_collide(lb->target));
  return 0;
}

/* physics time, as far as the hot path reads it */

int physics_control_next_step(physics_t * phys) { phys->t_current += 1; return 0; }
int physics_control_init_time(physics_t * phys, int nstart, int nstep) {
  (void) nstep;
  phys->t_start = nstart;
  phys->t_current = nstart;
  return 0;
}
int physics_control_timestep(physics_t * phys) { return phys->t_current; }
int physics_control_time(physics_t * phys, double * t) {
  *t = 1.0*(phys->t_start + phys->t_current - 1.0);
  return 0;
}

/* lees_edw_t: plane positions, reprojection and displacement live in the library */

int lees_edw_create(lb_t * lb, const lees_edw_options_t * opts, lees_edw_t ** ple) {
  lees_edw_t * le = (lees_edw_t *) calloc(1, sizeof(lees_edw_t));
  if (le == NULL) return -1;
  le->lb = lb;
  if (opts) le->opts = *opts;
  if (b200lb_le_commit(lb->target, le->opts.nplanes, le->opts.uy,
		       le->opts.type == LE_SHEAR_TYPE_OSCILLATORY, le->opts.period, le->opts.nt0)) {
    free(le);
    return host_fail("b200lb_le_commit");
  }
  le->nplanelocal = le->opts.nplanes/lb->cs->mpi_cartsz[0];
  *ple = le;
  return 0;
}

int lees_edw_free(lees_edw_t * le) { free(le); return 0; }
int lees_edw_nplane_total(lees_edw_t * le) { return le->opts.nplanes; }
int lees_edw_nplane_local(lees_edw_t * le) { return le->nplanelocal; }
int lees_edw_plane_location(lees_edw_t * le, int plane) {
  return b200lb_le_plane_location(le->lb->target, plane);
}

/* The reference copies the lattice to the host, works there and copies it back
 * (src/model_le.c:70-83); here the planes are handled where the lattice is. */
int lb_le_apply_boundary_conditions(lb_t * lb, lees_edw_t * le) {
  if (lees_edw_nplane_local(le) == 0) return 0;
  /* the distributions are displaced for the step number (src/model_le.c:133,245) */
  TRY(b200lb_le_apply(lb->target, 1.0*physics_control_timestep(lb->phys)));
  return 0;
}

/* wall_t: links, map and the momentum tally live in the library */

int wall_create(lb_t * lb, wall_t ** pwall) {
  wall_t * wall = (wall_t *) calloc(1, sizeof(wall_t));
  if (wall == NULL) return -1;
  wall->lb = lb;
  *pwall = wall;
  return 0;
}

int wall_free(wall_t * wall) { free(wall); return 0; }

/* the face entries of the reference's table; the library derives edges and corners */
wall_slip_t wall_slip(double sbot[3], double stop[3]) {
  wall_slip_t ws;
  memset(&ws, 0, sizeof(ws));
  for (int a = 0; a < 3; a++) {
    ws.s[WALL_SLIP_XBOT + 2*a] = sbot[a];
    ws.s[WALL_SLIP_XTOP + 2*a] = stop[a];
    if (sbot[a] != 0.0 || stop[a] != 0.0) ws.active = 1;
  }
  return ws;
}

int wall_commit(wall_t * wall, wall_param_t * param) {
  wall->param = *param;
  wall->param.iswall = (param->isboundary[0] || param->isboundary[1] || param->isboundary[2]);
  TRY(b200lb_wall_commit(wall->lb->target, param->isboundary, param->ubot, param->utop));
  if (param->slip.active) {
    double sbot[3], stop[3];
    for (int a = 0; a < 3; a++) {
      sbot[a] = param->slip.s[WALL_SLIP_XBOT + 2*a];
      stop[a] = param->slip.s[WALL_SLIP_XTOP + 2*a];
    }
    TRY(b200lb_wall_slip_set(wall->lb->target, sbot, stop, 1));
  }
  return 0;
}

int wall_present(wall_t * wall) { return wall != NULL && wall->param.iswall; }

/* The reference seeds the wall layer for colloid-wall exchange here; wall_bbl then
 * overwrites every one of those populations (same site, same velocity). Nothing to do. */
int wall_set_wall_distributions(wall_t * wall) { (void) wall; return 0; }

int wall_bbl(wall_t * wall) {
  if (!wall_present(wall)) return 0;
  TRY(b200lb_wall_bbl(wall->lb->target));
  return 0;
}

int wall_momentum(wall_t * wall, double f[3]) {
  TRY(b200lb_wall_momentum(wall->lb->target, f));
  return 0;
}

int wall_momentum_add(wall_t * wall, const double f[3]) {
  TRY(b200lb_wall_momentum_add(wall->lb->target, f));
  return 0;
}

int lb_collision_ghost_modes_on(lb_t * lb)  { lb->isghost = LB_GHOST_ON;  return 0; }
int lb_collision_ghost_modes_off(lb_t * lb) { lb->isghost = LB_GHOST_OFF; return 0; }

/* noise_t: the generator state lives on the device; the host array is the interchange
 * copy noise_memcpy moves, as for lb->f. */

int noise_create(lb_t * lb, noise_t ** pnoise) {
  noise_t * obj = (noise_t *) calloc(1, sizeof(noise_t));
  if (obj == NULL) return -1;
  obj->lb = lb;
  obj->nsites = lb->nsite;
  *pnoise = obj;
  return 0;
}

int noise_free(noise_t * obj) { if (obj) { free(obj->state); free(obj); } return 0; }

int noise_init(noise_t * obj, int master_seed) {
  if (obj->state == NULL) {
    obj->state = (unsigned int *) calloc((size_t) NNOISE_STATE*obj->nsites, sizeof(unsigned int));
    if (obj->state == NULL) return -1;
  }
  TRY(b200lb_noise_init(obj->lb->target, master_seed));
  return noise_memcpy(obj, tdpMemcpyDeviceToHost);
}

int noise_present_set(noise_t * obj, noise_enum_t type, int present) {
  obj->on[type] = present;
  return 0;
}

int noise_present(noise_t * obj, noise_enum_t type, int * present) {
  *present = obj->on[type];
  return 0;
}

int noise_memcpy(noise_t * obj, tdpMemcpyKind flag) {
  if (flag == tdpMemcpyHostToDevice) {
    TRY(b200lb_noise_state_upload(obj->lb->target, obj->state, B200LB_HOST_AOS));
  }
  else {
    TRY(b200lb_noise_state_download(obj->lb->target, obj->state, B200LB_HOST_AOS));
  }
  return 0;
}

int lb_halo(lb_t * lb) { return lb_halo_swap(lb, lb->haloscheme); }

int lb_halo_swap(lb_t * lb, lb_halo_enum_t flag) {
  int scheme = (flag == LB_HALO_OPENMP_REDUCED) ? B200LB_HALO_REDUCED : B200LB_HALO_FULL;
  TRY(b200lb_halo_swap(lb->target, scheme));
  return 0;
}

int lb_propagation(lb_t * lb) {
  TRY(b200lb_propagate(lb->target));
  return 0;
}

/* hydro */

static int field_init(field_t * f, lb_t * lb, int nf, int id) {
  f->nf = nf; f->nsites = lb->nsite; f->lb = lb; f->id = id;
  f->data = (double *) calloc((size_t) nf*lb->nsite, sizeof(double));
  return (f->data == NULL) ? -1 : 0;
}

int hydro_create(lb_t * lb, hydro_t ** ph) {
  hydro_t * h = (hydro_t *) calloc(1, sizeof(hydro_t));
  if (h == NULL) return -1;
  h->nsite = lb->nsite; h->lb = lb;
  if (field_init(&h->rho, lb, 1, B200LB_FIELD_RHO) || field_init(&h->u, lb, 3, B200LB_FIELD_U)
      || field_init(&h->force, lb, 3, B200LB_FIELD_FORCE)) return -1;
  *ph = h;
  return 0;
}

int hydro_free(hydro_t * h) {
  if (h == NULL) return 0;
  free(h->rho.data); free(h->u.data); free(h->force.data); free(h);
  return 0;
}

static int field_xfer(field_t * f, tdpMemcpyKind flag) {
  if (flag == tdpMemcpyHostToDevice) TRY(b200lb_field_upload(f->lb->target, f->id, f->data, B200LB_HOST_AOS));
  else TRY(b200lb_field_download(f->lb->target, f->id, f->data, B200LB_HOST_AOS));
  return 0;
}

int hydro_memcpy(hydro_t * h, tdpMemcpyKind flag) {
  /* src/hydro.c:154-178 copies rho, u, force (eta is not on this path) */
  if (flag == tdpMemcpyHostToDevice) {
    if (field_xfer(&h->rho, flag) || field_xfer(&h->u, flag) || field_xfer(&h->force, flag)) return -1;
    return 0;
  }
  if (field_xfer(&h->rho, flag) || field_xfer(&h->u, flag) || field_xfer(&h->force, flag)) return -1;
  return 0;
}

int hydro_f_zero(hydro_t * h, const double fzero[3]) {
  TRY(b200lb_hydro_f_zero(h->lb->target, fzero));
  return 0;
}

int hydro_u_zero(hydro_t * h, const double uzero[3]) {
  TRY(b200lb_hydro_u_zero(h->lb->target, uzero));
  return 0;
}

int hydro_rho0(hydro_t * h, double rho0) {
  TRY(b200lb_hydro_rho0(h->lb->target, rho0));
  return 0;
}

int hydro_correct_momentum(hydro_t * h) {
  TRY(b200lb_hydro_correct_momentum(h->lb->target));
  return 0;
}

/* order parameter, gradient, free energy */

/* The order parameter as a finite-difference field ("free_energy symmetric"):
 * phi_ch_t + pth_t of the reference are one object in the library (b200lb_ch_create). */

static int advection_order_ = 1;             /* src/advection.c:85: a file-scope default */

int advection_order_set(const int order) { advection_order_ = order; return 0; }
int advection_order(int * order) { *order = advection_order_; return 0; }

static int fd_commit(lb_t * lb) {
  const int method = lb->fd_method ? lb->fd_method : FE_FORCE_METHOD_STRESS_DIVERGENCE;
  if (lb->fd_committed && lb->fd_order == advection_order_ && lb->fd_method_applied == method) return 0;
  TRY(b200lb_ch_create(lb->target, advection_order_, method));
  lb->fd_committed = 1; lb->fd_order = advection_order_; lb->fd_method_applied = method;
  return 0;
}

int phi_ch_create(lb_t * lb, phi_ch_info_t * info, phi_ch_t ** ppch) {
  phi_ch_t * pch = NULL;
  if (info != NULL && info->conserve != 0) {
    snprintf(host_err, sizeof(host_err), "phi_ch_create: the compensated update (conserve = %d) is not on this path", info->conserve);
    return -1;
  }
  if (fd_commit(lb)) return -1;
  pch = (phi_ch_t *) calloc(1, sizeof(phi_ch_t));
  if (pch == NULL) return -1;
  pch->lb = lb;
  *ppch = pch;
  return 0;
}
int phi_ch_free(phi_ch_t * pch) { free(pch); return 0; }

int pth_create(lb_t * lb, int method, pth_t ** ppth) {
  pth_t * pth = (pth_t *) calloc(1, sizeof(pth_t));
  if (pth == NULL) return -1;
  pth->method = method; pth->lb = lb;
  lb->fd_method = method;
  if (lb->fd_committed) { if (fd_commit(lb)) { free(pth); return -1; } }
  *ppth = pth;
  return 0;
}
int pth_free(pth_t * pth) { free(pth); return 0; }

static int fd_params(lb_t * lb, fe_t * fe) {
  b200lb_param_t p;
  if (fd_commit(lb)) return -1;
  if (gather_params(lb, fe, &p)) return -1;
  TRY(b200lb_set_param(lb->target, &p));
  TRY(b200lb_ch_grad_mu_set(lb->target, lb->phys->grad_mu));
  return 0;
}

/* src/phi_force.c:74; without hydrodynamics or with "no_force" there is nothing to do */
int phi_force_calculation(pth_t * pth, fe_t * fe, field_t * phi, hydro_t * hydro) {
  (void) phi;
  if (hydro == NULL) return 0;
  if (pth->method == FE_FORCE_METHOD_NO_FORCE) return 0;
  if (fd_params(pth->lb, fe)) return -1;
  TRY(b200lb_phi_force(pth->lb->target));
  return 0;
}

/* src/phi_cahn_hilliard.c:206 (no order-parameter noise, all-fluid map) */
int phi_cahn_hilliard(phi_ch_t * pch, fe_t * fe, field_t * phi, hydro_t * hydro, map_t * map, noise_t * noise) {
  (void) phi; (void) map;
  if (noise != NULL && noise->on[NOISE_PHI]) {
    snprintf(host_err, sizeof(host_err), "phi_cahn_hilliard: order-parameter noise is not on this path");
    return -1;
  }
  if (hydro == NULL) {
    snprintf(host_err, sizeof(host_err), "phi_cahn_hilliard: the purely diffusive case (no hydrodynamics) is not on this path");
    return -1;
  }
  if (fd_params(pch->lb, fe)) return -1;
  TRY(b200lb_cahn_hilliard(pch->lb->target));
  return 0;
}

/* viscosity_model arrhenius (src/visc_arrhenius.c): needs the order parameter */
int visc_arrhenius_create(lb_t * lb, field_t * phi, visc_arrhenius_param_t param, visc_t ** pvisc) {
  visc_t * visc = NULL;
  if (phi == NULL) {
    snprintf(host_err, sizeof(host_err), "visc_arrhenius_create: requires a composition");
    return -1;
  }
  TRY(b200lb_visc_arrhenius_commit(lb->target, param.eta_plus, param.eta_minus, param.phistar));
  visc = (visc_t *) calloc(1, sizeof(visc_t));
  if (visc == NULL) return -1;
  visc->param = param; visc->lb = lb;
  *pvisc = visc;
  return 0;
}
int visc_arrhenius_free(visc_t * visc) { free(visc); return 0; }
int visc_arrhenius_update(visc_t * visc, hydro_t * hydro) {
  (void) hydro;
  TRY(b200lb_visc_update(visc->lb->target));
  return 0;
}

int hydro_u_halo(hydro_t * hydro) {
  TRY(b200lb_hydro_u_halo(hydro->lb->target));
  return 0;
}

int field_create(lb_t * lb, const char * name, field_t ** pphi) {
  field_t * phi = (field_t *) calloc(1, sizeof(field_t));
  (void) name;
  if (phi == NULL || field_init(phi, lb, 1, B200LB_FIELD_PHI)) return -1;
  *pphi = phi;
  return 0;
}

int field_free(field_t * phi) { if (phi) { free(phi->data); free(phi); } return 0; }
int field_memcpy(field_t * phi, tdpMemcpyKind flag) { return field_xfer(phi, flag); }

int field_halo(field_t * phi) {
  TRY(b200lb_phi_halo(phi->lb->target));
  return 0;
}

int phi_lb_to_field(field_t * phi, lb_t * lb) {
  (void) phi;
  TRY(b200lb_phi_from_g(lb->target));
  return 0;
}

int field_grad_create(field_t * phi, field_grad_t ** pg) {
  field_grad_t * g = (field_grad_t *) calloc(1, sizeof(field_grad_t));
  if (g == NULL) return -1;
  g->field = phi; g->lb = phi->lb;
  g->d2 = grad_3d_7pt_fluid_d2;
  g->grad = (double *) calloc((size_t) 3*phi->nsites, sizeof(double));
  g->delsq = (double *) calloc((size_t) phi->nsites, sizeof(double));
  if (g->grad == NULL || g->delsq == NULL) return -1;
  *pg = g;
  return 0;
}

int field_grad_free(field_grad_t * g) { if (g) { free(g->grad); free(g->delsq); free(g); } return 0; }

/* The reference's function-pointer plug point (src/field_grad.h:22,47): gradient_rt.c
 * registers one of these per "fd_gradient_calculation" key. */
int grad_3d_7pt_fluid_d2(field_grad_t * g) {
  TRY(b200lb_grad7(g->lb->target));
  return 0;
}

int grad_3d_27pt_fluid_d2(field_grad_t * g) {
  TRY(b200lb_grad27(g->lb->target));
  return 0;
}

int field_grad_set(field_grad_t * g, grad_ft d2, grad_ft d4) {
  (void) d4;                      /* no fourth-order consumer on this path */
  if (g == NULL || d2 == NULL) return -1;
  g->d2 = d2;
  return 0;
}

int field_grad_compute(field_grad_t * g) {
  /* field_leesedwards(obj->field) (src/field_grad.c:310) displaces the order parameter
   * across the planes for physics_control_time (src/field.c:553): in the library that is
   * part of the gradient call */
  double t = 0.0;
  physics_control_time(g->lb->phys, &t);
  TRY(b200lb_le_field_time_set(g->lb->target, t));
  /* obj->d2(obj) (src/field_grad.c:312) */
  return g->d2(g);
}

int field_grad_memcpy(field_grad_t * g, tdpMemcpyKind flag) {
  if (flag == tdpMemcpyHostToDevice) {
    TRY(b200lb_field_upload(g->lb->target, B200LB_FIELD_GRAD, g->grad, B200LB_HOST_AOS));
    TRY(b200lb_field_upload(g->lb->target, B200LB_FIELD_DELSQ, g->delsq, B200LB_HOST_AOS));
  }
  else {
    TRY(b200lb_field_download(g->lb->target, B200LB_FIELD_GRAD, g->grad, B200LB_HOST_AOS));
    TRY(b200lb_field_download(g->lb->target, B200LB_FIELD_DELSQ, g->delsq, B200LB_HOST_AOS));
  }
  return 0;
}

int fe_symm_create(field_t * phi, field_grad_t * dphi, fe_symm_t ** pfe) {
  fe_symm_t * fe = (fe_symm_t *) calloc(1, sizeof(fe_symm_t));
  if (fe == NULL) return -1;
  fe->phi = phi; fe->dphi = dphi;
  *pfe = fe;
  return 0;
}

int fe_symm_free(fe_symm_t * fe) { free(fe); return 0; }
int fe_symm_param_set(fe_symm_t * fe, fe_symm_param_t values) { fe->param = values; return 0; }

int stats_distribution_sums(lb_t * lb, double sums[8]) {
  TRY(b200lb_stats(lb->target, sums));
  return 0;
}

/* stats_velocity_minmax (src/stats_velocity.c:53-130) without hydro_memcpy: extrema
 * start from FLT_MAX / FLT_MIN as there (:75-76); usum is the volume flux. */
int stats_velocity_minmax(stats_vel_t * stat, hydro_t * hydro, map_t * map) {
  double q[12];
  (void) map;
  TRY(b200lb_field_stats(hydro->lb->target, B200LB_FIELD_U, q));
  for (int ia = 0; ia < 3; ia++) {
    stat->usum[ia] = q[4*ia + 0];
    stat->umin[ia] = (q[4*ia + 2] < FLT_MAX) ? q[4*ia + 2] : FLT_MAX;
    stat->umax[ia] = (q[4*ia + 3] > FLT_MIN) ? q[4*ia + 3] : FLT_MIN;
  }
  return 0;
}

/* stats_field_info for a scalar (src/phi_stats.c:89-132): sum, mean, variance, min, max */
int stats_field_info_scalar(field_t * phi, map_t * map, double out[5]) {
  double q[4];
  const cs_t * cs = phi->lb->cs;
  const double vol = (double) cs->nlocal[0]*cs->nlocal[1]*cs->nlocal[2];
  (void) map;
  TRY(b200lb_field_stats(phi->lb->target, phi->id, q));
  out[0] = q[0];
  out[1] = q[0]/vol;
  out[2] = q[1]/vol - out[1]*out[1];
  out[3] = q[2];
  out[4] = q[3];
  return 0;
}

/* lb_io_aggr_pack / _unpack, binary records (src/model.c:1498-1560) */
int lb_io_aggr_pack(const lb_t * lb, io_aggregator_t * aggr) {
  if (aggr->szelement != (size_t) b200lb_io_record_doubles(lb->target)*sizeof(double)) {
    snprintf(host_err, sizeof(host_err), "lb_io_aggr_pack: record size %zu", aggr->szelement);
    return -1;
  }
  TRY(b200lb_io_pack(lb->target, (double *) aggr->buf));
  return 0;
}

int lb_io_aggr_unpack(lb_t * lb, const io_aggregator_t * aggr) {
  TRY(b200lb_io_unpack(lb->target, (const double *) aggr->buf));
  return 0;
}

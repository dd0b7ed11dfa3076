/*****************************************************************************
 *
 *  ref_driver.c -- TEST INFRASTRUCTURE ONLY.
 *
 *  A thin ctypes-facing driver linked against the UNMODIFIED reference
 *  (Ludwig v0.20.1 objects compiled in place from /root/reference by
 *  oracle/Makefile). It builds the reference's own pe/cs/lb/hydro/map/
 *  noise(/phi/phi_grad/fe_symm) objects on a periodic single-rank box and
 *  calls the reference's own hot-path entry points one stage at a time so
 *  that tests can compare raw fp64 arrays:
 *
 *    lb_collide           src/collision.c:144
 *    lb_halo_swap         src/model.c:570
 *    lb_propagation       src/propagation.c:43
 *    phi_lb_to_field      src/phi_lb_coupler.c:39
 *    field_halo           src/field.c:434
 *    grad_3d_7pt_fluid_d2 src/gradient_3d_7pt_fluid.c:75
 *    grad_3d_27pt_fluid_d2 src/gradient_3d_27pt_fluid.c:75 (refdrv_grad_select(27))
 *    noise_init / noise_reap_n / noise_present_set  src/noise.c:136,398,557
 *    wall_create / wall_commit / wall_bbl / wall_momentum  src/wall.c:59,168,960,1299
 *    lees_edw_create, lb_le_apply_boundary_conditions  src/leesedwards.c:80, src/model_le.c:57
 *    hydro_f_zero/u_zero  src/hydro.c:306,279
 *    phi_force_calculation  src/phi_force.c:74 (pth_stress_compute src/phi_force_stress.c:171,
 *                           pth_force_fluid_driver src/phi_force_colloid.c:280, phi_grad_mu_fluid
 *                           src/phi_grad_mu.c:52)
 *    phi_cahn_hilliard      src/phi_cahn_hilliard.c:206 (hydro_u_halo, advection_x, phi_ch_flux_mu1 ...)
 *
 *  The step sequence of refdrv_step() follows src/ludwig.c:530-862 for the
 *  "free_energy none", "symmetric_lb" and (refdrv_fd_options) the finite-difference
 *  "symmetric" cases (periodic, no colloids).
 *
 *  Nothing under ludwig_b200/ links or loads this file.
 *
 *****************************************************************************/

#include <assert.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "pe.h"
#include "coords.h"
#include "physics.h"
#include "lb_data.h"
#include "collision.h"
#include "propagation.h"
#include "hydro.h"
#include "map.h"
#include "noise.h"
#include "field.h"
#include "field_grad.h"
#include "symmetric.h"
#include "leesedwards.h"
#include "phi_lb_coupler.h"
#include "gradient_3d_7pt_fluid.h"
#include "gradient_3d_27pt_fluid.h"
#include "gradient_3d_27pt_solid.h"
#include "lb_bc_inflow_rhou.h"
#include "lb_bc_outflow_rhou.h"
#include "timer.h"
#include "wall.h"
#include "model_le.h"
#include "lees_edwards_options.h"
#include "phi_cahn_hilliard.h"
#include "phi_force.h"
#include "phi_force_stress.h"
#include "advection.h"
#include "fe_force_method.h"
#include "visc.h"
#include "visc_arrhenius.h"

typedef struct refdrv_s {
  pe_t * pe;
  cs_t * cs;
  lees_edw_t * le;
  physics_t * phys;
  lb_t * lb;
  hydro_t * hydro;
  map_t * map;
  noise_t * noise;
  field_t * phi;
  field_grad_t * phi_grad;
  fe_symm_t * fe;
  wall_t * wall;
  pth_t * pth;
  phi_ch_t * pch;
  visc_t * visc;
  int nhalo;
  int nplanes;
  int ndist;
  int mpi_up;
} refdrv_t;

static refdrv_t rd = {0};

int refdrv_nvel(void) { return NVEL; }
int refdrv_ndim(void) { return NDIM; }

/* Create everything for a periodic box ntotal[3], halo width 1. */

static int next_periodic[3] = {1, 1, 1};
static lees_edw_options_t next_le = {0};

/* Lees-Edwards planes of the NEXT refdrv_init ("N_LE_plane", "LE_plane_vel",
 * "LE_oscillation_period", "LE_time_offset"; src/leesedwards_rt.c) */
int refdrv_le_options(int nplanes, double uy, int oscillatory, int period, int nt0) {
  lees_edw_options_t opts = {0};
  opts.nplanes = nplanes;
  opts.uy = uy;
  opts.type = oscillatory ? LE_SHEAR_TYPE_OSCILLATORY : LE_SHEAR_TYPE_STEADY;
  opts.period = period;
  opts.nt0 = nt0;
  next_le = opts;
  return 0;
}

/* periodicity of the NEXT refdrv_init ("periodicity" input key, src/coords_rt.c) */
int refdrv_periodicity(const int periodic[3]) {
  for (int ia = 0; ia < 3; ia++) next_periodic[ia] = periodic[ia];
  return 0;
}

/* "free_energy symmetric" for the NEXT refdrv_init (src/ludwig.c:1159-1250): nhalo = 2,
 * the order parameter is a finite-difference field updated by phi_cahn_hilliard, the
 * fluid feels it through phi_force_calculation. order = "fd_advection_scheme_order",
 * method = "fe_force_method" (fe_force_method_enum_t). */
static int next_fd = 0, next_fd_order = 1, next_fd_method = FE_FORCE_METHOD_STRESS_DIVERGENCE;
int refdrv_fd_options(int order, int method) {
  next_fd = 1;
  next_fd_order = order;
  next_fd_method = method;
  return 0;
}
int refdrv_nhalo(void) { return rd.nhalo; }

int refdrv_init(const int ntotal[3], int ndist, int halo_scheme) {

  int nhalo = next_fd ? 2 : 1;
  rd.nhalo = nhalo;

  if (rd.mpi_up == 0) {
    MPI_Init(NULL, NULL);
    rd.mpi_up = 1;
  }

  pe_create(MPI_COMM_WORLD, PE_QUIET, &rd.pe);
  cs_create(rd.pe, &rd.cs);
  cs_nhalo_set(rd.cs, nhalo);
  cs_ntotal_set(rd.cs, ntotal);
  cs_periodicity_set(rd.cs, next_periodic);
  next_periodic[0] = next_periodic[1] = next_periodic[2] = 1;
  cs_init(rd.cs);
  physics_create(rd.pe, &rd.phys);       /* lees_edw_init takes physics_ref() */
  if (next_le.nplanes > 0) {
    lees_edw_create(rd.pe, rd.cs, &next_le, &rd.le);
    rd.nplanes = next_le.nplanes;
  }
  else {
    lees_edw_create(rd.pe, rd.cs, NULL, &rd.le);
    rd.nplanes = 0;
  }
  memset(&next_le, 0, sizeof(next_le));


  {
    lb_data_options_t opts = lb_data_options_default();
    opts.ndim  = NDIM;
    opts.nvel  = NVEL;
    opts.ndist = ndist;
    opts.halo  = (lb_halo_enum_t) halo_scheme;
    lb_data_create(rd.pe, rd.cs, &opts, &rd.lb);
  }
  rd.ndist = ndist;

  {
    hydro_options_t hopts = hydro_options_default();
    hydro_create(rd.pe, rd.cs, rd.le, &hopts, &rd.hydro);
  }

  map_create(rd.pe, rd.cs, 0, &rd.map);
  noise_create(rd.pe, rd.cs, &rd.noise);
  noise_init(rd.noise, 0);

  if (ndist == 2 || next_fd) {
    field_options_t fopts = field_options_ndata_nhalo(1, nhalo);
    field_create(rd.pe, rd.cs, rd.le, "phi", &fopts, &rd.phi);
    field_grad_create(rd.pe, rd.phi, 2, &rd.phi_grad);
    field_grad_set(rd.phi_grad, grad_3d_7pt_fluid_d2, NULL);
    fe_symm_create(rd.pe, rd.cs, rd.phi, rd.phi_grad, &rd.fe);
  }
  if (next_fd) {
    phi_ch_info_t ch_options = {0};
    assert(ndist == 1);
    phi_ch_create(rd.pe, rd.cs, rd.le, &ch_options, &rd.pch);
    pth_create(rd.pe, rd.cs, next_fd_method, &rd.pth);
    advection_order_set(next_fd_order);
    /* phi_force_calculation asks the wall object whether there are walls
     * (wall_create leaves wall->param zeroed: none) */
    wall_create(rd.pe, rd.cs, rd.map, rd.lb, &rd.wall);
    next_fd = 0;
  }

  return 0;
}

int refdrv_finish(void) {
  refdrv_open_free();

  if (rd.visc) rd.visc->func->free(rd.visc);
  if (rd.pch) phi_ch_free(rd.pch);
  if (rd.pth) pth_free(rd.pth);
  if (rd.wall) wall_free(rd.wall);
  if (rd.fe) fe_symm_free(rd.fe);
  if (rd.phi_grad) field_grad_free(rd.phi_grad);
  if (rd.phi) field_free(rd.phi);
  if (rd.noise) noise_free(rd.noise);
  if (rd.map) map_free(rd.map);
  if (rd.hydro) hydro_free(rd.hydro);
  if (rd.lb) lb_free(rd.lb);
  if (rd.phys) physics_free(rd.phys);
  if (rd.le) lees_edw_free(rd.le);
  if (rd.cs) cs_free(rd.cs);
  if (rd.pe) pe_free(rd.pe);
  {
    int up = rd.mpi_up;
    memset(&rd, 0, sizeof(rd));
    rd.mpi_up = up;
  }
  return 0;
}

/* Parameters */

int refdrv_physics(double rho0, double eta_shear, double eta_bulk,
		   const double fbody[3], double mobility) {
  double f[3] = {fbody[0], fbody[1], fbody[2]};
  physics_rho0_set(rd.phys, rho0);
  physics_eta_shear_set(rd.phys, eta_shear);
  physics_eta_bulk_set(rd.phys, eta_bulk);
  physics_fbody_set(rd.phys, f);
  physics_mobility_set(rd.phys, mobility);
  return 0;
}

int refdrv_relaxation(int nrelax) {
  return lb_collision_relaxation_set(rd.lb, (lb_relaxation_enum_t) nrelax);
}

int refdrv_fe_symm(double a, double b, double kappa) {
  fe_symm_param_t p = {0};
  p.a = a; p.b = b; p.kappa = kappa;
  assert(rd.fe);
  return fe_symm_param_set(rd.fe, p);
}

/* Raw host arrays of the reference objects. The host layout of a serial
 * (non -DADDR_SOA) build is array-of-structures: src/memory.h:124-130,
 *   f[(index*ndist + n)*nvel + p], u[index*3 + ia], grad[(index*nf+n)*3+ia].
 * refdrv_addr_f() exposes LB_ADDR itself so tests can confirm that. */

int refdrv_nsite(void) { return rd.lb->nsite; }
int refdrv_nall(int nall[3]) { return cs_nall(rd.cs, nall); }
int refdrv_index(int ic, int jc, int kc) { return cs_index(rd.cs, ic, jc, kc); }
int refdrv_addr_f(int index, int n, int p) {
  return LB_ADDR(rd.lb->nsite, rd.lb->ndist, rd.lb->nvel, index, n, p);
}

double * refdrv_f(void)     { return rd.lb->f; }
double * refdrv_rho(void)   { return rd.hydro->rho->data; }
double * refdrv_u(void)     { return rd.hydro->u->data; }
double * refdrv_force(void) { return rd.hydro->force->data; }
double * refdrv_phi(void)   { return rd.phi ? rd.phi->data : NULL; }
int refdrv_phi_nsites(void) { return rd.phi ? rd.phi->nsites : 0; }   /* with the Lees-Edwards buffer planes */
double * refdrv_grad(void)  { return rd.phi_grad ? rd.phi_grad->grad : NULL; }
double * refdrv_delsq(void) { return rd.phi_grad ? rd.phi_grad->delsq : NULL; }
char   * refdrv_status(void) { return rd.map->status; }

/* Model constants as the reference holds them after lb_data_create
 * (src/model.c:358-390) */

int refdrv_model(double * wv, double * ma, double * mi, int * cv) {
  int p, m, ia;
  for (p = 0; p < NVEL; p++) {
    wv[p] = rd.lb->param->wv[p];
    for (ia = 0; ia < 3; ia++) cv[3*p + ia] = rd.lb->param->cv[p][ia];
    for (m = 0; m < NVEL; m++) {
      ma[p*NVEL + m] = rd.lb->param->ma[p][m];
      mi[p*NVEL + m] = rd.lb->param->mi[p][m];
    }
  }
  return 0;
}

int refdrv_rtau(double * rtau) {
  int p;
  lb_collision_relaxation_times_set(rd.lb);
  for (p = 0; p < NVEL; p++) rtau[p] = rd.lb->param->rtau[p];
  return 0;
}

int refdrv_equilib_set(int index, double rho, const double u[3]) {
  double uu[3] = {u[0], u[1], u[2]};
  return lb_1st_moment_equilib_set(rd.lb, index, rho, uu);
}

/* Individual stages: each is the reference's public entry point */

int refdrv_collide(void) {
  return lb_collide(rd.lb, rd.hydro, rd.map, rd.noise, (fe_t *) rd.fe, rd.visc);
}

/* "viscosity_model arrhenius" (src/ludwig.c:2001-2040): needs the order parameter */
int refdrv_visc_arrhenius(double eta_plus, double eta_minus, double phistar) {
  visc_arrhenius_param_t param = {0};
  visc_arrhenius_t * visc = NULL;
  if (rd.phi == NULL) return -1;
  param.eta_plus = eta_plus; param.eta_minus = eta_minus; param.phistar = phistar;
  visc_arrhenius_create(rd.pe, rd.cs, rd.phi, param, &visc);
  rd.visc = (visc_t *) visc;
  return 0;
}
int refdrv_visc_update(void) { return rd.visc->func->update(rd.visc, rd.hydro); }
double * refdrv_eta(void) { return rd.hydro->eta->data; }

int refdrv_halo(int scheme) {
  if (scheme < 0) return lb_halo(rd.lb);
  return lb_halo_swap(rd.lb, (lb_halo_enum_t) scheme);
}

int refdrv_propagate(void)   { return lb_propagation(rd.lb); }
int refdrv_phi_from_lb(void) { return phi_lb_to_field(rd.phi, rd.lb); }
int refdrv_phi_halo(void)    { return field_halo(rd.phi); }
int refdrv_grad_compute(void) { return field_grad_compute(rd.phi_grad); }

/* The reference's function-pointer plug point (src/field_grad.c field_grad_set):
 * which d2 stencil field_grad_compute() runs, as gradient_rt.c would select. */
/* Open boundaries for the fluid: what lb_bc_open_rt (src/lb_bc_open_rt.c:48-170) builds for
 * "lb_bc_inflow_type rhou" / "lb_bc_outflow_type rhou"; refdrv_open_apply = the four calls
 * after lb_halo in src/ludwig.c:820-831 */
static lb_bc_inflow_rhou_t * rd_inflow = NULL;
static lb_bc_outflow_rhou_t * rd_outflow = NULL;

int refdrv_open_bc(int flow_dir, const double u0[3], double rho0) {
  lb_bc_inflow_opts_t oi = lb_bc_inflow_opts_default();
  lb_bc_outflow_opts_t oo = lb_bc_outflow_opts_default();
  oi.nvel = NVEL; oo.nvel = NVEL;
  for (int a = 0; a < 3; a++) { oi.flow[a] = (a == flow_dir); oo.flow[a] = (a == flow_dir); oi.u0[a] = u0[a]; }
  oo.rho0 = rho0;
  if (rd_inflow) { lb_bc_inflow_rhou_free(rd_inflow); rd_inflow = NULL; }
  if (rd_outflow) { lb_bc_outflow_rhou_free(rd_outflow); rd_outflow = NULL; }
  if (lb_bc_inflow_rhou_create(rd.pe, rd.cs, &oi, &rd_inflow)) return -1;
  if (lb_bc_outflow_rhou_create(rd.pe, rd.cs, &oo, &rd_outflow)) return -1;
  return 0;
}

int refdrv_open_apply(void) {
  if (rd_inflow == NULL || rd_outflow == NULL) return -1;
  lb_bc_inflow_rhou_update(rd_inflow, rd.hydro);
  lb_bc_inflow_rhou_impose(rd_inflow, rd.hydro, rd.lb);
  lb_bc_outflow_rhou_update(rd_outflow, rd.hydro);
  lb_bc_outflow_rhou_impose(rd_outflow, rd.hydro, rd.lb);
  return 0;
}

int refdrv_open_free(void) {
  if (rd_inflow) { lb_bc_inflow_rhou_free(rd_inflow); rd_inflow = NULL; }
  if (rd_outflow) { lb_bc_outflow_rhou_free(rd_outflow); rd_outflow = NULL; }
  return 0;
}

int refdrv_grad_select(int npt) {
  if (rd.phi_grad == NULL) return -1;
  if (npt == 7)  return field_grad_set(rd.phi_grad, grad_3d_7pt_fluid_d2, NULL);
  if (npt == 27) return field_grad_set(rd.phi_grad, grad_3d_27pt_fluid_d2, NULL);
  if (npt == 28) {
    /* fd_gradient_calculation 3d_27pt_solid (src/gradient_rt.c): needs the map the walls
     * have marked (refdrv_wall first) and the free energy for 1/kappa */
    grad_3d_27pt_solid_map_set(rd.map);
    if (rd.fe) grad_3d_27pt_solid_fe_set(rd.fe);
    return field_grad_set(rd.phi_grad, grad_3d_27pt_solid_d2, NULL);
  }
  return -1;
}

/* Isothermal fluctuations: what "isothermal_fluctuations on" + "temperature kt" do
 * (src/collision_rt.c:55, src/physics_rt.c) and the per-site generator state
 * (src/noise.c:136-233), host layout state[index*4 + ia] in an AOS build.
 * reseed > 0 re-runs noise_init with that master seed. */
int refdrv_noise(int on, double kt, int reseed) {
  if (reseed > 0) {
    free(rd.noise->state);
    rd.noise->state = NULL;
    noise_init(rd.noise, reseed);
  }
  physics_kt_set(rd.phys, kt);
  return noise_present_set(rd.noise, NOISE_RHO, on);
}
/* lb_le_apply_boundary_conditions (src/model_le.c:57) at time step t
 * (physics_control_timestep, which the time loop advances before each step) */
int refdrv_le_apply(int t) {
  physics_control_init_time(rd.phys, t, 1);
  return lb_le_apply_boundary_conditions(rd.lb, rd.le);
}
int refdrv_le_plane_location(int np) { return lees_edw_plane_location(rd.le, np); }
int refdrv_time_set(int t) { return physics_control_init_time(rd.phys, t, 1); }
double refdrv_time(void) { double t; physics_control_time(rd.phys, &t); return t; }

/* Flat walls: what wall_rt_init does for "boundary_walls" / "boundary_speed_*"
 * (src/wall_rt.c:40-120): wall_create + wall_commit (map, links, wall velocities). */
int refdrv_wall(const int isboundary[3], const double ubot[3], const double utop[3]) {
  wall_param_t p = {0};
  for (int ia = 0; ia < 3; ia++) {
    p.isboundary[ia] = isboundary[ia];
    p.ubot[ia] = ubot[ia];
    p.utop[ia] = utop[ia];
  }
  p.iswall = (p.isboundary[0] || p.isboundary[1] || p.isboundary[2]);
  wall_create(rd.pe, rd.cs, rd.map, rd.lb, &rd.wall);
  return wall_commit(rd.wall, &p);
}
/* the same with partial slip: wall_slip() table + "boundary_walls_slip_active"
 * (src/wall_rt.c:86-104) */
int refdrv_wall_slip(const int isboundary[3], const double sbot[3], const double stop[3],
		     int force_active) {
  wall_param_t p = {0};
  double sb[3] = {sbot[0], sbot[1], sbot[2]};
  double st[3] = {stop[0], stop[1], stop[2]};
  for (int ia = 0; ia < 3; ia++) p.isboundary[ia] = isboundary[ia];
  p.iswall = (p.isboundary[0] || p.isboundary[1] || p.isboundary[2]);
  p.slip = wall_slip(sb, st);
  if (wall_slip_valid(&p.slip) == 0) return -1;
  if (force_active) p.slip.active = 1;
  wall_create(rd.pe, rd.cs, rd.map, rd.lb, &rd.wall);
  return wall_commit(rd.wall, &p);
}
int refdrv_wall_nlink(void) { return rd.wall ? rd.wall->nlink : 0; }
int refdrv_wall_bbl(void) {
  wall_set_wall_distributions(rd.wall);
  return wall_bbl(rd.wall);
}
int refdrv_wall_momentum(double f[3]) { return wall_momentum(rd.wall, f); }

/* "lb_ghost_modes off" (src/collision_rt.c:87-102) */
int refdrv_ghost_modes(int on) {
  return on ? lb_collision_ghost_modes_on(rd.lb) : lb_collision_ghost_modes_off(rd.lb);
}
unsigned int * refdrv_noise_state(void) { return rd.noise->state; }
unsigned int refdrv_noise_uniform(unsigned int state[4]) { return noise_uniform(state); }
int refdrv_noise_reap_n(int index, int nmax, double * reap) {
  return noise_reap_n(rd.noise, index, nmax, reap);
}

/* The finite-difference order parameter: stage by stage (src/ludwig.c:727-768) */
int refdrv_phi_force(void) {
  return phi_force_calculation(rd.pe, rd.cs, rd.le, rd.wall, rd.pth, (fe_t *) rd.fe,
			       rd.map, rd.phi, rd.hydro);
}
int refdrv_cahn_hilliard(void) {
  return phi_cahn_hilliard(rd.pch, (fe_t *) rd.fe, rd.phi, rd.hydro, rd.map, NULL);
}
int refdrv_u_halo(void) { return hydro_u_halo(rd.hydro); }
int refdrv_grad_mu(const double gm[3]) {       /* "grad_mu" key, src/physics_rt.c */
  double g[3] = {gm[0], gm[1], gm[2]};
  return physics_grad_mu_set(rd.phys, g);
}

int refdrv_hydro_f_zero(const double f[3]) { return hydro_f_zero(rd.hydro, f); }
int refdrv_hydro_u_zero(const double u[3]) { return hydro_u_zero(rd.hydro, u); }

/* The time loop, in the order of src/ludwig.c:530-862 for a periodic,
 * colloid-free, wall-free system; returns wall-clock seconds. */

double refdrv_step(int nsteps) {

  int n;
  double t0, t1;
  const double zero[3] = {0.0, 0.0, 0.0};

  t0 = MPI_Wtime();

  for (n = 0; n < nsteps; n++) {
    if (rd.nplanes > 0) physics_control_next_step(rd.phys);     /* src/ludwig.c:530 */
    hydro_f_zero(rd.hydro, zero);
    if (rd.ndist == 2) {
      phi_lb_to_field(rd.phi, rd.lb);
      field_halo(rd.phi);
      field_grad_compute(rd.phi_grad);
    }
    if (rd.pch) {                       /* src/ludwig.c:558-575, 727-768 */
      field_halo(rd.phi);
      field_grad_compute(rd.phi_grad);
      phi_force_calculation(rd.pe, rd.cs, rd.le, rd.wall, rd.pth, (fe_t *) rd.fe,
			    rd.map, rd.phi, rd.hydro);
      phi_cahn_hilliard(rd.pch, (fe_t *) rd.fe, rd.phi, rd.hydro, rd.map, NULL);
    }
    hydro_u_zero(rd.hydro, zero);
    if (rd.visc) rd.visc->func->update(rd.visc, rd.hydro);      /* src/ludwig.c:794 */
    lb_collide(rd.lb, rd.hydro, rd.map, rd.noise, (fe_t *) rd.fe, rd.visc);
    if (rd.nplanes > 0) lb_le_apply_boundary_conditions(rd.lb, rd.le);   /* src/ludwig.c:810 */
    lb_halo(rd.lb);
    if (rd_inflow && rd_outflow) refdrv_open_apply();            /* src/ludwig.c:820-831 */
    if (rd.wall && wall_present(rd.wall)) {                      /* src/ludwig.c:837-842 */
      wall_set_wall_distributions(rd.wall);
      wall_bbl(rd.wall);
    }
    lb_propagation(rd.lb);
  }

  t1 = MPI_Wtime();

  return (t1 - t0);
}

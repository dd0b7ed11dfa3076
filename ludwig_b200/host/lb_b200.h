/*
 * lb_b200.h -- host-side C mirror of the reference call surface for the LB hot path,
 * implemented entirely on top of the C ABI (include/b200lb.h -> libb200lb.so).
 *
 * Host code stays in C, as in the reference (zazu29/ludwig v0.20.1). The names,
 * argument meaning, return convention (0 = success) and the host array layouts are
 * those of the reference, so code written against
 *   lb_collide            src/collision.h:27       lb_propagation  src/propagation.h:21
 *   lb_halo, lb_halo_swap src/lb_data.h:159-160    lb_memcpy       src/lb_data.h:156
 *   hydro_f_zero/u_zero   src/hydro.h:64-65        hydro_memcpy    src/hydro.h:56
 *   phi_lb_to_field       src/phi_lb_coupler.h:22  field_halo      src/field.h:97
 *   field_grad_compute    src/field_grad.h:49      field_memcpy, field_grad_memcpy
 * reads the same here. The objects are slim: they hold the host arrays (the
 * interchange format, LB_ADDR / addr_rank1 AOS, src/memory.h:124-130) and share ONE
 * device handle; everything outside the hot path (pe_t, io, colloids ...) is absent.
 * Where the reference pulls state from the physics_ref() singleton, this mirror takes
 * a physics_t argument at creation. Failures: the reference aborts through pe_fatal;
 * here every function returns non-zero and lb_b200_error() gives the message.
 */

#ifndef LUDWIG_LB_B200_H
#define LUDWIG_LB_B200_H

#include "../../include/b200lb.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef enum tdpMemcpyKind_enum { tdpMemcpyHostToDevice = 1, tdpMemcpyDeviceToHost = 2 } tdpMemcpyKind;

typedef enum lb_relaxation_enum { LB_RELAXATION_M10, LB_RELAXATION_BGK, LB_RELAXATION_TRT } lb_relaxation_enum_t;
typedef enum lb_halo_enum { LB_HALO_TARGET, LB_HALO_OPENMP_FULL, LB_HALO_OPENMP_REDUCED } lb_halo_enum_t;

/* the part of cs_t the path needs (src/coords.h:30-63) */
typedef struct cs_s {
  int ntotal[3];
  int nlocal[3];
  int nhalo;                 /* 1 */
  int mpi_cartsz[3];
  int mpi_cartcoords[3];
  int periodic[3];
  int mpi_cart_neighbours[2][3];   /* [CS_BACKWARD/CS_FORWARD][X,Y,Z] */
  int mpi_cartrank;
  int nranks;
} cs_t;

/* what lb_collide reads from physics_ref() (src/collision.c:1197-1205, :1945-1973) */
typedef struct physics_s {
  double rho0;
  double eta_shear;
  double eta_bulk;
  double fbody[3];
  double mobility;
  double kt;                       /* physics_kt: isothermal fluctuations */
  int t_current;                   /* physics_control_timestep (src/physics.c) */
  double grad_mu[3];               /* physics_grad_mu: external chemical potential gradient */
  int t_start;                     /* the step this run started from (0 unless restarted) */
} physics_t;
int physics_control_next_step(physics_t * phys);                   /* src/physics.c:670 */
int physics_control_init_time(physics_t * phys, int nstart, int nstep);  /* src/physics.c:687 */
int physics_control_timestep(physics_t * phys);                    /* src/physics.c:635: the step number */
int physics_control_time(physics_t * phys, double * t);            /* src/physics.c:651: t_start + t_current - 1 */

typedef struct lb_data_options_s {
  int ndim;
  int nvel;
  int ndist;
  lb_relaxation_enum_t nrelax;
  lb_halo_enum_t halo;
} lb_data_options_t;

typedef enum lb_mode_enum_type {LB_GHOST_ON = 0, LB_GHOST_OFF = 1} lb_mode_enum_t;

typedef struct fe_symm_param_s { double a, b, kappa, c, h; } fe_symm_param_t;

typedef struct lb_data_s {
  int ndim, nvel, ndist, nsite;
  cs_t * cs;
  double * f;                      /* host distributions, LB_ADDR(nsite,ndist,nvel,index,n,p) */
  lb_relaxation_enum_t nrelax;
  int isghost;                     /* LB_GHOST_ON / LB_GHOST_OFF (src/lb_data.h:152) */
  lb_halo_enum_t haloscheme;
  physics_t * phys;                /* stands in for physics_ref() */
  b200lb_t * target;               /* the device side (reference: lb->target) */
  /* finite-difference order parameter: what phi_ch_create / pth_create / advection_order_set
   * have asked for and what the library has been told */
  int fd_method, fd_committed, fd_order, fd_method_applied;
} lb_t;

typedef struct field_s { int nf; int nsites; double * data; lb_t * lb; int id; } field_t;

typedef struct hydro_s {
  int nsite;
  field_t rho, u, force;           /* host copies, addr_rank1(nsite, nf, index, ia) */
  lb_t * lb;
} hydro_t;

typedef struct field_grad_s field_grad_t;
typedef int (* grad_ft)(field_grad_t * fgrad);           /* src/field_grad.h:22 */
struct field_grad_s { field_t * field; double * grad; double * delsq; lb_t * lb; grad_ft d2; };

typedef struct fe_symm_s { fe_symm_param_t param; field_t * phi; field_grad_t * dphi; } fe_symm_t;

typedef void map_t;     /* configs are all-fluid: not consulted */
/* noise_t (src/noise.h:38-50): per-site generator state and the on/off switches */
typedef enum {NOISE_RHO = 0, NOISE_PHI, NOISE_QAB, NOISE_END} noise_enum_t;
#define NNOISE_STATE 4
typedef struct noise_s {
  int nsites;
  int on[NOISE_END];
  unsigned int * state;            /* host copy, state[index*NNOISE_STATE + ia] */
  lb_t * lb;
} noise_t;
/* visc_t (src/visc.h) with the one model of the reference, visc_arrhenius_t */
typedef struct visc_arrhenius_param_s { double eta_minus, eta_plus, phistar; } visc_arrhenius_param_t;   /* src/visc_arrhenius.h */
typedef struct visc_s { visc_arrhenius_param_t param; struct lb_data_s * lb; } visc_t;
typedef void fe_t;

const char * lb_b200_error(void);

lb_data_options_t lb_data_options_default(void);
int lb_data_create(cs_t * cs, physics_t * phys, const lb_data_options_t * opts, lb_t ** lb);
int lb_free(lb_t * lb);
int lb_memcpy(lb_t * lb, tdpMemcpyKind flag);
int lb_collision_relaxation_set(lb_t * lb, lb_relaxation_enum_t nrelax);
int lb_collision_ghost_modes_on(lb_t * lb);                        /* src/collision.c:1120 */
int lb_collision_ghost_modes_off(lb_t * lb);                       /* src/collision.c:1137 */

int lb_collide(lb_t * lb, hydro_t * hydro, map_t * map, noise_t * noise, fe_t * fe, visc_t * visc);
int lb_halo(lb_t * lb);
int lb_halo_swap(lb_t * lb, lb_halo_enum_t flag);
int lb_propagation(lb_t * lb);

/* wall_t (src/wall.h:47-76): flat no-slip walls */
enum { WALL_NO_SLIP = 0, WALL_SLIP_XBOT, WALL_SLIP_XTOP, WALL_SLIP_YBOT, WALL_SLIP_YTOP,
       WALL_SLIP_ZBOT, WALL_SLIP_ZTOP, WALL_SLIP_MAX = 19 };        /* src/wall.h:25-45 */
typedef struct wall_slip_s { int active; double s[WALL_SLIP_MAX]; } wall_slip_t;
typedef struct wall_param_s {
  int iswall;
  int isboundary[3];
  double ubot[3];
  double utop[3];
  wall_slip_t slip;
} wall_param_t;
wall_slip_t wall_slip(double sbot[3], double stop[3]);             /* src/wall.c:284 */
typedef struct wall_s { wall_param_t param; lb_t * lb; } wall_t;
int wall_create(lb_t * lb, wall_t ** wall);                        /* src/wall.c:59 */
int wall_free(wall_t * wall);
int wall_commit(wall_t * wall, wall_param_t * param);              /* src/wall.c:168 */
int wall_present(wall_t * wall);
int wall_set_wall_distributions(wall_t * wall);                    /* src/wall.c:901 */
int wall_bbl(wall_t * wall);                                       /* src/wall.c:960 */
int wall_momentum(wall_t * wall, double f[3]);                     /* src/wall.c:1299 */
int wall_momentum_add(wall_t * wall, const double f[3]);           /* src/wall.c:1279 */

/* lees_edw_t (src/leesedwards.c): the planes live in the library; the host side keeps the options */
typedef enum lees_edw_enum {LE_SHEAR_TYPE_INVALID, LE_SHEAR_TYPE_STEADY, LE_SHEAR_TYPE_OSCILLATORY} lees_edw_enum_t;
typedef struct lees_edw_options_s {                                /* src/lees_edwards_options.h:25-34 */
  int nplanes;
  lees_edw_enum_t type;
  int period;
  int nt0;
  double uy;
} lees_edw_options_t;
typedef struct lees_edw_s { lees_edw_options_t opts; int nplanelocal; lb_t * lb; } lees_edw_t;
int lees_edw_create(lb_t * lb, const lees_edw_options_t * opts, lees_edw_t ** le);   /* src/leesedwards.c:80 */
int lees_edw_free(lees_edw_t * le);
int lees_edw_nplane_total(lees_edw_t * le);
int lees_edw_nplane_local(lees_edw_t * le);
int lees_edw_plane_location(lees_edw_t * le, int plane);           /* src/leesedwards.c:599 */
int lb_le_apply_boundary_conditions(lb_t * lb, lees_edw_t * le);   /* src/model_le.c:57 */

int noise_create(lb_t * lb, noise_t ** noise);                      /* src/noise.c:56 */
int noise_free(noise_t * noise);
int noise_init(noise_t * noise, int master_seed);                  /* src/noise.c:136 */
int noise_present_set(noise_t * noise, noise_enum_t type, int present);   /* src/noise.c:557 */
int noise_present(noise_t * noise, noise_enum_t type, int * present);
int noise_memcpy(noise_t * noise, tdpMemcpyKind flag);             /* src/noise.c:253 */

int hydro_create(lb_t * lb, hydro_t ** hydro);
int hydro_free(hydro_t * hydro);
int hydro_memcpy(hydro_t * hydro, tdpMemcpyKind flag);
int hydro_f_zero(hydro_t * hydro, const double fzero[3]);
int hydro_u_zero(hydro_t * hydro, const double uzero[3]);
int hydro_rho0(hydro_t * hydro, double rho0);                      /* src/hydro.h:66 */
int hydro_correct_momentum(hydro_t * hydro);                       /* src/hydro.h:63 */

int field_create(lb_t * lb, const char * name, field_t ** phi);      /* the order parameter */
int field_free(field_t * phi);
int field_memcpy(field_t * phi, tdpMemcpyKind flag);
int field_halo(field_t * phi);
int phi_lb_to_field(field_t * phi, lb_t * lb);

int field_grad_create(field_t * phi, field_grad_t ** grad);           /* d2 = grad_3d_7pt_fluid_d2 */
int field_grad_free(field_grad_t * grad);
int field_grad_compute(field_grad_t * grad);
int field_grad_set(field_grad_t * grad, grad_ft d2, grad_ft d4);   /* src/field_grad.h:47 */
int grad_3d_7pt_fluid_d2(field_grad_t * grad);            /* src/gradient_3d_7pt_fluid.c:75 (default) */
int grad_3d_27pt_fluid_d2(field_grad_t * grad);           /* src/gradient_3d_27pt_fluid.c:75 */
int field_grad_memcpy(field_grad_t * grad, tdpMemcpyKind flag);

/* The order parameter as a finite-difference field ("free_energy symmetric",
 * src/ludwig.c:1159-1250, time step src/ludwig.c:727-768). Create the phi_ch_t before the
 * field_t of the order parameter (the library allocates phi with it). */
typedef enum fe_force_method_enum {                                /* src/fe_force_method.h:20-28 */
  FE_FORCE_METHOD_INVALID, FE_FORCE_METHOD_NO_FORCE, FE_FORCE_METHOD_STRESS_DIVERGENCE,
  FE_FORCE_METHOD_PHI_GRADMU, FE_FORCE_METHOD_PHI_GRADMU_CORRECTION,
  FE_FORCE_METHOD_RELAXATION_SYMM, FE_FORCE_METHOD_RELAXATION_ANTI, FE_FORCE_METHOD_MAX
} fe_force_method_enum_t;
typedef struct phi_ch_info_s { int conserve; } phi_ch_info_t;     /* src/phi_cahn_hilliard.h:29 */
typedef struct phi_ch_s { phi_ch_info_t info; lb_t * lb; } phi_ch_t;
typedef struct pth_s { int method; lb_t * lb; } pth_t;            /* src/phi_force_stress.h */
int advection_order_set(const int order);                          /* src/advection.c:93 */
int advection_order(int * order);
int phi_ch_create(lb_t * lb, phi_ch_info_t * info, phi_ch_t ** pch);            /* src/phi_cahn_hilliard.c:121 */
int phi_ch_free(phi_ch_t * pch);
int pth_create(lb_t * lb, int method, pth_t ** pth);                            /* src/phi_force_stress.c:41 */
int pth_free(pth_t * pth);
int phi_force_calculation(pth_t * pth, fe_t * fe, field_t * phi, hydro_t * hydro);   /* src/phi_force.c:74 */
int phi_cahn_hilliard(phi_ch_t * pch, fe_t * fe, field_t * phi, hydro_t * hydro,
		      map_t * map, noise_t * noise);                             /* src/phi_cahn_hilliard.c:206 */
int hydro_u_halo(hydro_t * hydro);                                              /* src/hydro.c */

int visc_arrhenius_create(lb_t * lb, field_t * phi, visc_arrhenius_param_t param, visc_t ** visc);   /* src/visc_arrhenius.c:52 */
int visc_arrhenius_free(visc_t * visc);
int visc_arrhenius_update(visc_t * visc, hydro_t * hydro);          /* src/visc_arrhenius.c:149; visc->func->update */

int fe_symm_create(field_t * phi, field_grad_t * dphi, fe_symm_t ** fe);
int fe_symm_free(fe_symm_t * fe);
int fe_symm_param_set(fe_symm_t * fe, fe_symm_param_t values);

/* stats_vel_t as filled by stats_velocity_minmax (src/stats_velocity.h) */
typedef struct stats_vel_s { double umin[3], umax[3], usum[3]; } stats_vel_t;
int stats_velocity_minmax(stats_vel_t * stat, hydro_t * hydro, map_t * map);
/* the numbers stats_field_info prints for a scalar field: total mean variance min max */
int stats_field_info_scalar(field_t * phi, map_t * map, double out[5]);

/* io_aggregator_t (src/io_aggregator.h:23-29), the members the LB records use */
typedef struct io_aggregator_s { size_t szelement; size_t szbuf; char * buf; } io_aggregator_t;
int lb_io_aggr_pack(const lb_t * lb, io_aggregator_t * aggr);      /* src/model.c:1498 */
int lb_io_aggr_unpack(lb_t * lb, const io_aggregator_t * aggr);    /* src/model.c:1534 */

/* stats_distribution_print / stats_distribution_momentum sums (src/stats_distribution.c) */
int stats_distribution_sums(lb_t * lb, double sums[8]);

#ifdef __cplusplus
}
#endif

#endif

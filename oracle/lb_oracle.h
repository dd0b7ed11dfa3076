/*
 * lb_oracle.h -- TEST INFRASTRUCTURE ONLY. Not part of the product.
 *
 * Plain-C CPU restatement of the Ludwig (zazu29/ludwig v0.20.1) lattice
 * Boltzmann time-step hot path, used as the parity checker for the CUDA
 * path in ludwig_b200/. Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load the library built from it.
 *
 * Parity status: PINNED. tests/test_oracle_vs_ref.py compares every function
 * here with the unmodified reference compiled in place (oracle/_ref, built by
 * oracle/Makefile) and tests/golden/ holds vectors dumped from that build
 * (generator: tests/golden/make_golden.py).
 *
 * All arrays use the reference's HOST layout (serial build, src/memory.h:124-130):
 *   f[(index*ndist + n)*nvel + p]        LB_ADDR, src/lb_data.h:136
 *   u/force[index*3 + a], rho[index], phi[index], grad[index*3 + a], delsq[index]
 * with index = (ic*nall[Y] + jc)*nall[Z] + kc over the halo-inclusive local
 * box, ic = 0..nall[X]-1 (src/coords.c:617-631 with nhalo = 1).
 */

#ifndef LB_ORACLE_H
#define LB_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_NVELMAX 27

enum { ORC_RELAX_M10 = 0, ORC_RELAX_BGK = 1, ORC_RELAX_TRT = 2 };

typedef struct orc_model_s {
  int ndim;
  int nvel;
  int nhydro;                              /* 1 + ndim + ndim(ndim+1)/2 */
  int cv[ORC_NVELMAX][3];
  double wv[ORC_NVELMAX];
  double na[ORC_NVELMAX];                  /* mode normalisers */
  double ma[ORC_NVELMAX][ORC_NVELMAX];     /* modes from f   */
  double mi[ORC_NVELMAX][ORC_NVELMAX];     /* f from modes   */
  /* Forward matrix as the reference's collision actually applies it. It is
   * ma except for D3Q19, whose hand-unrolled d3q19_f2mode_chunk has a zero
   * where ma[13][4] = -1 (src/collision.c:2300, "fchunk[4]*c0"). Invisible
   * with the default M10 scheme (ghost modes are fully relaxed) but part of
   * the reference's BGK/TRT results, so parity needs it. */
  double maf[ORC_NVELMAX][ORC_NVELMAX];
} orc_model_t;

typedef struct orc_param_s {
  double rho0;
  double eta_shear;
  double eta_bulk;
  double force_global[3];
  int    nrelax;
  /* binary fluid (symmetric free energy) */
  double mobility;
  double a, b, kappa;
  /* lb_collision_ghost_modes_off (src/collision.c:1137, input key "lb_ghost_modes off"):
   * no ghost-mode noise is drawn, and the binary collision relaxes the ghosts
   * fully whatever the scheme (src/collision.c:1607-1614). 0 = ghost modes on. */
  int ghost_off;
} orc_param_t;

typedef struct orc_box_s {
  int nlocal[3];
  int nall[3];       /* nlocal + 2 */
  long nsite;
} orc_box_t;

int  orc_model_init(int nvel, orc_model_t * model);
int  orc_box_init(const int nlocal[3], orc_box_t * box);
long orc_index(const orc_box_t * box, int ic, int jc, int kc); /* 1-based */

int orc_rtau(const orc_model_t * model, const orc_param_t * prm,
	     double rtau[ORC_NVELMAX]);

int orc_equilib_set(const orc_model_t * model, int ndist, double * f,
		    long index, double rho, const double u[3]);

int orc_collide_mrt(const orc_model_t * model, const orc_box_t * box,
		    const orc_param_t * prm, int ndist, double * f,
		    const double * force, double * rho, double * u);

int orc_collide_binary(const orc_model_t * model, const orc_box_t * box,
		       const orc_param_t * prm, double * f,
		       const double * force, double * u,
		       const double * phi, const double * grad,
		       const double * delsq);

int orc_halo_periodic(const orc_box_t * box, int nfel, double * data);
int orc_propagate(const orc_model_t * model, const orc_box_t * box,
		  int ndist, const double * f, double * fprime);

int orc_phi_from_g(const orc_model_t * model, const orc_box_t * box,
		   const double * f, double * phi);
int orc_grad_7pt(const orc_box_t * box, const double * phi,
		 double * grad, double * delsq);
int orc_grad_27pt(const orc_box_t * box, const double * phi,
		  double * grad, double * delsq);
/* which stencil orc_step uses for the binary fluid: 7 (default) or 27 */
int orc_gradient_select(int npt);

/* Isothermal fluctuations (lb_oracle_noise.c). state[index*4 + ia] holds the
 * KISS99 words of every site of the halo-inclusive box. While a state is
 * attached, orc_collide_mrt / orc_collide_binary / orc_step add the random stress
 * and ghost-mode terms at temperature kt and advance the generators of the
 * sites they visit; orc_noise_attach(NULL, 0) switches them off again. */
unsigned int orc_noise_uniform(unsigned int state[4]);
int orc_noise_init(const orc_box_t * box, int master_seed,
		   const int noffset[3], const int ntotal[3],
		   unsigned int * state);
int orc_noise_reap_n(unsigned int * state, long index, int nmax, double * reap);
int orc_noise_attach(unsigned int * state, double kt);
unsigned int * orc_noise_attached(double * kt);
int orc_fluctuations(const orc_model_t * model, const double * rtau, double kt,
		     int ghost_off, unsigned int * state, long index,
		     double shat[3][3], double * ghat);

/* Flat no-slip walls (lb_oracle_wall.c; src/wall.c). While a wall is attached
 * orc_step runs the bounce-back on links between halo and propagation. */
typedef struct orc_wall_s {
  int isboundary[3];      /* boundary_walls */
  double ubot[3];         /* boundary_speed_bottom (used for a single pair of walls) */
  double utop[3];
  double fnet[3];         /* momentum handed to the walls so far (wall_momentum) */
  /* partial slip (boundary_walls_slip_*; src/wall.c:284-322,489-600,1118-1210): fraction
   * s of specular reflection per wall; edges take the mean of their two walls, corners
   * and links without a tangential component stay no-slip. Walls do not move then. */
  int slip_active;
  double sbot[3];
  double stop[3];
} orc_wall_t;
int orc_wall_attach(orc_wall_t * wall);
orc_wall_t * orc_wall_attached(void);
int orc_wall_nlink(const orc_model_t * model, const orc_box_t * box,
		   const orc_wall_t * wall);
int orc_wall_bbl(const orc_model_t * model, const orc_box_t * box,
		 const orc_param_t * prm, int ndist, orc_wall_t * wall, double * f);

/* grad_3d_27pt_solid_d2 next to flat walls with neutral wetting (lb_oracle_wall.c);
 * orc_gradient_select(28) makes orc_step use it (fd_gradient_calculation 3d_27pt_solid) */
int orc_grad_27pt_solid(const orc_box_t * box, const orc_wall_t * wall, const double * phi,
			double * grad, double * delsq);

/* Open boundaries for the fluid, inflow/outflow of type "rhou" along the one direction
 * without walls (lb_oracle_wall.c; src/lb_bc_inflow_rhou.c, src/lb_bc_outflow_rhou.c).
 * While attached, orc_step applies them after the halo swap (src/ludwig.c:820-831). */
typedef struct orc_open_s {
  int id;                 /* flow direction (0, 1, 2) */
  double u0[3];           /* lb_bc_inflow_rhou_u0 */
  double rho0;            /* lb_bc_outflow_rhou_rho0 */
} orc_open_t;
int orc_open_attach(orc_open_t * open);
orc_open_t * orc_open_attached(void);
int orc_open_apply(const orc_model_t * model, const orc_box_t * box, int ndist,
		   const orc_open_t * open, double * rho, double * u, double * f);

/* Lees-Edwards planes (lb_oracle_le.c; src/model_le.c, src/leesedwards.c). While
 * attached, orc_step advances t_current by one at the start of every step
 * (physics_control_next_step, src/ludwig.c:530) and applies the boundary conditions
 * to the post-collision populations (src/ludwig.c:810). */
typedef struct orc_le_s {
  int nplanes;            /* N_LE_plane */
  double uy;              /* LE_plane_vel: velocity jump across a plane */
  int oscillatory;        /* 0 = steady shear */
  int period;             /* LE_oscillation_period (time steps) */
  int nt0;                /* LE_time_offset */
  int t_current;          /* physics_control_timestep */
  int t_start;            /* first step of this run (0 unless restarted): the order parameter
			   * is displaced for the time t_start + t_current - 1 of
			   * physics_control_time (src/physics.c:651-659), the distributions for
			   * the step number t_current */
} orc_le_t;
int orc_le_attach(orc_le_t * le);
orc_le_t * orc_le_attached(void);
int orc_le_plane_location(const orc_box_t * box, const orc_le_t * le, int np);
int orc_le_apply(const orc_model_t * model, const orc_box_t * box, int ndist,
		 const orc_le_t * le, double t, double * f);
/* the order parameter across the planes (binary fluid, one-site halo): the buffer
 * planes of field_leesedwards, buf[2 nplanes][nall_y][nall_z], and field_grad_compute
 * with them (npt = 7 or 27) */
int orc_le_field_buffers(const orc_box_t * box, const orc_le_t * le, double t,
			 const double * phi, double * buf);
int orc_le_field_grad(const orc_box_t * box, const orc_le_t * le, double t, int npt,
		      const double * phi, double * grad, double * delsq);

/* Finite-difference order parameter (lb_oracle_ch.c; "free_energy symmetric":
 * src/phi_force.c, src/phi_force_stress.c, src/phi_grad_mu.c, src/advection.c,
 * src/phi_cahn_hilliard.c). While attached, orc_step with ndist = 1 runs field_halo,
 * the gradient, phi_force_calculation and phi_cahn_hilliard ahead of the collision;
 * A, B, kappa and the mobility come from orc_param_t. */
enum { ORC_FORCE_NONE = 1, ORC_FORCE_STRESS_DIVERGENCE = 2, ORC_FORCE_PHI_GRADMU = 3 };  /* fe_force_method_enum_t */
typedef struct orc_ch_s {
  int order;              /* fd_advection_scheme_order: 1, 2 or 3 */
  int method;             /* fe_force_method */
  double grad_mu[3];      /* external chemical potential gradient ("grad_mu") */
} orc_ch_t;
int orc_ch_attach(orc_ch_t * ch);
orc_ch_t * orc_ch_attached(void);
int orc_ch_stress(const orc_box_t * box, const orc_param_t * prm,
		  const double * phi, const double * grad, const double * delsq,
		  double * pth);
int orc_ch_force_divergence(const orc_box_t * box, const double * pth, double * force);
int orc_ch_force_gradmu(const orc_box_t * box, const orc_param_t * prm,
			const double grad_mu[3], const double * phi,
			const double * delsq, double * force);
int orc_ch_advection(const orc_box_t * box, int order, const double * u,
		     const double * phi, double * flux);
int orc_ch_flux_mu(const orc_box_t * box, const orc_param_t * prm,
		   const double grad_mu[3], const double * phi, const double * delsq,
		   double * flux);
int orc_ch_update(const orc_box_t * box, const double * flux, double * phi);
int orc_phi_force(const orc_box_t * box, const orc_param_t * prm, const orc_ch_t * ch,
		  const double * phi, double * grad, double * delsq, double * force);
int orc_cahn_hilliard(const orc_box_t * box, const orc_param_t * prm, const orc_ch_t * ch,
		      double * phi, double * delsq, double * u);
/* the same route across Lees-Edwards planes (lb_oracle_ch.c, second half): one step's
 * field_halo + field_grad_compute [+ phi_force_calculation] [+ phi_cahn_hilliard] at
 * t = physics_control_time; interior values of phi, u, force in the halo-1 layout */
int orc_ch_le_step(const orc_box_t * box, const orc_param_t * prm, const orc_ch_t * ch,
		   const orc_le_t * le, double t, int npt, int do_force, int do_update,
		   double * phi, const double * u, double * force, double * grad, double * delsq);


/* Viscosity model "arrhenius" (src/visc_arrhenius.c): eta = eta_minus^((1 - phi/phistar)/2)
 * eta_plus^((1 + phi/phistar)/2). While an eta array is attached, orc_collide_mrt takes each
 * site's viscosity from it (src/collision.c:386-404); with parameters attached as well,
 * orc_step refreshes eta from phi ahead of every collision (src/ludwig.c:794). */
typedef struct orc_visc_s {
  double eta_plus;
  double eta_minus;
  double phistar;
} orc_visc_t;
int orc_visc_attach(const orc_visc_t * visc, const double * eta);
const orc_visc_t * orc_visc_attached(const double ** eta);
int orc_visc_arrhenius_update(const orc_box_t * box, const orc_visc_t * visc,
			      const double * phi, double * eta);

int orc_fill(const orc_box_t * box, int nf, const double * v, double * data);

/* stats[0..7] = sum rho, sum rho^2 (about mean is left to caller), min, max,
 * momentum x, y, z, sum phi. Interior sites only. */
int orc_stats(const orc_model_t * model, const orc_box_t * box, int ndist,
	      const double * f, double stats[8]);

/* One or more full time steps in the order of src/ludwig.c:530-862. All
 * work arrays are caller-provided; f and fprime are swapped internally and
 * the result is always left in f. Returns 0. */
int orc_step(const orc_model_t * model, const orc_box_t * box,
	     const orc_param_t * prm, int ndist, int nsteps,
	     double * f, double * fprime, double * rho, double * u,
	     double * force, double * phi, double * grad, double * delsq);

#ifdef __cplusplus
}
#endif

#endif

/*
 * b200lb.h -- C ABI of libb200lb.so: the B200 (sm_100a) implementation of
 * Ludwig's lattice-Boltzmann time-step hot path.
 *
 * This is the drop-in boundary. Every entry point is plain C (pointers, ints,
 * doubles; no CUDA or torch types) and replaces one reference function or
 * one tdpLaunchKernel dispatch (reference: zazu29/ludwig v0.20.1, paths relative
 * to its root). INTEGRATION.md shows the few lines a Ludwig maintainer adds to
 * src/collision.c, src/propagation.c, src/model.c ... to forward to these.
 *
 * Conventions
 *   - all functions return 0 on success, non-zero on failure; the message of
 *     the last failure on the calling thread is b200lb_last_error().
 *   - there is NO CPU fallback: without a CUDA device b200lb_create() fails.
 *   - host arrays use the reference's host layouts (src/memory.h:124-130,187):
 *       B200LB_HOST_AOS  f[(index*ndist + n)*nvel + p]   (default host build)
 *       B200LB_HOST_SOA  f[(n*nvel + p)*nsite + index]   (-DADDR_SOA builds)
 *       B200LB_HOST_AOSOA(V)  f[((index/V)*ndist*nvel + n*nvel + p)*V + index%V]
 *                        (-DADDR_AOSOA -DVVL=V builds, src/memory.h:260-284: blocks of V sites;
 *                        nsite must be a multiple of V)
 *     rank-1 fields (u, force, grad) likewise [index*3 + a] or [a*nsite + index];
 *     index = cs_index(ic,jc,kc) over the halo-inclusive local box with
 *     nhalo = 1 (src/coords.c:617-631). The device layout is private.
 *   - work is enqueued on the handle's own CUDA streams; functions that hand
 *     data to the host synchronise first, so callers see the reference's
 *     synchronous semantics.
 */

#ifndef B200LB_H
#define B200LB_H

#ifdef __cplusplus
extern "C" {
#endif

#define B200LB_NVELMAX 27

enum { B200LB_HOST_AOS = 0, B200LB_HOST_SOA = 1, B200LB_HOST_AOSOA_BASE = 256 };
#define B200LB_HOST_AOSOA(V) (B200LB_HOST_AOSOA_BASE + (V))

/* lb_relaxation_enum_t, src/lb_data_options.h:18-23 */
enum { B200LB_RELAX_M10 = 0, B200LB_RELAX_BGK = 1, B200LB_RELAX_TRT = 2 };

/* lb_halo_enum_t, src/lb_data_options.h:26-28. TARGET = all velocities on six
 * faces, x then y then z (src/halo_swap.c:709-1063). REDUCED = only the
 * velocities that propagate into the receiving domain (src/model.c:1192-1219). */
enum { B200LB_HALO_FULL = 0, B200LB_HALO_REDUCED = 2 };

/* Fields that can be moved across the boundary (hydro_t src/hydro.h:33-49,
 * field_t phi src/field.h:55-79, field_grad_t src/field_grad.h:24-43). */
enum {
  B200LB_FIELD_RHO = 0,      /* 1 double / site */
  B200LB_FIELD_U = 1,        /* 3 */
  B200LB_FIELD_FORCE = 2,    /* 3 */
  B200LB_FIELD_PHI = 3,      /* 1 */
  B200LB_FIELD_GRAD = 4,     /* 3 */
  B200LB_FIELD_DELSQ = 5,    /* 1 */
  B200LB_FIELD_ETA = 6       /* 1: hydro->eta, the per-site shear viscosity of a viscosity model */
};

typedef struct b200lb_s b200lb_t;        /* opaque handle: one local lattice */

/* What lb_data_create + cs_t know (src/model.c:56, src/coords.h:30-63). */
typedef struct b200lb_desc_s {
  int ndim;              /* 2 or 3 */
  int nvel;              /* 9, 19 or 27 */
  int ndist;             /* 1, or 2 for the binary fluid (f and g) */
  int nlocal[3];         /* local interior extent; nlocal[2] = 1 when ndim = 2 */
  int nhalo;             /* must be 1 */
  int cartsz[3];         /* ranks per direction (cs_cartsz) */
  int cartcoords[3];     /* this rank's Cartesian coordinates */
  int periodic[3];       /* cs_periodic */
  int nbr_lo[3];         /* rank of the lower neighbour per direction */
  int nbr_hi[3];         /* rank of the upper neighbour per direction */
  int rank;              /* this rank in the communicator */
  int nranks;
  int device;            /* CUDA device ordinal, or -1 for the current one */
  int zalign;            /* 0 = default; interior rows start on zalign doubles */
  int lattices;          /* 0 or 2: f and fprime; 1: single-lattice streaming from the start
                          * (option "lattices"): the second array is never allocated */
} b200lb_desc_t;

/* lb_collide_param_t (src/lb_data.h:53-69), collide_param_t
 * (src/collision.c:79-89) and fe_symm_param_t (src/symmetric.h:40-46):
 * what lb_collide gathers from physics_ref() and commits every step. */
typedef struct b200lb_param_s {
  double rho0;
  double eta_shear;
  double eta_bulk;
  double force_global[3];    /* fbody + fpulse sin(...) (src/collision.c:1945-1970) */
  int    nrelax;             /* B200LB_RELAX_* */
  double mobility;           /* binary: rtau2 = 2/(1 + 2 mobility) */
  double a, b, kappa;        /* symmetric free energy */
  double kt;                 /* physics_kt: temperature of the isothermal fluctuations */
  int    ghost_off;          /* lb_collision_ghost_modes_off (src/collision.c:1137, key
                              * "lb_ghost_modes off"): no ghost-mode noise; the binary
                              * collision relaxes the ghosts fully (src/collision.c:1607) */
} b200lb_param_t;

const char * b200lb_last_error(void);
int  b200lb_device_count(void);
const char * b200lb_version(void);

int  b200lb_create(const b200lb_desc_t * desc, b200lb_t ** h);  /* lb_data_create + hydro_create (+ field/field_grad when ndist = 2) */
int  b200lb_destroy(b200lb_t * h);                              /* lb_free, src/model.c:180 */
long b200lb_nsite(const b200lb_t * h);                          /* host nsite = prod(nlocal + 2) */

/* run-time options; recognised names:
 *   "fused"       0/1  collide-and-stream in one pass (lazy propagation)
 *   "lattices"    1/2  2 (default): two lattices, f and fprime of lb_t. 1: single-lattice
 *                      streaming (AA pattern): half the memory; lb_collide alternates between
 *                      an in-place sweep and a sweep that pulls, collides and stores one
 *                      propagation ahead. For the time loop's cycle lb_collide - lb_halo -
 *                      lb_propagation on a single-fluid lattice periodic within the rank;
 *                      the lattice can be looked at (lb_memcpy, statistics) after
 *                      lb_propagation, other sequences are refused. Same results as 2
 *                      to round-off of nothing: bit-identical
 *   "halo"        B200LB_HALO_FULL / B200LB_HALO_REDUCED
 *   "hydro_write" 0/1/2  1: the collision stores rho,u every step (the reference's
 *                      behaviour, src/collision.c:570-579); 0: never -- they are recomputed
 *                      from the post-collision lattice when somebody asks (identical to
 *                      round-off); 2 (default): the collision stores them when hydro
 *                      quantities were asked for since the previous collision, so a caller
 *                      that looks every step gets 1 and a time loop that only looks at
 *                      statistics steps gets 0 in between
 *   "vec"         1/2  lattice sites per thread
 *   "block"       threads per block along the contiguous direction
 *   "threads"     threads per block of the sweep in all: 0 (default: 256, or 128 for the
 *                      binary fluid and the schemes that keep ghost modes), 64, 128, 256
 *   "halo_lazy"   0/1  defer the local periodic halo copies until the host looks
 *   "overlap"     0/1  exchange halos on a second stream next to the interior sweep
 *   "direct"      0/1  reduced halo as one message per neighbour direction
 *   "diag"        0/1  gather conserved sums inside the collision sweep
 *   "grad_lazy"   0/1  binary fluid: evaluate the 7-point stencil inside the collision
 *                      and fill the grad/delsq arrays only when they are asked for
 *   "gradient"    7/27 which stencil b200lb_step uses for the order parameter
 *                      (fd_gradient_calculation 3d_7pt_fluid / 3d_27pt_fluid, src/gradient_rt.c:81-98)
 *   "ch_fuse"     0/1  finite-difference order parameter: phi_force_calculation is
 *                      evaluated by the pass of the phi_cahn_hilliard that follows it
 *   "ch_planes"   n    ... x-planes one block of that kernel marches (default 8)
 *   "ch_exact"    0/1  ... 1: the kernel build without fused multiply-add that keeps the
 *                      reference's order of operations (force and phi bit-identical to
 *                      the reference's); 0 (default): fused multiply-add, a few ulp away
 *   "g_moment"    0/1  binary fluid, fused mode: phi_lb_to_field also stores the first
 *                      moment of the (pulled) order-parameter populations, and the
 *                      collision that streams in the same pass reads that instead of
 *                      pulling g a second time (same bits, 104 B/site less traffic;
 *                      default 1)
 *   "binary_walls" 0/1 allow flat walls with the binary fluid (ndist = 2): the bounce-back
 *                      of both distributions is folded into a propagation of its own,
 *                      the collision runs on the streamed lattice, the gradient is
 *                      b200lb_grad27_solid ("gradient" = 28). Default 1; with 0
 *                      b200lb_wall_commit refuses ndist = 2
 *   "le_fd"       0/1  allow Lees-Edwards planes together with the finite-difference order
 *                      parameter (buffer-region gradient, flux form of the force with its
 *                      plane correction, phi_ch_le_fix_fluxes: ch_le_pass_a/_b). Default 1;
 *                      with 0 b200lb_le_commit / b200lb_ch_create refuse the combination
 *   "pdl"         0/1  peer-memory halo: launch the sweep that follows the synchronisation
 *                      kernel with programmatic stream serialisation, so that its interior
 *                      starts under that kernel and only threads on a decomposed face wait
 *                      (griddepcontrol.wait). Correct (tests pass with it) but measured to gain
 *                      nothing on B200 (0.8066 -> 0.8057 ms): default 0
 *   "zpack"       0/1/2 peer-memory halo, single fluid on two lattices: what crosses a face of
 *                      the contiguous direction goes to dense planes behind the neighbour's
 *                      lattice array and is pulled from one's own (1, default: one word
 *                      per column; 2: a 32-byte sector per column) instead of the row-end
 *                      halo words (0). Same results; anything that looks at halo sites
 *                      gets them unpacked first
 *   "loopback"    0/1  single rank: send periodic directions through the exchange
 *                      buffers with this GPU as its own neighbour (tests, profiling) */
int  b200lb_set_option(b200lb_t * h, const char * name, long value);
long b200lb_get_option(const b200lb_t * h, const char * name);

/* lb_memcpy (src/model.c:228), hydro_memcpy (src/hydro.c:154), field_memcpy,
 * field_grad_memcpy: layout boundary between host arrays and the device. */
int  b200lb_f_upload(b200lb_t * h, const double * f_host, int host_layout);
/* lb_io_read piecewise (src/model.c:1568-1649 goes through an aggregator buffer, never one
 * array for the whole lattice): host sites [first, first + count) in the AOS layout,
 * f[(site - first)*ndist*nvel + n*nvel + p]. Pieces follow each other from site 0 to
 * b200lb_nsite(); a 514^3 lattice (20.6 GB) then needs no 20.6 GB host array. */
int  b200lb_f_upload_range(b200lb_t * h, const double * f, long first, long count);
int  b200lb_f_download(b200lb_t * h, double * f_host, int host_layout);
int  b200lb_field_upload(b200lb_t * h, int field, const double * host, int host_layout);
int  b200lb_field_download(b200lb_t * h, int field, double * host, int host_layout);

/* lb_collide_param_commit + lb_collision_parameters_commit
 * (src/model.c:342, src/collision.c:1928-1980) */
int  b200lb_set_param(b200lb_t * h, const b200lb_param_t * prm);

/* The hot path. Names follow the reference call surface (src/ludwig.c:530-862). */
int  b200lb_hydro_f_zero(b200lb_t * h, const double f[3]);   /* hydro_f_zero  src/hydro.c:306 */
int  b200lb_hydro_u_zero(b200lb_t * h, const double u[3]);   /* hydro_u_zero  src/hydro.c:279 */
int  b200lb_hydro_rho0(b200lb_t * h, double rho0);           /* hydro_rho0    src/hydro.c:335 */
int  b200lb_hydro_correct_momentum(b200lb_t * h);            /* hydro_correct_momentum src/hydro.c:829 */
int  b200lb_collide(b200lb_t * h);                           /* lb_collide    src/collision.c:144 */
int  b200lb_halo(b200lb_t * h);                              /* lb_halo       src/model.c:553 */
int  b200lb_halo_swap(b200lb_t * h, int scheme);             /* lb_halo_swap  src/model.c:570 */
int  b200lb_propagate(b200lb_t * h);                         /* lb_propagation src/propagation.c:43 */
int  b200lb_phi_from_g(b200lb_t * h);                        /* phi_lb_to_field src/phi_lb_coupler.c:39 */
int  b200lb_phi_halo(b200lb_t * h);                          /* field_halo(phi) src/field.c:434 */
int  b200lb_grad7(b200lb_t * h);                             /* grad_3d_7pt_fluid_d2 src/gradient_3d_7pt_fluid.c:75 */
int  b200lb_grad27(b200lb_t * h);                            /* grad_3d_27pt_fluid_d2 src/gradient_3d_27pt_fluid.c:75 */
int  b200lb_grad27_solid(b200lb_t * h);                                /* grad_3d_27pt_solid_d2 src/gradient_3d_27pt_solid.c:131: next to flat walls, neutral wetting */

/* Flat no-slip walls ("boundary_walls", "boundary_speed_bottom/_top"; src/wall.c,
 * src/wall_rt.c:40-120). A wall in a non-periodic direction occupies the halo layer at
 * global coordinate 0 and ntotal + 1 (wall_init_map). wall_bbl belongs between lb_halo
 * and lb_propagation (src/ludwig.c:837-842): the propagation that follows -- on its own
 * or fused into the next collision -- bounces the populations that would cross into a
 * wall back to the site they left (bounce-back on links, wall_bbl_kernel
 * src/wall.c:997-1107, moving-wall term for a single pair of walls included), and the
 * momentum handed to the walls is tallied (wall->fnet). Single fluid only; porous media
 * and colloids are not on this path. */
int  b200lb_wall_commit(b200lb_t * h, const int isboundary[3], const double ubot[3],
                        const double utop[3]);                         /* wall_create + wall_commit src/wall.c:59,168 */
/* partial slip, "boundary_walls_slip_fraction_bot/_top", "boundary_walls_slip_active":
 * fraction s of specular reflection per wall (faces: that wall's s, edges: the mean of
 * the two, corners: none), wall_bbl_slip_kernel src/wall.c:1118-1210. After wall_commit. */
int  b200lb_wall_slip_set(b200lb_t * h, const double sbot[3], const double stop[3],
                          int force_active);                           /* wall_slip src/wall.c:284 */
int  b200lb_wall_bbl(b200lb_t * h);                                    /* wall_bbl src/wall.c:960 */
/* Open boundaries for the fluid ("lb_bc_open yes", lb_bc_inflow_type / lb_bc_outflow_type
 * rhou: lb_bc_inflow_rhou_create / lb_bc_outflow_rhou_create, chosen in
 * src/lb_bc_open_rt.c:48-170): flow along the one direction without walls (commit the walls
 * first), inlet at its low end with velocity u0 and the density of the fluid behind it,
 * outlet at the high end with density rho0 and the velocity of the fluid in front of it.
 * b200lb_open_bc_apply replaces inflow/outflow->func->update + ->impose between lb_halo
 * and lb_propagation (src/ludwig.c:820-831; host loops over link lists in the reference):
 * the populations the propagation will carry in from the two halo layers are set to
 * equilibrium. flow_dir < 0 switches them off. b200lb_step applies them itself. */
int  b200lb_open_bc_commit(b200lb_t * h, int flow_dir, const double u0[3], double rho0);
int  b200lb_open_bc_apply(b200lb_t * h);
int  b200lb_wall_momentum(b200lb_t * h, double f[3]);                  /* wall_momentum src/wall.c:1299 */
int  b200lb_wall_momentum_add(b200lb_t * h, const double f[3]);        /* wall_momentum_add src/wall.c:1279 */

/* Lees-Edwards sliding periodic boundaries for the distributions ("N_LE_plane",
 * "LE_plane_vel", "LE_oscillation_period", "LE_time_offset": lees_edw_options_t,
 * src/lees_edwards_options.h:25-34; lees_edw_create src/leesedwards.c:80).
 * nplanes_total planes normal to x, uniformly spaced, velocity jump uy along y;
 * oscillatory != 0: uy cos(2 pi (t - nt0)/period). b200lb_le_apply replaces
 * lb_le_apply_boundary_conditions (src/model_le.c:57): called between lb_collide and
 * lb_halo (src/ludwig.c:810) with t = the physics time (the step number), it reprojects
 * and displaces the populations about to cross a plane -- on the device, where the
 * reference copies the whole lattice to the host and back. Refused, as the checks of
 * src/leesedwards.c:416-456 do: planes that do not divide the system or the ranks in x,
 * a plane within a halo width of a rank boundary; and a decomposition in y (the
 * reference's message-passing variant, src/model_le.c:399) which is not on this path.
 * b200lb_step advances its own step counter (b200lb_le_time_set, 0 at le_commit) and
 * applies the planes after every collision, as the time loop of src/ludwig.c does. */
int  b200lb_le_commit(b200lb_t * h, int nplanes_total, double uy, int oscillatory,
                      int period, int nt0);
int  b200lb_le_apply(b200lb_t * h, double t);                          /* lb_le_apply_boundary_conditions src/model_le.c:57 */
int  b200lb_le_plane_location(const b200lb_t * h, int np);             /* lees_edw_plane_location src/leesedwards.c:599 */
int  b200lb_le_time_set(b200lb_t * h, int nstep);                      /* physics_control_init_time src/physics.c:687: the run (re)starts at this step */
/* Binary fluid (ndist = 2) with planes: field_grad_compute (b200lb_grad7 / b200lb_grad27)
 * first does what field_leesedwards does (src/field.c:500-600, called from
 * src/field_grad.c:310): two buffer planes of phi per plane, displaced along y by what the
 * plane has slid at physics_control_time -- t_start + t_current - 1 (src/physics.c:651-659),
 * one less than the step number the distributions are displaced by -- with four-point
 * interpolation; the stencil of the sites next to a plane then takes its x neighbours from
 * them (lees_edw_ic_to_buff, src/gradient_3d_7pt_fluid.c:264-265,
 * src/gradient_3d_27pt_fluid.c:262-265). b200lb_le_field_time_set hands over that time
 * ahead of the gradient call; b200lb_step keeps it itself. Planes closer than three sites
 * are refused for the order parameter. */
int  b200lb_le_field_time_set(b200lb_t * h, double t);                 /* physics_control_time as read by field_leesedwards, src/field.c:553 */

/* The order parameter as a finite-difference field ("free_energy symmetric",
 * src/ludwig.c:1159-1250): ndist = 1, phi is a field_t that follows Cahn-Hilliard with the
 * velocity of the previous collision, the fluid feels it as a body force. Per time step
 * (src/ludwig.c:558-575, 727-768):
 *   field_halo(phi); field_grad_compute; phi_force_calculation; phi_cahn_hilliard
 * = b200lb_phi_halo; b200lb_grad7 | b200lb_grad27; b200lb_phi_force; b200lb_cahn_hilliard.
 * The reference makes seven passes over the lattice for the last two (stress, its
 * divergence, advective fluxes, diffusive fluxes twice, no-flux mask, update) and keeps
 * a halo two sites wide for them; here they are ONE kernel reading phi, grad phi,
 * delsq phi and u at a site and its six neighbours (option "ch_fuse" = 0: two
 * launches). Free-energy parameters A, B, kappa and the mobility are those of
 * b200lb_set_param. With option "ch_exact" results are bit-identical to the reference's;
 * the default build uses fused multiply-add (a few ulp away, 1.6x faster).
 * On this path: three-dimensional models (the lattice may be one site thick), fully periodic, fluid only, Lees-Edwards planes (option "le_fd");
 * advection orders 1-4 (3 and 4 only where every direction is periodic within the rank:
 * they read phi two sites away; 4 is a host loop behind two copies of the flux arrays in
 * the reference, src/advection.c:465-470); force methods no_force, stress_divergence, phi_gradmu.
 * Fields: B200LB_FIELD_PHI / _GRAD / _DELSQ move the order parameter and its gradient,
 * index over the usual one-site halo box (the interior is what the reference's
 * nhalo = 2 arrays hold inside their halo). */
enum { B200LB_FORCE_NONE = 1, B200LB_FORCE_STRESS_DIVERGENCE = 2, B200LB_FORCE_PHI_GRADMU = 3 };  /* fe_force_method_enum_t src/fe_force_method.h:20-28 */
int  b200lb_ch_create(b200lb_t * h, int advection_order, int force_method);   /* phi_ch_create src/phi_cahn_hilliard.c:121, pth_create src/phi_force_stress.c:41, advection_order_set src/advection.c:93 */
int  b200lb_ch_grad_mu_set(b200lb_t * h, const double grad_mu[3]);            /* physics_grad_mu_set ("grad_mu") */
int  b200lb_phi_force(b200lb_t * h);                                          /* phi_force_calculation src/phi_force.c:74 */
int  b200lb_cahn_hilliard(b200lb_t * h);                                      /* phi_cahn_hilliard src/phi_cahn_hilliard.c:206 */
int  b200lb_hydro_u_halo(b200lb_t * h);                                       /* hydro_u_halo src/hydro.c */

/* Viscosity model ("viscosity_model arrhenius", src/ludwig.c:2001-2040; visc_t src/visc.h).
 * While a model is present lb_collide (single fluid) takes the shear viscosity of every
 * site from hydro->eta, the bulk viscosity in the Newtonian ratio eta_bulk/eta_shear, and
 * all relaxation times that depend on them (src/collision.c:386-404,
 * lb_relaxation_time_shear_v/_bulk_v/_ghosts_v). b200lb_visc_update is the model's update
 * (visc->func->update, src/ludwig.c:794): eta = eta_minus^((1 - phi/phistar)/2)
 * eta_plus^((1 + phi/phistar)/2) from the order parameter (src/visc_arrhenius.c:185-220).
 * b200lb_visc_model_set(h, 1) alone makes the collision read whatever the caller put in
 * B200LB_FIELD_ETA (a model evaluated on the host). Not with isothermal fluctuations
 * (their variances follow one viscosity). b200lb_step runs the update before each collision. */
int  b200lb_visc_arrhenius_commit(b200lb_t * h, double eta_plus, double eta_minus, double phistar);   /* visc_arrhenius_create src/visc_arrhenius.c:52 */
int  b200lb_visc_update(b200lb_t * h);                                  /* visc_arrhenius_update src/visc_arrhenius.c:149 */
int  b200lb_visc_model_set(b200lb_t * h, int on);                       /* lb_collide(..., visc): p.have_visc_model src/collision.c:1954 */

/* Isothermal fluctuations ("isothermal_fluctuations on", src/collision_rt.c:55):
 * every site owns a KISS99 generator seeded from its global position, so results do
 * not depend on the decomposition (noise_create + noise_init, src/noise.c:56,136;
 * ludwig.c:243 calls it with master_seed 0). While present, lb_collide adds the random
 * stress and ghost-mode terms of src/collision.c:476-515 / :1663-1745 at temperature
 * param.kt and advances the generators of the interior sites it collides (the
 * reference also advances the y,z halo images, which nothing reads).
 * D2Q9 and D3Q19; D3Q27 has more ghost modes than one draw yields (src/noise.h:18).
 * state: the reference's host array noise->state, state[index*4 + ia] (AOS) or
 * state[ia*nsite + index] (SOA), index over the halo-inclusive box (noise_memcpy,
 * noise I/O for restarts). */
int  b200lb_noise_init(b200lb_t * h, int master_seed);                /* noise_init src/noise.c:136 */
int  b200lb_noise_present_set(b200lb_t * h, int on);                  /* noise_present_set(NOISE_RHO) src/noise.c:557 */
int  b200lb_noise_state_upload(b200lb_t * h, const unsigned int * state, int host_layout);   /* noise_memcpy H2D src/noise.c:253 */
int  b200lb_noise_state_download(b200lb_t * h, unsigned int * state, int host_layout);       /* noise_memcpy D2H */

/* nsteps iterations of the src/ludwig.c:530-862 loop restricted to the hot
 * path: f_zero; [phi_from_g; phi_halo; grad7 | grad27]; [after b200lb_ch_create: phi_halo;
 * grad7 | grad27; phi_force; cahn_hilliard]; u_zero; collide; halo; [wall_bbl]; propagate */
int  b200lb_step(b200lb_t * h, int nsteps);

/* Conserved-quantity diagnostics over interior sites of this rank
 * (stats_distribution_print / stats_distribution_momentum,
 * src/stats_distribution.c:55-118,201-265):
 * out = { sum rho, sum rho^2, min rho, max rho, gx, gy, gz, sum phi } */
int  b200lb_stats(b200lb_t * h, double out[8]);

/* The same sums gathered inside the last lb_collide sweep (option "diag" = 1):
 * warp-shuffle + per-block partials in the sweep itself, folded here. They describe
 * the state that ENTERED that collision (= the lattice after the previous
 * lb_propagation); momentum is sum_p c_p f_p. Costs no extra pass over the lattice. */
int  b200lb_step_diag(b200lb_t * h, double out[8]);

/* Device reductions of a field over the interior sites of this rank, per component c:
 * out[4c + 0..3] = sum, sum of squares, min, max. B200LB_FIELD_U gives what
 * stats_velocity_minmax reports (src/stats_velocity.c:43-107), B200LB_FIELD_PHI what
 * stats_field_info / stats_field_q_reduce sums (src/phi_stats.c:89-230), without the
 * full-field D2H of src/ludwig.c:955,2335; the caller finishes with its MPI_Reduce. */
int  b200lb_field_stats(b200lb_t * h, int field, double * out);

/* b200lb_field_stats that allocates nothing. NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid it). Same kernels, same result
 * (out[4*c + 0..3] = sum, sum of squares, min, max per component over the interior sites);
 * the per-block partial sums, the 256 x 8 intermediate sums of their fold and the result live
 * in a file-scope device array of 8*(B200LB_NOALLOC_MAXBLOCK + 1) + 2048 doubles (4 MiB +
 * 16 KiB) that every GPU loading the library holds until exit. Limits and rules:
 *   - the lattice must need at most B200LB_NOALLOC_MAXBLOCK blocks of 128 x 2 sites,
 *     ceil(n_z/128)*ceil(n_y/2)*n_x (256^3 fits exactly); larger lattices are refused;
 *   - the field's device planes must exist and hold current values: a field that was never
 *     uploaded / written, or is still recorded lazily (rho,u with hydro_write = 0, the
 *     uniform force of hydro_f_zero, a gradient not yet materialised), is refused -- making
 *     it current could allocate. b200lb_field_stats or b200lb_field_download does that;
 *   - all calls on a device share the array, which keeps the last call's contents: calls
 *     that overlap on the GPU (two handles or threads at once) race. Calls on one handle
 *     are ordered by its stream.
 * Calls no cudaMalloc, cudaFree or allocating library routine. */
#define B200LB_NOALLOC_MAXBLOCK 65536
int  b200lb_field_stats_noalloc(b200lb_t * h, int field, double * out);

/* lb_io_aggr_pack / lb_io_aggr_unpack with binary records (src/model.c:1498-1560): the
 * interior sites in (ic,jc,kc) order, kc fastest, b200lb_io_record_doubles() =
 * ndist*nvel doubles per site ordered [n][p] (lb_write_buf, src/model.c:1385) -- the
 * io_aggregator_t buffer lb_io_write hands to MPI-IO, filled straight from the device. */
long b200lb_io_record_doubles(const b200lb_t * h);
int  b200lb_io_pack(b200lb_t * h, double * buf);
int  b200lb_io_unpack(b200lb_t * h, const double * buf);

int  b200lb_sync(b200lb_t * h);          /* tdpDeviceSynchronize for this handle */

/* Multi-GPU: one rank per GPU; halo planes go device-to-device with NCCL
 * send/recv (replaces the MPI_Isend/Irecv of host buffers, src/halo_swap.c:762-1035).
 * Rank 0 makes the id, the caller's launcher (MPI_Bcast, torch.distributed ...)
 * distributes its 128 bytes, every rank calls comm_init. */
int  b200lb_nccl_unique_id(char id[128]);
int  b200lb_comm_init(b200lb_t * h, const char id[128]);

/* Peer-memory halo (one node, NVLink/NVSwitch): every rank exports CUDA IPC handles
 * of its lattices (192 bytes), the launcher gathers them from all ranks in rank
 * order, every rank imports. From then on the collision sweep stores the outgoing
 * populations of boundary sites directly into the neighbours' halo sites and
 * lb_halo (reduced scheme) has nothing left to send; option "p2p" = 0 goes back to
 * NCCL messages. Ranks must issue the same sequence of calls (SPMD, as Ludwig does). */
int  b200lb_p2p_export(b200lb_t * h, char out[192]);
int  b200lb_p2p_import(b200lb_t * h, const char * all, int nranks);

/* Timing on the handle's compute stream (CUDA events), for bench.py */
int  b200lb_timer_start(b200lb_t * h);
int  b200lb_timer_stop(b200lb_t * h, float * ms);     /* synchronises */
long b200lb_launch_count(const b200lb_t * h);         /* kernels launched so far */
/* TIMER_COLLIDE_KERNEL / TIMER_PROP_KERNEL (src/collision.c:198-207, src/propagation.c:84-91):
 * time of the sweep kernels ALONE. Between _start and _stop every sweep launch (collision,
 * propagation, the fused pass; each wall or overlap slab is one) is bracketed by CUDA events
 * on the launching stream; _stop synchronises and returns their summed duration and count.
 * Other launches of a step (halo kernels, peer synchronisation) are not in it. */
int  b200lb_kernel_timer_start(b200lb_t * h, int max_launches);
int  b200lb_kernel_timer_stop(b200lb_t * h, double * ms, long * nlaunch);

/* Pinned host memory for lb->f and friends (optional; speeds up lb_memcpy) */
int  b200lb_host_alloc(void ** ptr, long nbytes);
int  b200lb_host_free(void * ptr);

#ifdef __cplusplus
}
#endif

#endif

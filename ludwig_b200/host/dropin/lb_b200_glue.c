/* lb_b200_glue.c -- see lb_b200_glue.h. Everything here is a forward to the C ABI of
 * include/b200lb.h; no arithmetic of the path is done on the host. */

#include <assert.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "b200lb.h"
#include "lb_b200_glue.h"
#include "physics.h"
#include "symmetric.h"
#include "gradient_3d_7pt_fluid.h"
#include "gradient_3d_27pt_fluid.h"
#include "util.h"

static b200lb_t * b200_ = NULL;
static pe_t * pe_ = NULL;
static long nsite_ = 0;
static int noise_uploaded_ = 0;
static int sync_ = 0;         /* LUDWIG_B200_SYNC=1: the device is idle when a call returns,
			       * so that Ludwig's own wall-clock timers (src/timer.c) bracket
			       * the work of the call and not only its launch */

#define CK(call) do { if ((call) != 0) pe_fatal(pe_, "libb200lb: %s\n", b200lb_last_error()); } while (0)

#ifdef ADDR_SOA
#define HOST_LAYOUT B200LB_HOST_SOA
#else
#define HOST_LAYOUT B200LB_HOST_AOS
#endif

static int done(void) {
  if (sync_) CK(b200lb_sync(b200_));
  return 0;
}

int lb_b200_active(void) { return b200_ != NULL; }

int lb_b200_create(pe_t * pe, cs_t * cs, int ndim, int nvel, int ndist) {

  b200lb_desc_t d;
  int nhalo = 0;
  int nsites = 0;
  MPI_Comm comm;

  assert(pe);
  assert(cs);
  if (b200_) pe_fatal(pe, "lb_halo_b200: one distribution object per process\n");
  pe_ = pe;
  memset(&d, 0, sizeof(d));
  d.ndim = ndim; d.nvel = nvel; d.ndist = ndist;
  cs_nhalo(cs, &nhalo);
  cs_nsites(cs, &nsites);
  d.nhalo = nhalo;
  cs_nlocal(cs, d.nlocal);
  cs_cartsz(cs, d.cartsz);
  cs_cart_coords(cs, d.cartcoords);
  cs_periodic(cs, d.periodic);
  for (int ia = 0; ia < 3; ia++) {
    d.nbr_lo[ia] = cs_cart_neighb(cs, CS_BACK, ia);
    d.nbr_hi[ia] = cs_cart_neighb(cs, CS_FORW, ia);
    /* a non-periodic edge of the Cartesian communicator has no neighbour */
    if (!d.periodic[ia] && d.cartcoords[ia] == 0) d.nbr_lo[ia] = -1;
    if (!d.periodic[ia] && d.cartcoords[ia] == d.cartsz[ia] - 1) d.nbr_hi[ia] = -1;
  }
  d.rank = pe_mpi_rank(pe);
  d.nranks = pe_mpi_size(pe);
  d.device = -1;                          /* node-local rank, as src/ludwig.c:467-492 */
  if (nhalo != 1) pe_fatal(pe, "lb_halo_b200: halo width %d (the device path carries one site)\n", nhalo);
  CK(b200lb_create(&d, &b200_));
  nsite_ = b200lb_nsite(b200_);
  if (nsite_ != nsites) pe_fatal(pe, "lb_halo_b200: %ld sites against Ludwig's %d\n", nsite_, nsites);

  if (d.nranks > 1) {
    /* one rank per GPU: NCCL communicator, and on one node the neighbours' lattices
     * mapped so that the sweep writes their halo sites directly */
    char id[128];
    char mine[192];
    char * all = (char *) malloc((size_t) 192*d.nranks);
    pe_mpi_comm(pe, &comm);
    if (d.rank == 0) CK(b200lb_nccl_unique_id(id));
    MPI_Bcast(id, 128, MPI_CHAR, 0, comm);
    CK(b200lb_comm_init(b200_, id));
    CK(b200lb_p2p_export(b200_, mine));
    MPI_Allgather(mine, 192, MPI_CHAR, all, 192, MPI_CHAR, comm);
    CK(b200lb_p2p_import(b200_, all, d.nranks));
    free(all);
  }
  {
    const char * s = getenv("LUDWIG_B200_SYNC");
    sync_ = (s && atoi(s) != 0);
  }
  pe_info(pe, "Distribution back end: %s (lb_halo_b200)\n", b200lb_version());
  return 0;
}

int lb_b200_free(void) {
  if (b200_) b200lb_destroy(b200_);
  b200_ = NULL;
  return 0;
}

int lb_b200_memcpy(lb_t * lb, tdpMemcpyKind flag) {
  assert(lb);
  switch (flag) {
  case tdpMemcpyHostToDevice:
    CK(b200lb_f_upload(b200_, lb->f, HOST_LAYOUT));
    break;
  case tdpMemcpyDeviceToHost:
    CK(b200lb_f_download(b200_, lb->f, HOST_LAYOUT));
    break;
  default:
    /* device to device (src/ludwig.c:926): the lattice already is where it is used */
    break;
  }
  return done();
}

int lb_b200_halo(lb_t * lb) {
  assert(lb);
  /* the populations that stream into the receiver: the message set of
   * lb_halo_openmp_reduced (src/model.c:1192-1219). Host code behind lb_memcpy D2H
   * (statistics, i/o) reads interior sites only. */
  CK(b200lb_halo_swap(b200_, B200LB_HALO_REDUCED));
  return done();
}

int lb_b200_propagation(lb_t * lb) {
  assert(lb);
  CK(b200lb_propagate(b200_));
  return done();
}

/* lb_collide: the parameter block of lb_collision_relaxation_times_set and
 * lb_collision_parameters_commit (src/collision.c:1181-1264, 1928-1980), then the sweep */
int lb_b200_collide(lb_t * lb, hydro_t * hydro, noise_t * noise, fe_t * fe, visc_t * visc) {

  b200lb_param_t p;
  physics_t * phys = NULL;
  double fbody[3], fpulse[3], freq = 0.0;
  int noise_on = 0;
  PI_DOUBLE(pi);

  assert(lb);
  if (hydro == NULL) return 0;                     /* as src/collision.c:149 */
  if (visc) pe_fatal(pe_, "lb_halo_b200: viscosity model not wired into this build (b200lb_visc_*)\n");
  if (noise) noise_present(noise, NOISE_RHO, &noise_on);
  if (noise_on && !noise_uploaded_) {
    /* the generators Ludwig seeded (noise_init, src/ludwig.c:243) or read from a restart
     * file move to the device once; from here on they advance there */
    CK(b200lb_noise_state_upload(b200_, noise->state, HOST_LAYOUT));
    noise_uploaded_ = 1;
  }
  CK(b200lb_noise_present_set(b200_, noise_on));

  memset(&p, 0, sizeof(p));
  physics_ref(&phys);
  physics_rho0(phys, &p.rho0);
  physics_eta_shear(phys, &p.eta_shear);
  physics_eta_bulk(phys, &p.eta_bulk);
  physics_mobility(phys, &p.mobility);
  physics_kt(phys, &p.kt);
  physics_fbody(phys, fbody);
  physics_fpulse(phys, fpulse);
  physics_fpulse_frequency(phys, &freq);
  {
    const double t = 1.0*physics_control_timestep(phys);
    for (int ia = 0; ia < 3; ia++) {
      p.force_global[ia] = fbody[ia] + fpulse[ia]*sin(2.0*pi*freq*t);
    }
  }
  p.nrelax = (int) lb->nrelax;                     /* M10, BGK, TRT: same numbering */
  p.ghost_off = (lb->param->isghost == LB_GHOST_OFF);
  if (lb->ndist == 2) {
    fe_symm_param_t fp;
    assert(fe);
    if (fe->id != FE_SYMMETRIC) pe_fatal(pe_, "lb_halo_b200: two distributions go with free_energy symmetric_lb\n");
    fe_symm_param((fe_symm_t *) fe, &fp);
    p.a = fp.a; p.b = fp.b; p.kappa = fp.kappa;
  }
  CK(b200lb_set_param(b200_, &p));
  CK(b200lb_collide(b200_));
  return done();
}

int lb_b200_hydro_f_zero(const double fzero[3]) {
  CK(b200lb_hydro_f_zero(b200_, fzero));
  return 0;
}

int lb_b200_hydro_u_zero(const double uzero[3]) {
  CK(b200lb_hydro_u_zero(b200_, uzero));
  return 0;
}

/* hydro_memcpy: rho, u and the force field keep Ludwig's host arrays as interchange */
int lb_b200_hydro_memcpy(hydro_t * hydro, tdpMemcpyKind flag) {
  assert(hydro);
  if (flag == tdpMemcpyDeviceToHost) {
    CK(b200lb_field_download(b200_, B200LB_FIELD_RHO, hydro->rho->data, HOST_LAYOUT));
    CK(b200lb_field_download(b200_, B200LB_FIELD_U, hydro->u->data, HOST_LAYOUT));
    CK(b200lb_field_download(b200_, B200LB_FIELD_FORCE, hydro->force->data, HOST_LAYOUT));
  }
  /* host to device: the host arrays are zero until the first collision (src/hydro.c:56);
   * the library's fields start the same way */
  return done();
}

int lb_b200_is_phi(const field_t * field) {
  return b200_ != NULL && field != NULL && field->nf == 1 && field->name != NULL
    && strcmp(field->name, "phi") == 0;
}

int lb_b200_phi_lb_to_field(field_t * phi, lb_t * lb) {
  assert(phi);
  assert(lb);
  CK(b200lb_phi_from_g(b200_));
  return done();
}

int lb_b200_field_halo(field_t * phi) {
  assert(phi);
  CK(b200lb_phi_halo(b200_));
  return done();
}

int lb_b200_field_memcpy(field_t * phi, tdpMemcpyKind flag) {
  assert(phi);
  if (phi->nsites != nsite_) pe_fatal(pe_, "lb_halo_b200: order parameter on %d sites (Lees-Edwards buffers?) against %ld\n", phi->nsites, nsite_);
  if (flag == tdpMemcpyHostToDevice) CK(b200lb_field_upload(b200_, B200LB_FIELD_PHI, phi->data, HOST_LAYOUT));
  if (flag == tdpMemcpyDeviceToHost) CK(b200lb_field_download(b200_, B200LB_FIELD_PHI, phi->data, HOST_LAYOUT));
  return done();
}

/* field_grad_compute: the function-pointer plug point (src/field_grad.h:22) says which
 * stencil the input file asked for (src/gradient_rt.c:81-98) */
int lb_b200_field_grad_compute(field_grad_t * grad) {
  assert(grad);
  if (grad->d2 == grad_3d_7pt_fluid_d2) CK(b200lb_grad7(b200_));
  else if (grad->d2 == grad_3d_27pt_fluid_d2) CK(b200lb_grad27(b200_));
  else pe_fatal(pe_, "lb_halo_b200: fd_gradient_calculation 3d_7pt_fluid or 3d_27pt_fluid\n");
  return done();
}

int lb_b200_field_grad_memcpy(field_grad_t * grad, tdpMemcpyKind flag) {
  assert(grad);
  if (flag == tdpMemcpyDeviceToHost) {
    CK(b200lb_field_download(b200_, B200LB_FIELD_GRAD, grad->grad, HOST_LAYOUT));
    CK(b200lb_field_download(b200_, B200LB_FIELD_DELSQ, grad->delsq, HOST_LAYOUT));
  }
  return done();
}

/* flat walls: the library folds the bounce-back into the propagation that follows wall_bbl
 * and tallies the walls' momentum on the device */
int lb_b200_wall_commit(wall_t * wall) {
  assert(wall);
  if (wall->param->isporousmedia) pe_fatal(pe_, "lb_halo_b200: porous media are not on the device path\n");
  if (wall->param->iswall == 0) return 0;
  CK(b200lb_wall_commit(b200_, wall->param->isboundary, wall->param->ubot, wall->param->utop));
  if (wall->param->slip.active) {
    /* the faces of the slip table (src/wall.h:25-28); the library derives edges (the mean
     * of the two walls) and corners (none) the way wall_init_boundaries_slip lays them out */
    const double * s = wall->param->slip.s;
    const double sbot[3] = {s[WALL_SLIP_XBOT], s[WALL_SLIP_YBOT], s[WALL_SLIP_ZBOT]};
    const double stop[3] = {s[WALL_SLIP_XTOP], s[WALL_SLIP_YTOP], s[WALL_SLIP_ZTOP]};
    CK(b200lb_wall_slip_set(b200_, sbot, stop, 1));
  }
  return 0;
}

int lb_b200_wall_bbl(wall_t * wall) {
  assert(wall);
  CK(b200lb_wall_bbl(b200_));
  return done();
}

int lb_b200_wall_momentum(wall_t * wall, double f[3]) {
  double fdev[3] = {0.0, 0.0, 0.0};
  assert(wall);
  CK(b200lb_wall_momentum(b200_, fdev));
  /* wall->fnet holds what host code added through wall_momentum_add (src/wall.c:1279) */
  for (int ia = 0; ia < 3; ia++) f[ia] = wall->fnet[ia] + fdev[ia];
  return 0;
}

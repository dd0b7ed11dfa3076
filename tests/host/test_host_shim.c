/*
 * test_host_shim.c -- C test of the host-side mirror (ludwig_b200/host/lb_b200.h),
 * written the way the reference's own unit tests are (tests/unit/test_prop.c:81-262,
 * tests/unit/test_model.c): real objects on a periodic box, no mocks.
 *
 *   1. do_test_velocity            every site keeps its own p after halo + propagation
 *   2. do_test_source_destination  f = global site id lands at r + c_p (periodic)
 *   3. the src/ludwig.c:530-862 loop through lb_collide / lb_halo / lb_propagation
 *      against the CPU oracle (oracle/lb_oracle.h), single fluid and binary, 1e-12
 *
 * Build and run: see tests/test_host_shim.py. Exit code 0 = pass.
 */

#include <float.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../ludwig_b200/host/lb_b200.h"
#include "../../oracle/lb_oracle.h"

static int nfail = 0;
#define CHECK(cond, msg) do { if (!(cond)) { printf("FAIL %s:%d %s\n", __FILE__, __LINE__, msg); nfail++; } } while (0)
#define OK(call) do { if ((call) != 0) { printf("ERROR %s:%d %s -> %s\n", __FILE__, __LINE__, #call, lb_b200_error()); exit(2); } } while (0)

static void cs_serial(cs_t * cs, int nx, int ny, int nz) {
  memset(cs, 0, sizeof(cs_t));
  cs->ntotal[0] = cs->nlocal[0] = nx;
  cs->ntotal[1] = cs->nlocal[1] = ny;
  cs->ntotal[2] = cs->nlocal[2] = nz;
  cs->nhalo = 1;
  for (int a = 0; a < 3; a++) { cs->mpi_cartsz[a] = 1; cs->periodic[a] = 1; }
  cs->nranks = 1;
}

static long cs_index(const cs_t * cs, int ic, int jc, int kc) {   /* src/coords.c:617-631, nhalo = 1 */
  return ((long) ic*(cs->nlocal[1] + 2) + jc)*(cs->nlocal[2] + 2) + kc;
}

#define LB_ADDR(nsite, ndist, nvel, index, n, p) (((long) (index)*(ndist) + (n))*(nvel) + (p))

static void test_prop(int ndist, lb_halo_enum_t halo) {
  cs_t cs;
  physics_t phys = {1.0, 0.1, 0.1, {0.0, 0.0, 0.0}, 0.0};
  lb_data_options_t opts = lb_data_options_default();
  lb_t * lb = NULL;
  orc_model_t model;
  const int nx = 8, ny = 6, nz = 10;

  cs_serial(&cs, nx, ny, nz);
  opts.ndist = ndist; opts.halo = halo;
  OK(lb_data_create(&cs, &phys, &opts, &lb));
  orc_model_init(lb->nvel, &model);

  /* do_test_velocity */
  for (int ic = 1; ic <= nx; ic++) for (int jc = 1; jc <= ny; jc++) for (int kc = 1; kc <= nz; kc++) {
    long index = cs_index(&cs, ic, jc, kc);
    for (int n = 0; n < ndist; n++) for (int p = 0; p < lb->nvel; p++)
      lb->f[LB_ADDR(lb->nsite, ndist, lb->nvel, index, n, p)] = 1.0*(p + n*lb->nvel);
  }
  OK(lb_memcpy(lb, tdpMemcpyHostToDevice));
  OK(lb_halo(lb));
  OK(lb_propagation(lb));
  OK(lb_memcpy(lb, tdpMemcpyDeviceToHost));
  for (int ic = 1; ic <= nx; ic++) for (int jc = 1; jc <= ny; jc++) for (int kc = 1; kc <= nz; kc++) {
    long index = cs_index(&cs, ic, jc, kc);
    for (int n = 0; n < ndist; n++) for (int p = 0; p < lb->nvel; p++)
      CHECK(lb->f[LB_ADDR(lb->nsite, ndist, lb->nvel, index, n, p)] == 1.0*(p + n*lb->nvel), "velocity mixing");
  }

  /* do_test_source_destination */
  for (int ic = 1; ic <= nx; ic++) for (int jc = 1; jc <= ny; jc++) for (int kc = 1; kc <= nz; kc++) {
    long index = cs_index(&cs, ic, jc, kc);
    double id = (double) (((ic - 1)*ny + (jc - 1))*nz + (kc - 1));
    for (int n = 0; n < ndist; n++) for (int p = 0; p < lb->nvel; p++)
      lb->f[LB_ADDR(lb->nsite, ndist, lb->nvel, index, n, p)] = id;
  }
  OK(lb_memcpy(lb, tdpMemcpyHostToDevice));
  OK(lb_halo(lb));
  OK(lb_propagation(lb));
  OK(lb_memcpy(lb, tdpMemcpyDeviceToHost));
  for (int ic = 1; ic <= nx; ic++) for (int jc = 1; jc <= ny; jc++) for (int kc = 1; kc <= nz; kc++) {
    long index = cs_index(&cs, ic, jc, kc);
    for (int n = 0; n < ndist; n++) for (int p = 0; p < lb->nvel; p++) {
      int is = (ic - 1 - model.cv[p][0] + nx) % nx;
      int js = (jc - 1 - model.cv[p][1] + ny) % ny;
      int ks = (kc - 1 - model.cv[p][2] + nz) % nz;
      double id = (double) ((is*ny + js)*nz + ks);
      CHECK(lb->f[LB_ADDR(lb->nsite, ndist, lb->nvel, index, n, p)] == id, "source-destination");
    }
  }
  lb_free(lb);
  printf("test_prop ndist %d halo %d done\n", ndist, (int) halo);
}

static double urand(unsigned long * s) {
  *s = *s*6364136223846793005UL + 1442695040888963407UL;
  return (double) ((*s >> 11) & 0xFFFFFFFFFFFFFUL)/4503599627370496.0 - 0.5;
}

static void test_loop(int ndist, int nsteps, int npt, double kt) {
  cs_t cs;
  physics_t phys = {1.0, 0.1, 0.1, {0.0, 0.0, 0.0}, 0.0, 0.0};
  lb_data_options_t opts = lb_data_options_default();
  lb_t * lb = NULL; hydro_t * hydro = NULL;
  noise_t * noise = NULL;
  unsigned int * ostate = NULL;
  field_t * phi = NULL; field_grad_t * dphi = NULL; fe_symm_t * fe = NULL;
  orc_model_t model; orc_box_t box; orc_param_t prm;
  const int nl[3] = {10, 8, 6};
  const double zero[3] = {0.0, 0.0, 0.0};
  unsigned long seed = 8361235;

  cs_serial(&cs, nl[0], nl[1], nl[2]);
  opts.ndist = ndist;
  if (ndist == 2) { phys.eta_shear = phys.eta_bulk = 0.00625; phys.mobility = 3.75; }
  OK(lb_data_create(&cs, &phys, &opts, &lb));
  OK(hydro_create(lb, &hydro));
  /* ludwig.c:242-243 creates the generator always; collision_rt.c:55 switches it on */
  OK(noise_create(lb, &noise));
  OK(noise_init(noise, 0));
  if (kt > 0.0) {
    phys.kt = kt;
    OK(noise_present_set(noise, NOISE_RHO, 1));
  }
  if (ndist == 2) {
    fe_symm_param_t fp = {-0.00625, 0.00625, 0.004, 0.0, 0.0};
    OK(field_create(lb, "phi", &phi));
    OK(field_grad_create(phi, &dphi));
    /* the plug point gradient_rt.c uses (src/gradient_rt.c:81-98) */
    if (npt == 27) OK(field_grad_set(dphi, grad_3d_27pt_fluid_d2, NULL));
    OK(fe_symm_create(phi, dphi, &fe));
    OK(fe_symm_param_set(fe, fp));
  }
  orc_model_init(lb->nvel, &model);
  orc_box_init(nl, &box);
  memset(&prm, 0, sizeof(prm));
  prm.rho0 = phys.rho0; prm.eta_shear = phys.eta_shear; prm.eta_bulk = phys.eta_bulk;
  prm.mobility = phys.mobility; prm.a = -0.00625; prm.b = 0.00625; prm.kappa = 0.004;

  size_t nf = (size_t) lb->nsite*ndist*lb->nvel;
  double * f = (double *) calloc(nf, sizeof(double));
  double * fprime = (double *) calloc(nf, sizeof(double));
  double * rho = (double *) calloc(lb->nsite, sizeof(double));
  double * u = (double *) calloc(3*(size_t) lb->nsite, sizeof(double));
  double * force = (double *) calloc(3*(size_t) lb->nsite, sizeof(double));
  double * ph = (double *) calloc(lb->nsite, sizeof(double));
  double * gr = (double *) calloc(3*(size_t) lb->nsite, sizeof(double));
  double * dsq = (double *) calloc(lb->nsite, sizeof(double));

  for (int ic = 1; ic <= nl[0]; ic++) for (int jc = 1; jc <= nl[1]; jc++) for (int kc = 1; kc <= nl[2]; kc++) {
    long index = cs_index(&cs, ic, jc, kc);
    double r = 1.0 + 0.02*urand(&seed);
    double uu[3] = {0.02*urand(&seed), 0.02*urand(&seed), 0.02*urand(&seed)};
    orc_equilib_set(&model, ndist, f, index, r, uu);
    if (ndist == 2) {
      double p0 = urand(&seed);
      for (int p = 0; p < lb->nvel; p++) f[LB_ADDR(lb->nsite, 2, lb->nvel, index, 1, p)] = (p == 0) ? p0 : 0.001*model.wv[p]*urand(&seed);
    }
  }
  memcpy(lb->f, f, nf*sizeof(double));
  OK(lb_memcpy(lb, tdpMemcpyHostToDevice));

  /* the loop of src/ludwig.c:530-862 through the mirrored call surface */
  for (int n = 0; n < nsteps; n++) {
    OK(hydro_f_zero(hydro, zero));
    if (ndist == 2) {
      OK(phi_lb_to_field(phi, lb));
      OK(field_halo(phi));
      OK(field_grad_compute(dphi));
    }
    OK(hydro_u_zero(hydro, zero));
    OK(lb_collide(lb, hydro, NULL, noise, (fe_t *) fe, NULL));
    OK(lb_halo(lb));
    OK(lb_propagation(lb));
  }
  OK(lb_memcpy(lb, tdpMemcpyDeviceToHost));
  OK(hydro_memcpy(hydro, tdpMemcpyDeviceToHost));
  OK(noise_memcpy(noise, tdpMemcpyDeviceToHost));

  if (kt > 0.0) {
    const int noffset[3] = {0, 0, 0};
    ostate = (unsigned int *) calloc(4*(size_t) lb->nsite, sizeof(unsigned int));
    orc_noise_init(&box, 0, noffset, nl, ostate);
    orc_noise_attach(ostate, kt);
  }

  orc_gradient_select(npt);
  orc_step(&model, &box, &prm, ndist, nsteps, f, fprime, rho, u, force, ph, gr, dsq);
  orc_gradient_select(7);
  orc_noise_attach(NULL, 0.0);
  if (kt > 0.0) {
    int nbad = 0;
    for (int ic = 1; ic <= nl[0]; ic++) for (int jc = 1; jc <= nl[1]; jc++) for (int kc = 1; kc <= nl[2]; kc++) {
      long index = cs_index(&cs, ic, jc, kc);
      for (int ia = 0; ia < 4; ia++) nbad += (noise->state[4*index + ia] != ostate[4*index + ia]);
    }
    CHECK(nbad == 0, "generator state of interior sites, bit for bit");
    free(ostate);
  }

  double emax = 0.0, fmax = 0.0, umax = 0.0, mass0 = 0.0, mass1 = 0.0;
  for (int ic = 1; ic <= nl[0]; ic++) for (int jc = 1; jc <= nl[1]; jc++) for (int kc = 1; kc <= nl[2]; kc++) {
    long index = cs_index(&cs, ic, jc, kc);
    for (int n = 0; n < ndist; n++) for (int p = 0; p < lb->nvel; p++) {
      long i = LB_ADDR(lb->nsite, ndist, lb->nvel, index, n, p);
      if (fabs(lb->f[i] - f[i]) > emax) emax = fabs(lb->f[i] - f[i]);
      if (fabs(f[i]) > fmax) fmax = fabs(f[i]);
      if (n == 0) { mass0 += f[i]; mass1 += lb->f[i]; }
    }
    for (int a = 0; a < 3; a++) {
      double d = fabs(hydro->u.data[3*index + a] - u[3*index + a]);
      if (d > umax) umax = d;
    }
  }
  printf("test_loop ndist %d (%d-pt gradient, kT %.3e): %d steps, max|df|/max|f| = %.3e, max|du| = %.3e, mass diff %.3e\n",
	 ndist, npt, kt, nsteps, emax/fmax, umax, fabs(mass1 - mass0));
  CHECK(emax/fmax < 1.0e-12, "time loop vs oracle");
  CHECK(umax < 1.0e-12, "hydro u vs oracle");
  CHECK(fabs(mass1 - mass0) < 1.0e-10, "mass");
  {
    double sums[8];
    OK(stats_distribution_sums(lb, sums));
    CHECK(fabs(sums[0] - mass1) < 1.0e-9, "device mass sum");
  }
  {
    /* stats_velocity_minmax as a device reduction against the host copy of u, and the
     * binary file records of lb_io_write against lb->f (src/model.c:1498-1525) */
    stats_vel_t sv;
    io_aggregator_t aggr;
    double hmin[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hmax[3] = {FLT_MIN, FLT_MIN, FLT_MIN};
    int nbad = 0;
    long ib = 0;
    OK(stats_velocity_minmax(&sv, hydro, NULL));
    aggr.szelement = (size_t) ndist*lb->nvel*sizeof(double);
    aggr.szbuf = aggr.szelement*nl[0]*nl[1]*nl[2];
    aggr.buf = (char *) malloc(aggr.szbuf);
    OK(lb_io_aggr_pack(lb, &aggr));
    for (int ic = 1; ic <= nl[0]; ic++) for (int jc = 1; jc <= nl[1]; jc++) for (int kc = 1; kc <= nl[2]; kc++) {
      long index = cs_index(&cs, ic, jc, kc);
      const double * rec = (const double *) (aggr.buf + ib*aggr.szelement);
      for (int a = 0; a < 3; a++) {
	double ua = hydro->u.data[3*index + a];
	if (ua < hmin[a]) hmin[a] = ua;
	if (ua > hmax[a]) hmax[a] = ua;
      }
      for (int n = 0; n < ndist; n++) for (int p = 0; p < lb->nvel; p++) {
	nbad += (rec[n*lb->nvel + p] != lb->f[LB_ADDR(lb->nsite, ndist, lb->nvel, index, n, p)]);
      }
      ib++;
    }
    for (int a = 0; a < 3; a++) {
      CHECK(sv.umin[a] == hmin[a] && sv.umax[a] == hmax[a], "stats_velocity_minmax on the device");
    }
    CHECK(nbad == 0, "lb_io_aggr_pack records");
    free(aggr.buf);
  }

  free(f); free(fprime); free(rho); free(u); free(force); free(ph); free(gr); free(dsq);
  if (fe) fe_symm_free(fe);
  if (dphi) field_grad_free(dphi);
  if (phi) field_free(phi);
  noise_free(noise);
  hydro_free(hydro);
  lb_free(lb);
}

/* The loop of src/ludwig.c:795-860 with flat walls: collide, halo, wall_bbl, propagate --
 * a body-force driven duct flow between y and z walls, against the oracle. */
static void test_walls(int nsteps) {
  cs_t cs;
  physics_t phys = {1.0, 0.1666, 0.1666, {1.0e-4, 0.0, 0.0}, 0.0, 0.0};
  lb_data_options_t opts = lb_data_options_default();
  lb_t * lb = NULL; hydro_t * hydro = NULL; wall_t * wall = NULL;
  wall_param_t wp;
  orc_model_t model; orc_box_t box; orc_param_t prm; orc_wall_t owall;
  const int nl[3] = {6, 9, 34};
  const double zero[3] = {0.0, 0.0, 0.0};
  unsigned long seed = 20200424;
  double gw[3];

  cs_serial(&cs, nl[0], nl[1], nl[2]);
  cs.periodic[1] = 0; cs.periodic[2] = 0;
  for (int a = 1; a < 3; a++) { cs.mpi_cart_neighbours[0][a] = -1; cs.mpi_cart_neighbours[1][a] = -1; }
  OK(lb_data_create(&cs, &phys, &opts, &lb));
  OK(hydro_create(lb, &hydro));
  memset(&wp, 0, sizeof(wp));
  wp.isboundary[1] = 1; wp.isboundary[2] = 1;
  OK(wall_create(lb, &wall));
  OK(wall_commit(wall, &wp));

  orc_model_init(lb->nvel, &model);
  orc_box_init(nl, &box);
  memset(&prm, 0, sizeof(prm));
  prm.rho0 = phys.rho0; prm.eta_shear = phys.eta_shear; prm.eta_bulk = phys.eta_bulk;
  for (int a = 0; a < 3; a++) prm.force_global[a] = phys.fbody[a];
  memset(&owall, 0, sizeof(owall));
  owall.isboundary[1] = 1; owall.isboundary[2] = 1;

  size_t nf = (size_t) lb->nsite*lb->nvel;
  double * f = (double *) calloc(nf, sizeof(double));
  double * fprime = (double *) calloc(nf, sizeof(double));
  double * rho = (double *) calloc(lb->nsite, sizeof(double));
  double * u = (double *) calloc(3*(size_t) lb->nsite, sizeof(double));
  double * force = (double *) calloc(3*(size_t) lb->nsite, sizeof(double));
  for (int ic = 1; ic <= nl[0]; ic++) for (int jc = 1; jc <= nl[1]; jc++) for (int kc = 1; kc <= nl[2]; kc++) {
    long index = cs_index(&cs, ic, jc, kc);
    double r = 1.0 + 0.02*urand(&seed);
    double uu[3] = {0.02*urand(&seed), 0.02*urand(&seed), 0.02*urand(&seed)};
    orc_equilib_set(&model, 1, f, index, r, uu);
  }
  memcpy(lb->f, f, nf*sizeof(double));
  OK(lb_memcpy(lb, tdpMemcpyHostToDevice));

  for (int n = 0; n < nsteps; n++) {
    OK(hydro_f_zero(hydro, zero));
    OK(hydro_u_zero(hydro, zero));
    OK(lb_collide(lb, hydro, NULL, NULL, NULL, NULL));
    OK(lb_halo(lb));
    OK(wall_set_wall_distributions(wall));
    OK(wall_bbl(wall));
    OK(lb_propagation(lb));
  }
  OK(lb_memcpy(lb, tdpMemcpyDeviceToHost));
  OK(hydro_memcpy(hydro, tdpMemcpyDeviceToHost));
  OK(wall_momentum(wall, gw));

  orc_wall_attach(&owall);
  orc_step(&model, &box, &prm, 1, nsteps, f, fprime, rho, u, force, NULL, NULL, NULL);
  orc_wall_attach(NULL);

  double emax = 0.0, fmax = 0.0, umax = 0.0;
  for (int ic = 1; ic <= nl[0]; ic++) for (int jc = 1; jc <= nl[1]; jc++) for (int kc = 1; kc <= nl[2]; kc++) {
    long index = cs_index(&cs, ic, jc, kc);
    for (int p = 0; p < lb->nvel; p++) {
      long i = LB_ADDR(lb->nsite, 1, lb->nvel, index, 0, p);
      if (fabs(lb->f[i] - f[i]) > emax) emax = fabs(lb->f[i] - f[i]);
      if (fabs(f[i]) > fmax) fmax = fabs(f[i]);
    }
    for (int a = 0; a < 3; a++) {
      double d = fabs(hydro->u.data[3*index + a] - u[3*index + a]);
      if (d > umax) umax = d;
    }
  }
  printf("test_walls: %d steps, max|df|/max|f| = %.3e, max|du| = %.3e, wall momentum %.6e %.3e %.3e (oracle %.6e)\n",
	 nsteps, emax/fmax, umax, gw[0], gw[1], gw[2], owall.fnet[0]);
  CHECK(emax/fmax < 1.0e-12, "walls: time loop vs oracle");
  CHECK(umax < 1.0e-12, "walls: hydro u vs oracle");
  for (int a = 0; a < 3; a++) CHECK(fabs(gw[a] - owall.fnet[a]) < 1.0e-10, "wall momentum account");

  free(f); free(fprime); free(rho); free(u); free(force);
  wall_free(wall);
  hydro_free(hydro);
  lb_free(lb);
}

/* Lees-Edwards planes through the reference's call order (src/ludwig.c:552, 800-845) */
static void test_lees_edwards(int nsteps, int oscillatory) {
  cs_t cs;
  physics_t phys = {1.0, 0.1666, 0.1666, {0.0, 0.0, 0.0}, 0.0, 0.0};
  lb_data_options_t opts = lb_data_options_default();
  lb_t * lb = NULL; hydro_t * hydro = NULL; lees_edw_t * le = NULL;
  lees_edw_options_t leopts = {2, LE_SHEAR_TYPE_STEADY, 0, 0, 0.0375};
  orc_model_t model; orc_box_t box; orc_param_t prm; orc_le_t ole;
  const int nl[3] = {12, 10, 7};
  const double zero[3] = {0.0, 0.0, 0.0};
  unsigned long seed = 20101208;

  if (oscillatory) { leopts.type = LE_SHEAR_TYPE_OSCILLATORY; leopts.period = 17; leopts.nt0 = 2; }
  cs_serial(&cs, nl[0], nl[1], nl[2]);
  OK(lb_data_create(&cs, &phys, &opts, &lb));
  OK(hydro_create(lb, &hydro));
  OK(lees_edw_create(lb, &leopts, &le));
  CHECK(lees_edw_nplane_local(le) == 2, "planes per rank");
  CHECK(lees_edw_plane_location(le, 0) == 3 && lees_edw_plane_location(le, 1) == 9, "plane location");

  orc_model_init(lb->nvel, &model);
  orc_box_init(nl, &box);
  memset(&prm, 0, sizeof(prm));
  prm.rho0 = phys.rho0; prm.eta_shear = phys.eta_shear; prm.eta_bulk = phys.eta_bulk;
  memset(&ole, 0, sizeof(ole));
  ole.nplanes = leopts.nplanes; ole.uy = leopts.uy; ole.oscillatory = oscillatory;
  ole.period = leopts.period; ole.nt0 = leopts.nt0;

  size_t nf = (size_t) lb->nsite*lb->nvel;
  double * f = (double *) calloc(nf, sizeof(double));
  double * fprime = (double *) calloc(nf, sizeof(double));
  double * rho = (double *) calloc(lb->nsite, sizeof(double));
  double * u = (double *) calloc(3*(size_t) lb->nsite, sizeof(double));
  double * force = (double *) calloc(3*(size_t) lb->nsite, sizeof(double));
  for (int ic = 1; ic <= nl[0]; ic++) for (int jc = 1; jc <= nl[1]; jc++) for (int kc = 1; kc <= nl[2]; kc++) {
    long index = cs_index(&cs, ic, jc, kc);
    double r = 1.0 + 0.02*urand(&seed);
    double uu[3] = {0.02*urand(&seed), 0.02*urand(&seed), 0.02*urand(&seed)};
    orc_equilib_set(&model, 1, f, index, r, uu);
  }
  memcpy(lb->f, f, nf*sizeof(double));
  OK(lb_memcpy(lb, tdpMemcpyHostToDevice));

  for (int n = 0; n < nsteps; n++) {
    OK(physics_control_next_step(&phys));
    OK(hydro_f_zero(hydro, zero));
    OK(hydro_u_zero(hydro, zero));
    OK(lb_collide(lb, hydro, NULL, NULL, NULL, NULL));
    OK(lb_le_apply_boundary_conditions(lb, le));
    OK(lb_halo(lb));
    OK(lb_propagation(lb));
  }
  OK(lb_memcpy(lb, tdpMemcpyDeviceToHost));
  OK(hydro_memcpy(hydro, tdpMemcpyDeviceToHost));

  orc_le_attach(&ole);
  orc_step(&model, &box, &prm, 1, nsteps, f, fprime, rho, u, force, NULL, NULL, NULL);
  orc_le_attach(NULL);

  double emax = 0.0, fmax = 0.0, umax = 0.0;
  for (int ic = 1; ic <= nl[0]; ic++) for (int jc = 1; jc <= nl[1]; jc++) for (int kc = 1; kc <= nl[2]; kc++) {
    long index = cs_index(&cs, ic, jc, kc);
    for (int p = 0; p < lb->nvel; p++) {
      long i = LB_ADDR(lb->nsite, 1, lb->nvel, index, 0, p);
      if (fabs(lb->f[i] - f[i]) > emax) emax = fabs(lb->f[i] - f[i]);
      if (fabs(f[i]) > fmax) fmax = fabs(f[i]);
    }
    for (int a = 0; a < 3; a++) {
      double d = fabs(hydro->u.data[3*index + a] - u[3*index + a]);
      if (d > umax) umax = d;
    }
  }
  printf("test_lees_edwards (%s): %d steps, max|df|/max|f| = %.3e, max|du| = %.3e\n",
	 oscillatory ? "oscillatory" : "steady", nsteps, emax/fmax, umax);
  CHECK(emax/fmax < 1.0e-12, "Lees-Edwards: time loop vs oracle");
  CHECK(umax < 1.0e-12, "Lees-Edwards: hydro u vs oracle");

  free(f); free(fprime); free(rho); free(u); free(force);
  lees_edw_free(le);
  hydro_free(hydro);
  lb_free(lb);
}

/* The order parameter as a finite-difference field ("free_energy symmetric") through the
 * reference's call order (src/ludwig.c:537-575, 727-768, 791-862) */
static void test_cahn_hilliard(int nsteps, int order, int method, int npt, int with_visc) {
  cs_t cs;
  physics_t phys = {1.0, 0.00625, 0.00625, {0.0, 0.0, 0.0}, 1.25, 0.0};
  lb_data_options_t opts = lb_data_options_default();
  lb_t * lb = NULL; hydro_t * hydro = NULL; field_t * phi = NULL; field_grad_t * grad = NULL;
  fe_symm_t * fe = NULL; phi_ch_t * pch = NULL; pth_t * pth = NULL;
  phi_ch_info_t info = {0};
  fe_symm_param_t fp = {-0.00625, 0.00625, 0.004, 0.0, 0.0};
  visc_t * visc = NULL;
  visc_arrhenius_param_t vp = {0};
  orc_visc_t ovisc = {0.625, 0.00625, 1.0};
  orc_model_t model; orc_box_t box; orc_param_t prm; orc_ch_t och;
  const int nl[3] = {10, 8, 6};
  const double zero[3] = {0.0, 0.0, 0.0};
  unsigned long seed = 20240229;

  if (method == FE_FORCE_METHOD_PHI_GRADMU) { phys.grad_mu[0] = 1.0e-5; phys.grad_mu[1] = -2.0e-5; phys.grad_mu[2] = 3.0e-5; }
  cs_serial(&cs, nl[0], nl[1], nl[2]);
  OK(lb_data_create(&cs, &phys, &opts, &lb));
  OK(hydro_create(lb, &hydro));
  OK(advection_order_set(order));
  OK(phi_ch_create(lb, &info, &pch));
  OK(pth_create(lb, method, &pth));
  OK(field_create(lb, "phi", &phi));
  OK(field_grad_create(phi, &grad));
  OK(field_grad_set(grad, npt == 27 ? grad_3d_27pt_fluid_d2 : grad_3d_7pt_fluid_d2, NULL));
  OK(fe_symm_create(phi, grad, &fe));
  OK(fe_symm_param_set(fe, fp));
  if (with_visc) {                      /* viscosity_model arrhenius, src/ludwig.c:2001-2040 */
    vp.eta_plus = ovisc.eta_plus; vp.eta_minus = ovisc.eta_minus; vp.phistar = ovisc.phistar;
    OK(visc_arrhenius_create(lb, phi, vp, &visc));
  }

  orc_model_init(lb->nvel, &model);
  orc_box_init(nl, &box);
  memset(&prm, 0, sizeof(prm));
  prm.rho0 = phys.rho0; prm.eta_shear = phys.eta_shear; prm.eta_bulk = phys.eta_bulk;
  prm.mobility = phys.mobility; prm.a = fp.a; prm.b = fp.b; prm.kappa = fp.kappa;
  memset(&och, 0, sizeof(och));
  och.order = order; och.method = method;
  for (int a = 0; a < 3; a++) och.grad_mu[a] = phys.grad_mu[a];

  size_t nf = (size_t) lb->nsite*lb->nvel;
  double * f = (double *) calloc(nf, sizeof(double));
  double * fprime = (double *) calloc(nf, sizeof(double));
  double * rho = (double *) calloc(lb->nsite, sizeof(double));
  double * u = (double *) calloc(3*(size_t) lb->nsite, sizeof(double));
  double * force = (double *) calloc(3*(size_t) lb->nsite, sizeof(double));
  double * ophi = (double *) calloc(lb->nsite, sizeof(double));
  double * ograd = (double *) calloc(3*(size_t) lb->nsite, sizeof(double));
  double * odelsq = (double *) calloc(lb->nsite, sizeof(double));
  for (int ic = 1; ic <= nl[0]; ic++) for (int jc = 1; jc <= nl[1]; jc++) for (int kc = 1; kc <= nl[2]; kc++) {
    long index = cs_index(&cs, ic, jc, kc);
    double r = 1.0 + 0.01*urand(&seed);
    double uu[3] = {0.01*urand(&seed), 0.01*urand(&seed), 0.01*urand(&seed)};
    orc_equilib_set(&model, 1, f, index, r, uu);
    ophi[index] = 0.4*urand(&seed);
  }
  memcpy(lb->f, f, nf*sizeof(double));
  memcpy(phi->data, ophi, lb->nsite*sizeof(double));
  OK(lb_memcpy(lb, tdpMemcpyHostToDevice));
  OK(field_memcpy(phi, tdpMemcpyHostToDevice));

  for (int n = 0; n < nsteps; n++) {
    OK(hydro_f_zero(hydro, zero));
    OK(field_halo(phi));
    OK(field_grad_compute(grad));
    OK(phi_force_calculation(pth, (fe_t *) fe, phi, hydro));
    OK(phi_cahn_hilliard(pch, (fe_t *) fe, phi, hydro, NULL, NULL));
    OK(hydro_u_zero(hydro, zero));
    if (visc) OK(visc_arrhenius_update(visc, hydro));         /* src/ludwig.c:794 */
    OK(lb_collide(lb, hydro, NULL, NULL, (fe_t *) fe, visc));
    OK(lb_halo(lb));
    OK(lb_propagation(lb));
  }
  OK(lb_memcpy(lb, tdpMemcpyDeviceToHost));
  OK(hydro_memcpy(hydro, tdpMemcpyDeviceToHost));
  OK(field_memcpy(phi, tdpMemcpyDeviceToHost));

  double * oeta = (double *) calloc(lb->nsite, sizeof(double));
  orc_gradient_select(npt);
  orc_ch_attach(&och);
  if (with_visc) orc_visc_attach(&ovisc, oeta);
  orc_step(&model, &box, &prm, 1, nsteps, f, fprime, rho, u, force, ophi, ograd, odelsq);
  orc_visc_attach(NULL, NULL);
  free(oeta);
  orc_ch_attach(NULL);
  orc_gradient_select(7);

  double emax = 0.0, fmax = 0.0, umax = 0.0, pmax = 0.0;
  for (int ic = 1; ic <= nl[0]; ic++) for (int jc = 1; jc <= nl[1]; jc++) for (int kc = 1; kc <= nl[2]; kc++) {
    long index = cs_index(&cs, ic, jc, kc);
    for (int p = 0; p < lb->nvel; p++) {
      long i = LB_ADDR(lb->nsite, 1, lb->nvel, index, 0, p);
      if (fabs(lb->f[i] - f[i]) > emax) emax = fabs(lb->f[i] - f[i]);
      if (fabs(f[i]) > fmax) fmax = fabs(f[i]);
    }
    for (int a = 0; a < 3; a++) {
      double d = fabs(hydro->u.data[3*index + a] - u[3*index + a]);
      if (d > umax) umax = d;
    }
    if (fabs(phi->data[index] - ophi[index]) > pmax) pmax = fabs(phi->data[index] - ophi[index]);
  }
  printf("test_cahn_hilliard (order %d, method %d, %d-point gradient%s): %d steps, max|df|/max|f| = %.3e, "
	 "max|du| = %.3e, max|dphi| = %.3e\n", order, method, npt, with_visc ? ", viscosity model" : "",
	 nsteps, emax/fmax, umax, pmax);
  CHECK(emax/fmax < 1.0e-12, "Cahn-Hilliard: distributions vs oracle");
  CHECK(umax < 1.0e-12, "Cahn-Hilliard: hydro u vs oracle");
  CHECK(pmax < 1.0e-13, "Cahn-Hilliard: order parameter vs oracle");

  free(f); free(fprime); free(rho); free(u); free(force); free(ophi); free(ograd); free(odelsq);
  if (visc) visc_arrhenius_free(visc);
  fe_symm_free(fe);
  field_grad_free(grad);
  field_free(phi);
  pth_free(pth);
  phi_ch_free(pch);
  hydro_free(hydro);
  lb_free(lb);
}

int main(int argc, char ** argv) {
  if (argc > 1 && strcmp(argv[1], "fd") == 0) {
    /* the finite-difference order parameter on its own (tests/test_host_shim.py) */
    test_cahn_hilliard(6, 1, FE_FORCE_METHOD_STRESS_DIVERGENCE, 27, 0);
    test_cahn_hilliard(6, 2, FE_FORCE_METHOD_STRESS_DIVERGENCE, 7, 0);
    test_cahn_hilliard(6, 3, FE_FORCE_METHOD_PHI_GRADMU, 27, 0);
    test_cahn_hilliard(6, 2, FE_FORCE_METHOD_STRESS_DIVERGENCE, 27, 1);
    if (nfail) { printf("%d check(s) FAILED\n", nfail); return 1; }
    printf("host shim (fd): all passed\n");
    return 0;
  }
  test_prop(1, LB_HALO_TARGET);
  test_prop(2, LB_HALO_TARGET);
  test_prop(1, LB_HALO_OPENMP_REDUCED);
  test_loop(1, 7, 7, 0.0);
  test_loop(2, 7, 7, 0.0);
  test_loop(2, 7, 27, 0.0);
  test_loop(1, 7, 7, 2.13333e-05);          /* isothermal fluctuations */
  test_loop(2, 7, 7, 2.13333e-05);
  test_walls(9);
  test_lees_edwards(9, 0);
  test_lees_edwards(23, 1);
  if (nfail) { printf("%d check(s) FAILED\n", nfail); return 1; }
  printf("host shim: all passed\n");
  return 0;
}

/* lb_b200_glue.h -- the forwarding layer a Ludwig build links in to run its lattice
 * Boltzmann hot path on libb200lb.so (include/b200lb.h). Added to src/ of a scratch copy
 * of the reference by tools/make_ludwig_b200.py; the few lines that call it from the
 * reference's own functions are listed in INTEGRATION.md.
 *
 * One handle per process (Ludwig has one lb_t, one hydro_t and one order parameter per
 * rank; its own physics_ref() is a singleton too). Every function returns 0; a library
 * error ends the run through pe_fatal, as every failure on this path does in the
 * reference (src/pe.c:226-243). */

#ifndef LUDWIG_LB_B200_GLUE_H
#define LUDWIG_LB_B200_GLUE_H

#include "pe.h"
#include "coords.h"
#include "lb_data.h"
#include "hydro.h"
#include "field.h"
#include "field_grad.h"
#include "free_energy.h"
#include "noise.h"
#include "visc.h"
#include "wall.h"

int lb_b200_active(void);                 /* lb_halo_scheme lb_halo_b200 was selected */

int lb_b200_create(pe_t * pe, cs_t * cs, int ndim, int nvel, int ndist);  /* src/model.c:56 lb_data_create */
int lb_b200_free(void);                                                   /* src/model.c:186 lb_free */
int lb_b200_memcpy(lb_t * lb, tdpMemcpyKind flag);                        /* src/model.c:228 lb_memcpy */
int lb_b200_halo(lb_t * lb);                                              /* src/model.c:570 lb_halo_swap */
int lb_b200_collide(lb_t * lb, hydro_t * hydro, noise_t * noise, fe_t * fe, visc_t * visc);
                                                                          /* src/collision.c:144 lb_collide */
int lb_b200_propagation(lb_t * lb);                                       /* src/propagation.c:43 */

int lb_b200_wall_commit(wall_t * wall);                                   /* src/wall.c:168 wall_commit */
int lb_b200_wall_bbl(wall_t * wall);                                      /* src/wall.c:960 wall_bbl */
int lb_b200_wall_momentum(wall_t * wall, double f[3]);                    /* src/wall.c:1299 wall_momentum */

int lb_b200_hydro_f_zero(const double fzero[3]);                          /* src/hydro.c:306 */
int lb_b200_hydro_u_zero(const double uzero[3]);                          /* src/hydro.c:279 */
int lb_b200_hydro_memcpy(hydro_t * hydro, tdpMemcpyKind flag);            /* src/hydro.c:154 */

int lb_b200_is_phi(const field_t * field);     /* the order parameter of the binary fluid */
int lb_b200_phi_lb_to_field(field_t * phi, lb_t * lb);                    /* src/phi_lb_coupler.c:39 */
int lb_b200_field_halo(field_t * phi);                                    /* src/field.c:434 */
int lb_b200_field_memcpy(field_t * phi, tdpMemcpyKind flag);              /* src/field.c:261 */
int lb_b200_field_grad_compute(field_grad_t * grad);                      /* src/field_grad.c:305 */
int lb_b200_field_grad_memcpy(field_grad_t * grad, tdpMemcpyKind flag);   /* src/field_grad.c:154 */

#endif

// ch_kernels.cuh -- launchers of the finite-difference order-parameter kernel
// (ch_kernels.cu, compiled twice: ch_launch_exact with -fmad=false and the reference's
// order of operations -- bit parity --, ch_launch_fast with fused multiply-add).

#pragma once

#include <cuda_runtime.h>

#include "layout.cuh"

// fe_force_method_enum_t, src/fe_force_method.h:20-28
enum { CH_FORCE_NONE = 1, CH_FORCE_STRESS_DIVERGENCE = 2, CH_FORCE_PHI_GRADMU = 3 };

struct ChParams {
  Layout L;
  int wrap[3];            // device axis is periodic within this rank: neighbours wrap
  const double * phi;     // [nsd]    order parameter, halos of decomposed directions current
  const double * grad;    // [3][nsd] its gradient, likewise
  const double * delsq;   // [nsd]
  const double * u;       // [3][nsd] velocity of the last collision, likewise
  double * force;         // [3][nsd] body force on the fluid (written or added to)
  double * phi_out;       // [nsd]    updated order parameter (interior)
  double A, B, kappa;     // symmetric free energy
  double mobility;
  double gmu[3];          // external chemical potential gradient
  double fconst[3];       // value the force field holds everywhere when !accumulate
  int accumulate;         // 1: force += , 0: force = fconst +
  int order;              // advection scheme: 1, 2, 3
  int method;             // CH_FORCE_*
  int nx;                 // x-planes marched by one block
  int skipz;              // n_z = 1: the z fluxes stay out of the update (wz = 0, src/phi_cahn_hilliard.c:999)
};

// do_force: phi_force_calculation; do_update: phi_cahn_hilliard. One pass does both.
cudaError_t ch_launch_exact(const ChParams & P, bool do_force, bool do_update, cudaStream_t st);
cudaError_t ch_launch_fast(const ChParams & P, bool do_force, bool do_update, cudaStream_t st);

// ---------------------------------------------------------------------------
// ... across Lees-Edwards planes. ch_kernel runs over the whole lattice as if there were
// none; the layers of sites around every plane (loc - 1 .. loc + 2) are then redone with the
// buffer planes of le_field_kernels.cuh: pass A materialises the fluxes through the plane
// as either side sees them (they differ: src/phi_cahn_hilliard.c:604-616), a reduction forms
// the plane average of the stress-flux mismatch (phi_force_flux_fix_local,
// src/phi_force.c:595-671), pass B applies both corrections and rewrites force and phi.
// With planes the force is the divergence of face-interpolated stresses whatever the
// method (src/phi_force.c:89-96).
// ---------------------------------------------------------------------------

#define CH_LE_MAXPLANE 16

struct ChLeParams {
  int nplanes;
  int loc[CH_LE_MAXPLANE];    // halo-inclusive x of the layer of sites below plane p
  const double * bphi;        // [nplanes][4][s0] buffer planes of phi
  const double * bgrad;       // [nplanes][2][3][s0] gradient in buffers b = 1, 2
  const double * bdelsq;      // [nplanes][2][s0]
  const double * bux;         // [nplanes][2][s0] u_x in buffers b = 1, 2 (time t + 1)
  double * scr;               // [nplanes][8][s0]: fe(loc), fw(loc + 1), stress fluxe[3](loc), fluxw[3](loc + 1)
  double * fbar;              // [nplanes][3] sum over the plane of -fluxe(loc) + fluxw(loc + 1)
  int jdy_up, jdy_dn;         // phi_ch_le_fix_fluxes: displacement +t u_y (looking up), -t u_y
  double fr_up, fr_dn;
  double ra;                  // 1/2 / (L_y L_z)
};

cudaError_t ch_le_launch_exact(const ChParams & P, const ChLeParams & Q, bool do_force, bool do_update, cudaStream_t st);
cudaError_t ch_le_launch_fast(const ChParams & P, const ChLeParams & Q, bool do_force, bool do_update, cudaStream_t st);

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
};

// do_force: phi_force_calculation; do_update: phi_cahn_hilliard. One pass does both.
cudaError_t ch_launch_exact(const ChParams & P, bool do_force, bool do_update, cudaStream_t st);
cudaError_t ch_launch_fast(const ChParams & P, bool do_force, bool do_update, cudaStream_t st);

// model_ops.cuh -- per-velocity-set launchers, instantiated once per model in
// model_d3q19.cu / model_d3q27.cu / model_d2q9.cu so the three compile in parallel.

#pragma once

#include <cuda_runtime.h>

#include "layout.cuh"

struct SweepArgs {
  KParams P;
  bool pull, collide, binary, ghost;
  bool noise;         // isothermal fluctuations on (collide only)
  bool visc;          // viscosity model: per-site relaxation times from P.eta (collide, single fluid)
  bool gmom;          // binary fluid, pull + collide: first moment of g from P.jphi, g not read
  bool wall;          // bounce-back on links folded into the pull (pull only)
  bool diag;          // accumulate conserved sums in the sweep (collide only)
  bool pdl;           // peer-memory sweeps: launch with programmatic stream serialisation (the kernel
                      // before it in the stream is p2p_sync_kernel)
  int aa;             // single-lattice streaming: 0 off, 1 even step, 2 odd step, 3 un-swap and stream
  long long nblocks;  // out: blocks launched (= diagnostic partials written)
  int vec;            // sites per thread (1 or 2)
  int bx;             // threads per block along the contiguous axis
  int bt;             // threads per block (0: 256, or 128 for the kernels with many live values)
  cudaStream_t stream;
};

struct ModelOps {
  int nvel, ndim, nhydro;
  int dax[3];                                   // device axis -> physical axis
  int cv[27][3];
  double wv[27];
  double ghost_norm[B200LB_MAXGHOST];           // sum_p w_p ma[k][p]^2 per ghost mode
  cudaError_t (*sweep)(SweepArgs &);
  cudaError_t (*phi_from_g)(const Layout &, const int wrap[3], const double * g,
			    double * phi, double * jphi, bool pull, cudaStream_t);
  cudaError_t (*stats)(const Layout &, const double * f, int ndist,
		       double * partial, dim3 grid, dim3 block, cudaStream_t);
  cudaError_t (*hydro)(const Layout &, const double * f, int ndist, const double * force,
		       const double fg[3], double * rho, double * u, int mode, const int wrap[3],
		       cudaStream_t);
  cudaError_t (*ghost_upload)(void);
};

const ModelOps * model_ops_d3q19(void);
const ModelOps * model_ops_d3q27(void);
const ModelOps * model_ops_d2q9(void);

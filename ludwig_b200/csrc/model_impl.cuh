// model_impl.cuh -- body shared by the three model translation units.
// Define B200LB_MODEL (D3Q19 ...) and B200LB_OPS_FN before
// including. Binary-fluid kernels are built for 3-d models only
// (src/collision.c:618 asserts NDIM == 3 for lb_collision_binary).

#include <cstring>
#include "model_ops.cuh"
#include "model_tables.h"

static __constant__ GhostTab c_ghost;

#include "site_kernels.cuh"
#include "aux_kernels.cuh"

typedef B200LB_MODEL MODEL;

template <int VEC, bool PULL, bool COLLIDE, bool BINARY, int GM, bool WALL = false, int AA = 0>
static cudaError_t launch_one(SweepArgs & a) {
  // walls are folded into the pull of the single-fluid kernels, and into the pure
  // propagation of the binary fluid (whose collision then runs on the streamed lattice)
  if constexpr (!WALL && PULL && VEC == 1 && (!BINARY || !COLLIDE)) {
    if (a.wall) return launch_one<VEC, PULL, COLLIDE, BINARY, GM, true>(a);
  }
  else if constexpr (!WALL) {
    if (a.wall) return cudaErrorInvalidValue;
  }
  constexpr bool GHOST = ((GM & 3) != 0);
  const KParams & P = a.P;
  const int n2 = P.hi[2] - P.lo[2], n1 = P.hi[1] - P.lo[1], n0 = P.hi[0] - P.lo[0];
  a.nblocks = 0;
  if (n0 <= 0 || n1 <= 0 || n2 <= 0) return cudaSuccess;
  const int nt2 = (n2 + VEC - 1)/VEC;
  int bx = a.bx;
  while (bx > 1 && bx/2 >= nt2) bx /= 2;      // thin slabs: let the warp span rows instead
  int by = 256/bx;
  if (BINARY || GHOST) by = (128/bx > 0) ? 128/bx : 1;
  if (a.bt > 0) by = (a.bt/bx > 0) ? a.bt/bx : 1;
  while (by > 1 && by/2 >= n1) by /= 2;
  dim3 block(bx, by, 1);
  dim3 grid((nt2 + bx - 1)/bx, (n1 + by - 1)/by, n0);
  a.nblocks = (long long) grid.x*grid.y*grid.z;
  if constexpr (COLLIDE) {
    if (P.push) {
      // peer-memory halo: the variants that also store into the neighbours' halo sites
      // (single-lattice odd sweep: into the neighbours' interior)
      cudaLaunchConfig_t cfg;
      memset(&cfg, 0, sizeof(cfg));
      cfg.gridDim = grid; cfg.blockDim = block; cfg.stream = a.stream;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[0].val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = attr;
      cfg.numAttrs = a.pdl ? 1 : 0;
      if (a.diag) return cudaLaunchKernelEx(&cfg, lb_sweep_kernel<MODEL, VEC, PULL, COLLIDE, BINARY, GM, true, WALL, true, AA>, P);
      return cudaLaunchKernelEx(&cfg, lb_sweep_kernel<MODEL, VEC, PULL, COLLIDE, BINARY, GM, false, WALL, true, AA>, P);
    }
  }
  if (P.push) return cudaErrorInvalidValue;
  if (COLLIDE && a.diag) {
    lb_sweep_kernel<MODEL, VEC, PULL, COLLIDE, BINARY, GM, COLLIDE, WALL, false, AA>
      <<<grid, block, 0, a.stream>>>(P);
  }
  else {
    lb_sweep_kernel<MODEL, VEC, PULL, COLLIDE, BINARY, GM, false, WALL, false, AA>
      <<<grid, block, 0, a.stream>>>(P);
  }
  return cudaGetLastError();
}

template <bool PULL, bool COLLIDE>
static cudaError_t launch_pc(SweepArgs & a) {
  constexpr bool HAVE_BINARY = (MODEL::NDIM == 3);
  constexpr bool HAVE_NOISE = (MODEL::NVEL - MODEL::NHYDRO <= B200LB_MAXNOISE);
  if constexpr (!COLLIDE) {
    // pure streaming: collision flags are irrelevant
    if (a.binary) {
      if (!HAVE_BINARY) return cudaErrorInvalidValue;
      if (a.vec == 2) return launch_one<2, PULL, false, HAVE_BINARY, 0>(a);
      return launch_one<1, PULL, false, HAVE_BINARY, 0>(a);
    }
    if (a.vec == 2) return launch_one<2, PULL, false, false, 0>(a);
    return launch_one<1, PULL, false, false, 0>(a);
  }
  else {
  if (a.noise) {
    // fluctuating collision: ghost back-projection always, forward projection when the
    // scheme keeps ghost modes. One draw of the site generator yields ten variates
    // (src/noise.h:18), which is not enough for the 17 ghost modes of D3Q27.
    if constexpr (HAVE_NOISE) {
      if (a.binary) {
	if (!HAVE_BINARY) return cudaErrorInvalidValue;
	if constexpr (PULL && HAVE_BINARY) {
	  if (a.gmom) {
	    if (a.ghost) return launch_one<1, PULL, COLLIDE, HAVE_BINARY, 11>(a);
	    return launch_one<1, PULL, COLLIDE, HAVE_BINARY, 10>(a);
	  }
	}
	if (a.gmom) return cudaErrorInvalidValue;
	if (a.ghost) return launch_one<1, PULL, COLLIDE, HAVE_BINARY, 3>(a);
	return launch_one<1, PULL, COLLIDE, HAVE_BINARY, 2>(a);
      }
      if (a.ghost) return launch_one<1, PULL, COLLIDE, false, 3>(a);
      return launch_one<1, PULL, COLLIDE, false, 2>(a);
    }
    else {
      return cudaErrorInvalidValue;
    }
  }
  if (a.binary) {
    if (!HAVE_BINARY) return cudaErrorInvalidValue;
    if constexpr (PULL && HAVE_BINARY) {
      if (a.gmom) {
	if (a.ghost) return launch_one<1, PULL, COLLIDE, HAVE_BINARY, 9>(a);
	return launch_one<1, PULL, COLLIDE, HAVE_BINARY, 8>(a);
      }
    }
    if (a.gmom) return cudaErrorInvalidValue;
    if (a.ghost) return launch_one<1, PULL, COLLIDE, HAVE_BINARY, 1>(a);
    return launch_one<1, PULL, COLLIDE, HAVE_BINARY, 0>(a);
  }
  if (a.visc) {
    // viscosity model: one site per thread (two more divides and the site's eta live)
    if (a.ghost) return launch_one<1, PULL, COLLIDE, false, 5>(a);
    return launch_one<1, PULL, COLLIDE, false, 4>(a);
  }
  if (a.ghost) return launch_one<1, PULL, COLLIDE, false, 1>(a);
  if (a.vec == 2) return launch_one<2, PULL, COLLIDE, false, 0>(a);
  return launch_one<1, PULL, COLLIDE, false, 0>(a);
  }
}

// single-lattice streaming (option lattices = 1): plain single-fluid sweeps, M10 or a scheme
// that keeps ghost modes
static cudaError_t launch_aa(SweepArgs & a) {
  if (a.binary || a.noise || a.visc || a.wall || a.gmom) return cudaErrorInvalidValue;
  if (a.aa == 3 && a.P.push) return cudaErrorInvalidValue;
  if (a.aa == 1 && !a.pull && a.collide) {
    if (a.ghost) return launch_one<1, false, true, false, 1, false, 1>(a);
    if (a.vec == 2) return launch_one<2, false, true, false, 0, false, 1>(a);
    return launch_one<1, false, true, false, 0, false, 1>(a);
  }
  if (a.aa == 2 && a.pull && a.collide) {
    if (a.ghost) return launch_one<1, true, true, false, 1, false, 2>(a);
    if (a.vec == 2) return launch_one<2, true, true, false, 0, false, 2>(a);
    return launch_one<1, true, true, false, 0, false, 2>(a);
  }
  if (a.aa == 3 && a.pull && !a.collide) {
    if (a.vec == 2) return launch_one<2, true, false, false, 0, false, 3>(a);
    return launch_one<1, true, false, false, 0, false, 3>(a);
  }
  return cudaErrorInvalidValue;
}

static cudaError_t sweep_impl(SweepArgs & a) {
  if (a.aa) return launch_aa(a);
  if (a.pull && a.collide) return launch_pc<true, true>(a);
  if (a.pull) return launch_pc<true, false>(a);
  if (a.collide) return launch_pc<false, true>(a);
  return cudaErrorInvalidValue;
}

static cudaError_t phi_impl(const Layout & L, const int w[3], const double * g,
			    double * phi, double * jphi, bool pull, cudaStream_t s) {
  dim3 block(128, 2, 1);
  dim3 grid((L.n[2] + 127)/128, (L.n[1] + 1)/2, L.n[0]);
  if (pull && jphi) phi_from_g_kernel<MODEL, true, true><<<grid, block, 0, s>>>(L, w[0], w[1], w[2], g, phi, jphi);
  else if (pull) phi_from_g_kernel<MODEL, true, false><<<grid, block, 0, s>>>(L, w[0], w[1], w[2], g, phi, nullptr);
  else phi_from_g_kernel<MODEL, false, false><<<grid, block, 0, s>>>(L, 0, 0, 0, g, phi, nullptr);
  return cudaGetLastError();
}

static cudaError_t stats_impl(const Layout & L, const double * f, int ndist,
			      double * partial, dim3 grid, dim3 block, cudaStream_t s) {
  if (ndist == 2) stats_kernel<MODEL, 2><<<grid, block, 0, s>>>(L, f, partial);
  else stats_kernel<MODEL, 1><<<grid, block, 0, s>>>(L, f, partial);
  return cudaGetLastError();
}

static cudaError_t hydro_impl(const Layout & L, const double * f, int ndist,
			      const double * force, const double fg[3], double * rho,
			      double * u, int mode, const int wrap[3], cudaStream_t s) {
  dim3 block(128, 2, 1);
  dim3 grid((L.n[2] + 127)/128, (L.n[1] + 1)/2, L.n[0]);
  if (ndist == 2) hydro_from_fstar_kernel<MODEL, true><<<grid, block, 0, s>>>(L, f, force, fg[0], fg[1], fg[2], rho, u, mode, wrap[0], wrap[1], wrap[2]);
  else hydro_from_fstar_kernel<MODEL, false><<<grid, block, 0, s>>>(L, f, force, fg[0], fg[1], fg[2], rho, u, mode, wrap[0], wrap[1], wrap[2]);
  return cudaGetLastError();
}

static cudaError_t ghost_upload_impl(void) {
  static GhostTab T;
  build_ghost_tab<MODEL>(T);
  return cudaMemcpyToSymbol(c_ghost, &T, sizeof(GhostTab));
}

const ModelOps * B200LB_OPS_FN(void) {
  static ModelOps ops;
  static bool init = false;
  if (!init) {
    ops.nvel = MODEL::NVEL; ops.ndim = MODEL::NDIM; ops.nhydro = MODEL::NHYDRO;
    for (int d = 0; d < 3; d++) ops.dax[d] = MODEL::dax(d);
    for (int p = 0; p < MODEL::NVEL; p++) {
      for (int a = 0; a < 3; a++) ops.cv[p][a] = MODEL::cv(p, a);
      ops.wv[p] = MODEL::wv(p);
    }
    // sum_p w_p ma[m][p]^2 of the ghost rows: the reference's rna[] = 1/na[]
    // (src/model.c:378), which scales the ghost-mode noise (src/collision.c:1811)
    for (int k = 0; k < MODEL::NVEL - MODEL::NHYDRO; k++) {
      double wip = 0.0;
      for (int p = 0; p < MODEL::NVEL; p++) {
	const int c[3] = {MODEL::cv(p, 0), MODEL::cv(p, 1), MODEL::cv(p, 2)};
	const double r = ghost_row(MODEL::NVEL, k, c);
	wip += MODEL::wv(p)*r*r;
      }
      ops.ghost_norm[k] = wip;
    }
    ops.sweep = sweep_impl;
    ops.phi_from_g = phi_impl;
    ops.stats = stats_impl;
    ops.hydro = hydro_impl;
    ops.ghost_upload = ghost_upload_impl;
    init = true;
  }
  return &ops;
}

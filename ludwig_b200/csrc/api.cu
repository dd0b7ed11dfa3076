// api.cu -- the C ABI of include/b200lb.h: handle, layout, state machine,
// halo exchange (local periodic copies and NCCL send/recv), transfers.
//
// There is deliberately no CPU path in this file: every operation is a CUDA
// kernel launch or a CUDA/NCCL call, and b200lb_create fails without a device.

#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>
#include <nvtx3/nvToolsExt.h>     // header-only; ranges show up in Nsight Systems / ncu --nvtx

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "../../include/b200lb.h"
#include "common_kernels.cuh"
#include "zpack_kernels.cuh"
#include "le_field_kernels.cuh"
#include "open_bc_kernels.cuh"
#include "solid_grad_kernels.cuh"
#include "model_ops.cuh"
#include "ch_kernels.cuh"

// ---------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------

static thread_local char g_err[512] = "";

static int fail(const char * fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return -1;
}

#define CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
  return fail("%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); } while (0)

extern "C" const char * b200lb_last_error(void) { return g_err; }
extern "C" const char * b200lb_version(void) { return "b200lb 0.1 (sm_100a)"; }

extern "C" int b200lb_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

// ---------------------------------------------------------------------------
// NCCL, resolved at run time so the library loads on a box without it and
// shares whichever libnccl the process already has (e.g. the one torch loaded).
// ---------------------------------------------------------------------------

struct NcclApi {
  void * lib;
  decltype(&ncclGetUniqueId) GetUniqueId;
  decltype(&ncclCommInitRank) CommInitRank;
  decltype(&ncclCommDestroy) CommDestroy;
  decltype(&ncclSend) Send;
  decltype(&ncclRecv) Recv;
  decltype(&ncclAllReduce) AllReduce;
  decltype(&ncclGroupStart) GroupStart;
  decltype(&ncclGroupEnd) GroupEnd;
  decltype(&ncclGetErrorString) GetErrorString;
};

static NcclApi g_nccl = {};

static int nccl_load(void) {
  if (g_nccl.lib) return 0;
  void * lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) return fail("NCCL not found: %s", dlerror());
#define SYM(name) g_nccl.name = (decltype(g_nccl.name)) dlsym(lib, "nccl" #name); \
  if (!g_nccl.name) return fail("NCCL symbol nccl" #name " missing")
  SYM(GetUniqueId); SYM(CommInitRank); SYM(CommDestroy); SYM(Send); SYM(Recv); SYM(AllReduce);
  SYM(GroupStart); SYM(GroupEnd); SYM(GetErrorString);
#undef SYM
  g_nccl.lib = lib;
  return 0;
}

#define NC(call) do { ncclResult_t r_ = (call); if (r_ != ncclSuccess) \
  return fail("%s:%d %s: %s", __FILE__, __LINE__, #call, g_nccl.GetErrorString(r_)); } while (0)

// NVTX range over a scope: the stages of the time step by the names Ludwig's timers use
// (src/timer.h:47-63), so a timeline reads like its timer report
struct NvtxRange {
  explicit NvtxRange(const char * name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};

// ---------------------------------------------------------------------------
// handle
// ---------------------------------------------------------------------------

#define STAGE_BYTES ((size_t) 64 << 20)

struct HaloBuf {
  double * to_lo, * to_hi, * from_lo, * from_hi;
  i64 nface;
};

struct b200lb_s {
  b200lb_desc_t desc;
  const ModelOps * ops;
  Layout L;
  int hostdim[3];            // physical nall
  long nsite_host;
  int nplane;                // ndist*nvel

  double * lat[2];           // two lattices; lat[cur] is "f", the other "fprime"
  int cur;
  double * rho, * u, * force, * phi, * grad, * delsq;
  // flat walls (wall_t, src/wall.h)
  int wall_is[3];            // physical: isboundary
  double wall_ubot[3], wall_utop[3];
  int wall_slip;             // wall->param->slip.active
  double wall_sbot[3], wall_stop[3];
  bool wall_on;              // wall_commit was called with at least one boundary
  bool bbl_armed;            // wall_bbl was called: the next lb_propagation bounces at the walls
  bool prop_wall;            // ... and that propagation is the pending / running one
  double * fnet_dev;         // [3] momentum handed to the walls since the last wall_momentum
  double fnet_host[3];       // running total on the host (wall_momentum_add and read-outs)
  // Lees-Edwards planes (lees_edw_t, src/leesedwards.c)
  int le_nplanes;            // total over the system; 0 = none
  int le_nlocal;             // per rank
  double le_uy;
  int le_osc, le_period, le_nt0;
  long le_step;              // physics time step counter of b200lb_step
  long le_tstart;            // ... and the step this run started from (physics_control_init_time)
  double le_field_t;         // physics_control_time at which the order parameter is displaced
  double * le_phibuf;        // [2 x local planes][s0] buffer planes of phi (field_leesedwards)
  double * le_chbuf;         // finite-difference order parameter with planes: buffers and scratch (ch_le_run)
  bool open_on;              // open boundaries for the fluid (b200lb_open_bc_commit)
  int open_id;               // flow direction
  double open_u0[3], open_rho0;
  int last_grad_npt;         // stencil of the last field_grad_compute (7 or 27)
  double * le_buf;           // crossing populations of the plane sites
  // the order parameter as a finite-difference field ("free_energy symmetric":
  // phi_ch_t src/phi_cahn_hilliard.h:33, pth_t src/phi_force_stress.h)
  bool ch_on;
  int ch_order, ch_method;   // fd_advection_scheme_order, fe_force_method
  double ch_gmu[3];          // physics_grad_mu: external chemical potential gradient
  double * phi2;             // the buffer the next Cahn-Hilliard update writes
  bool ch_force_pending;     // phi_force_calculation recorded; evaluated by the update's pass
  bool ch_grad_valid;        // grad/delsq belong to the current phi
  bool ch_halo_done;         // ... and their halos in decomposed directions are current
  // viscosity model (visc_t, src/visc.h; hydro->eta)
  double * eta;              // [nsd] per-site shear viscosity, or null
  bool visc_on;              // lb_collide reads eta (p.have_visc_model)
  bool visc_arrhenius;       // b200lb_visc_update evaluates the Arrhenius model from phi
  double visc_eta_plus, visc_eta_minus, visc_phistar;
  double * noise;            // [2][nsd] per-site generator state (noise_init), or null
  bool noise_on;             // noise_present_set(NOISE_RHO)
  double nv_diag, nv_off, nv_bulk, nv_ghost[B200LB_MAXNOISE];

  // single-lattice streaming (option lattices = 1, AA pattern; site_kernels.cuh)
  int opt_lattices;          // 2 (default): f and fprime; 1: one array, even and odd sweeps alternate
  bool aa_swapped;           // lat[0] holds post-collision populations in their opposites' slots (after an even sweep)
  bool aa_ahead;             // lat[0] is one propagation ahead of the API (after an odd sweep, until lb_propagation)
  int fstar_mode;            // where ensure_hydro finds the post-collision populations: 0, 1, 2 (aux_kernels.cuh)

  // state machine (see b200lb_propagate)
  bool prop_pending;         // lat[cur] holds post-collision values still to be streamed
  double * jphi;             // [3][nsd] first moment of the pulled g (option g_moment), or null
  bool jphi_valid;           // ... written by phi_from_g from the pending lattice, which is unchanged since
  bool force_uniform;        // force field == fconst everywhere (hydro_f_zero)
  double fconst[3];
  bool u_fill_pending;       // hydro_u_zero recorded but not yet applied
  double uconst[3];
  bool hydro_asked;          // somebody wanted rho,u since the last collision (hydro_write = 2: the next one stores them)
  bool hydro_stale;          // collision ran without storing rho,u (recoverable from lat[cur], see ensure_hydro)
  bool last_ffield;          // the last collision read the per-site force field
  double last_fg[3];         // ... and this uniform part
  int  selfper[3];           // device axis is periodic within this rank (one rank in that direction)
  bool halo_deferred;        // lb_halo was called; self-periodic planes not yet copied (pulls wrap)
  bool phi_halo_deferred;    // the same for field_halo(phi): every reader of phi wraps
  int  grad_state;           // 0 = grad/delsq arrays hold the gradient of phi, 1 = gradient
                             // requested but computed inside the collision (materialised
                             // on demand), 2 = phi changed before that ever happened
  int  halo_deferred_scheme;

  b200lb_param_t prm;
  double rtau_shear, rtau_bulk, rtau2;
  double gkeep[B200LB_MAXGHOST];
  bool ghost;

  // options
  int opt_fused, opt_halo, opt_hydro_write, opt_vec, opt_block, opt_halo_lazy, opt_overlap, opt_direct, opt_diag, opt_loopback, opt_grad_lazy, opt_p2p, opt_gradient, opt_ch_fuse, opt_ch_planes, opt_ch_exact, opt_g_moment, opt_le_fd, opt_binary_walls, opt_threads;

  cudaStream_t s_main, s_x[2];
  cudaStream_t s_comm;       // halo exchange of decomposed directions (overlaps the interior sweep)
  cudaEvent_t ev0, ev1, ev_x[2];
  cudaEvent_t ev_bnd;        // boundary slabs of the current sweep written
  cudaEvent_t ev_halo;       // exchange on s_comm finished
  bool bnd_valid;            // ev_bnd belongs to the sweep that produced lat[cur]
  bool halo_inflight;        // s_main must wait for ev_halo before touching the lattice
  double * stage[2];
  double * partial; i64 npartial;
  double * diag; i64 diag_cap, diag_count;   // in-sweep diagnostics: per-block partials of the last collision
  double * stats_dev; double * stats_host;
  double * stats_chain;      // [2048] scratch of stats_final
  long launches;
  // CUDA events around the sweep launches alone (b200lb_kernel_timer_*)
  cudaEvent_t * tev; int tev_cap, tev_n; bool tev_on;

  ncclComm_t comm;
  HaloBuf hb[3];             // per device axis (multi-rank directions only)
  int hb_nel;

  // direct reduced exchange (one message per neighbour direction)
  int direct_state;          // 0 = not set up, 1 = ready, -1 = unavailable (fall back to x,y,z phases)
  int npack, nunpack;
  HaloMsg * msg_pack, * msg_unpack;        // device tables
  int pack_peer[B200LB_MAXMSG], unpack_peer[B200LB_MAXMSG];
  i64 pack_off[B200LB_MAXMSG], pack_cnt[B200LB_MAXMSG];
  i64 unpack_off[B200LB_MAXMSG], unpack_cnt[B200LB_MAXMSG];
  i64 pack_max, unpack_max;
  double * dsend, * drecv;

  // peer-memory halo (CUDA IPC mappings of the neighbours' lattices)
  int p2p_ready;
  double * peer_lat[2][27];
  P2PSync p2p;
  void * ipc_open[3*26 + 3]; int n_ipc_open;
  unsigned long long push_step;
  long launches_at_sync;      // launch counter right after the last p2p_sync_kernel: if it has not moved,
                              // that kernel is what the next sweep follows in the stream
  int opt_pdl;
  int opt_grad_direct;        // 27-point gradient: one thread per site through the caches (1) or planes staged in shared memory (0)
  bool pushed_valid;          // the halo data of lat[cur] was pushed by the sweep that wrote it
  // packed halo of the contiguous axis (KParams::zoff): every lattice array carries dense
  // planes behind its population planes for the values peer-memory sweeps push across z faces
  int opt_zpack;              // 0 off; arrangement of the packed planes, see zpack_geom
  int zpacked[2];             // != 0: the z-face halo of lat[i] is in its packed planes (in this arrangement), not in the row-end words
};

// A direction whose halo planes travel through send/receive buffers: decomposed over
// ranks, or (option "loopback", for testing and profiling the multi-rank machinery on
// one GPU) periodic within the rank with this rank as its own neighbour.
// option "loopback": 1 = every periodic direction, 8 | mask = the physical axes in mask
static bool loop_axis(const b200lb_s * h, int a) {
  if (!h->opt_loopback) return false;
  if (h->opt_loopback & 8) return ((h->opt_loopback >> a) & 1) != 0;
  return true;
}

static bool decomposed(const b200lb_s * h, int d) {
  const int a = h->ops->dax[d];
  if (h->L.h[d] == 0) return false;
  if (h->desc.cartsz[a] > 1) return true;
  return loop_axis(h, a) && h->desc.periodic[a];
}

static bool have_comm(const b200lb_s * h) { return h->comm != nullptr || h->opt_loopback; }

static void update_selfper(b200lb_s * h) {
  for (int d = 0; d < 3; d++) {
    const int a = h->ops->dax[d];
    h->selfper[d] = (h->L.h[d] == 1 && h->desc.cartsz[a] == 1 && h->desc.periodic[a]
		     && !loop_axis(h, a)) ? 1 : 0;
  }
}

static int site_kparams(b200lb_s * h, KParams & P) {
  memset(&P, 0, sizeof(P));
  P.L = h->L;
  for (int d = 0; d < 3; d++) { P.lo[d] = 0; P.hi[d] = h->L.n[d]; P.wrap[d] = h->selfper[d]; }
  P.rtau_shear = h->rtau_shear; P.rtau_bulk = h->rtau_bulk;
  P.rtau2 = h->rtau2; P.A = h->prm.a; P.B = h->prm.b; P.kappa = h->prm.kappa;
  for (int k = 0; k < B200LB_MAXGHOST; k++) P.gkeep[k] = h->gkeep[k];
  if (h->wall_on) {
    // wall_init_uw (src/wall.c:864-893): wall velocities only for a single pair of walls
    const int nwall = h->wall_is[0] + h->wall_is[1] + h->wall_is[2];
    int iw = -1;
    if (nwall == 1) { for (int a = 0; a < 3; a++) if (h->wall_is[a]) iw = a; }
    for (int d = 0; d < 3; d++) {
      const int a = h->ops->dax[d];
      const bool is = h->wall_is[a] && h->L.h[d] == 1;
      P.wall_lo[d] = is && h->desc.cartcoords[a] == 0;
      P.wall_hi[d] = is && h->desc.cartcoords[a] == h->desc.cartsz[a] - 1;
    }
    const int Q = h->ops->nvel;
    for (int q = 1; q < Q; q++) {
      // incoming q = reversed outgoing link ij = Q - q; the link points into the bottom
      // wall when c_ij[iw] = -1 (src/wall.c:886-887)
      const int ij = Q - q;
      const double * uw = nullptr;
      if (iw >= 0 && h->ops->cv[ij][iw] == -1) uw = h->wall_ubot;
      if (iw >= 0 && h->ops->cv[ij][iw] == +1) uw = h->wall_utop;
      double cdotu = 0.0;
      if (uw) cdotu = h->ops->cv[ij][0]*uw[0] + h->ops->cv[ij][1]*uw[1] + h->ops->cv[ij][2]*uw[2];
      P.wall_corr[q] = -(2.0*3.0*h->ops->wv[ij]*h->prm.rho0*cdotu);
    }
    P.wall_slip = h->wall_slip;
    for (int d = 0; d < 3; d++) {
      const int a = h->ops->dax[d];
      P.wall_s_lo[d] = h->wall_sbot[a];
      P.wall_s_hi[d] = h->wall_stop[a];
    }
    // mirror images: the velocity index of c_p with the masked device-axis components reversed
    for (int p = 0; p < Q; p++) {
      for (int mask = 0; mask < 8; mask++) {
	int want[3];
	for (int d = 0; d < 3; d++) {
	  const int c = h->ops->cv[p][h->ops->dax[d]];
	  want[d] = (mask & (1 << d)) ? -c : c;
	}
	int found = p;
	for (int q = 0; q < Q; q++) {
	  bool same = true;
	  for (int d = 0; d < 3; d++) if (h->ops->cv[q][h->ops->dax[d]] != want[d]) same = false;
	  if (same) found = q;
	}
	P.wall_q[p][mask] = (signed char) found;
      }
    }
  }
  P.noise = h->noise;
  P.nv_diag = h->nv_diag; P.nv_off = h->nv_off; P.nv_bulk = h->nv_bulk;
  for (int k = 0; k < B200LB_MAXNOISE; k++) P.nv_ghost[k] = h->nv_ghost[k];
  P.nghost_draw = h->prm.ghost_off ? 0 : 1;
  P.nlv_big = sqrt(2.0 + sqrt(2.0));          // rtable of src/noise.c:70-77
  P.nlv_small = sqrt(2.0 - sqrt(2.0));
  return 0;
}

static size_t plane_bytes(const b200lb_s * h) { return (size_t) h->L.nsd*sizeof(double); }

// one lattice array: the population planes, then the packed z-face halo planes (two faces,
// at most nine populations each, over the halo-inclusive x-y extent; 4.8 MB at 256^3)
static i64 zpack_doubles(const b200lb_s * h) {
  const i64 n = (i64) 2*9*h->L.na[0]*h->L.na[1]*4;      // room for a 32-byte sector per column
  return (n + 15)/16*16;
}

// arrangement of the packed planes (option zpack), y fastest: 1 = one word per column
// (default); 2 = a whole 32-byte sector per column, written as one 256-bit store. Measured
// (profiles/r2b_push_axes_zpack2.log, r2b_bench_2gpu_112_zpack*.json): 1 is the faster one,
// on one GPU and across NVLink; x fastest makes no difference.
static void zpack_geom(const b200lb_s * h, int mode, i64 * zsx, i64 * zsy, i64 * zsp, int * zvec) {
  const i64 na0 = h->L.na[0], na1 = h->L.na[1];
  const i64 w = (mode == 2) ? 4 : 1;
  *zvec = (mode == 2) ? 1 : 0;
  *zsy = w; *zsx = w*na1;
  *zsp = w*na0*na1;
}
static size_t lat_bytes(const b200lb_s * h) {
  return plane_bytes(h)*h->nplane + (size_t) zpack_doubles(h)*sizeof(double);
}

// lazily allocated field planes, zero-filled
static int need_planes(b200lb_s * h, double ** p, int nplanes) {
  if (*p) return 0;
  CU(cudaMalloc(p, plane_bytes(h)*nplanes));
  CU(cudaMemsetAsync(*p, 0, plane_bytes(h)*nplanes, h->s_main));
  return 0;
}
static int need_hydro(b200lb_s * h) {
  if (need_planes(h, &h->rho, 1)) return -1;
  return need_planes(h, &h->u, 3);
}

// ---------------------------------------------------------------------------
// create / destroy
// ---------------------------------------------------------------------------

static int derive_params(b200lb_s * h) {
  const b200lb_param_t & p = h->prm;
  const double cs2 = 1.0/3.0;
  const ModelOps * m = h->ops;
  const int ng = m->nvel - m->nhydro;
  // lb_collision_relaxation_times_set, src/collision.c:1181-1264
  h->rtau_shear = 1.0/(0.5 + p.eta_shear/(p.rho0*cs2));
  h->rtau_bulk  = 1.0/(0.5 + p.eta_bulk/(p.rho0*cs2));
  h->rtau2 = 2.0/(1.0 + 2.0*p.mobility);              // src/collision.c:1973
  double rt[B200LB_MAXGHOST];
  for (int k = 0; k < B200LB_MAXGHOST; k++) rt[k] = 1.0;
  if (p.nrelax == B200LB_RELAX_M10) {
    /* ghosts fully relaxed */
  }
  else if (p.nrelax == B200LB_RELAX_BGK) {
    for (int k = 0; k < ng; k++) rt[k] = h->rtau_shear;
    h->rtau_bulk = h->rtau_shear;
  }
  else if (p.nrelax == B200LB_RELAX_TRT) {
    if (m->nvel != 19) return fail("TRT relaxation is available for D3Q19 only (as in the reference)");
    const double tau = p.eta_shear/(p.rho0*cs2);
    double r = 0.5 + 2.0*tau/(tau + 3.0/8.0);
    if (r > 2.0) r = 2.0;
    // modes 10,14,18 (even) relax with the shear rate, the odd ones with r
    for (int k = 0; k < ng; k++) rt[k] = (k == 0 || k == 4 || k == 8) ? h->rtau_shear : r;
  }
  else {
    return fail("unknown relaxation scheme %d", p.nrelax);
  }
  if (p.ghost_off && h->desc.ndist == 2) {
    // lb_collision_noise_var_set (src/collision.c:1607-1614) resets lb->param->rtau of
    // the ghosts, which only the binary collision reads (src/collision.c:918)
    for (int k = 0; k < ng; k++) rt[k] = 1.0;
  }
  h->ghost = false;
  for (int k = 0; k < B200LB_MAXGHOST; k++) {
    h->gkeep[k] = (k < ng) ? 1.0 - rt[k] : 0.0;
    if (h->gkeep[k] != 0.0) h->ghost = true;
  }
  // Variances of the fluctuating stress and ghost modes at temperature kt:
  // lb_fluctuations_var_eta / _var_bulk / _var_ghost, src/collision.c:1756-1815
  // (kT is normalised by cs^2). With the bulk and shear relaxation times of the scheme.
  {
    const double ktn = p.kt*3.0;
    const double tau_s = 1.0/h->rtau_shear, tau_b = 1.0/h->rtau_bulk;
    const double var_eta = sqrt(ktn)*sqrt(1.0/9.0)*sqrt((tau_s + tau_s - 1.0)/(tau_s*tau_s));
    h->nv_diag = var_eta*sqrt(2.0);
    h->nv_off  = var_eta;
    h->nv_bulk = sqrt(ktn)*sqrt(2.0/9.0)*sqrt((tau_b + tau_b - 1.0)/(tau_b*tau_b));
    for (int k = 0; k < B200LB_MAXNOISE; k++) {
      h->nv_ghost[k] = 0.0;
      if (k < ng) {
	const double tau_g = 1.0/rt[k];
	h->nv_ghost[k] = sqrt(ktn*m->ghost_norm[k])*sqrt((tau_g + tau_g - 1.0)/(tau_g*tau_g));
      }
    }
  }
  return 0;
}

static int create_impl(const b200lb_desc_t * desc, b200lb_s * h, const ModelOps * ops);

extern "C" int b200lb_create(const b200lb_desc_t * desc, b200lb_t ** out) {

  if (!desc || !out) return fail("null argument");
  *out = nullptr;
  int ndev = b200lb_device_count();
  if (ndev <= 0) return fail("no CUDA device: libb200lb has no CPU path");
  if (desc->nhalo != 1) return fail("nhalo must be 1 (got %d)", desc->nhalo);
  if (desc->ndist != 1 && desc->ndist != 2) return fail("ndist must be 1 or 2");
  if (desc->lattices < 0 || desc->lattices > 2) return fail("desc.lattices: 0 or 2 (two lattices), 1 (single-lattice streaming)");
  if (desc->lattices == 1 && desc->ndist != 1) return fail("desc.lattices = 1: the single-fluid distribution (ndist = 1)");

  const ModelOps * ops = nullptr;
  if (desc->nvel == 19) ops = model_ops_d3q19();
  if (desc->nvel == 27) ops = model_ops_d3q27();
  if (desc->nvel == 9)  ops = model_ops_d2q9();
  if (!ops) return fail("unsupported nvel %d (9, 19, 27)", desc->nvel);
  if (ops->ndim != desc->ndim) return fail("ndim %d does not match nvel %d", desc->ndim, desc->nvel);
  if (desc->ndist == 2 && ops->ndim != 3) return fail("two distributions need a 3-d model");
  if (ops->ndim == 2 && desc->nlocal[2] != 1) return fail("2-d model needs nlocal[2] = 1");
  for (int a = 0; a < 3; a++) if (desc->nlocal[a] < 1) return fail("bad nlocal");

  if (desc->device >= 0) CU(cudaSetDevice(desc->device));

  b200lb_s * h = (b200lb_s *) calloc(1, sizeof(b200lb_s));
  if (!h) return fail("out of host memory");
  if (create_impl(desc, h, ops) != 0) {
    char keep[sizeof(g_err)];
    memcpy(keep, g_err, sizeof(keep));
    b200lb_destroy(h);                     // releases whatever was allocated so far
    memcpy(g_err, keep, sizeof(keep));
    return -1;
  }
  *out = h;
  return 0;
}

static int create_impl(const b200lb_desc_t * desc, b200lb_s * h, const ModelOps * ops) {
  h->desc = *desc;
  h->ops = ops;
  h->nplane = desc->ndist*desc->nvel;

  Layout & L = h->L;
  int zalign = desc->zalign > 0 ? desc->zalign : 16;
  if (zalign < 2 || (zalign & 1)) return fail("zalign must be even");
  h->nsite_host = 1;
  for (int a = 0; a < 3; a++) {
    h->hostdim[a] = desc->nlocal[a] + 2;
    h->nsite_host *= h->hostdim[a];
  }
  for (int d = 0; d < 3; d++) {
    const int a = ops->dax[d];
    L.n[d] = desc->nlocal[a];
    L.h[d] = (ops->ndim == 2 && a == 2) ? 0 : 1;
    L.na[d] = L.n[d] + 2*L.h[d];
  }
  update_selfper(h);
  L.z0 = zalign - L.h[2];
  {
    i64 row = L.z0 + L.na[2];
    row = (row + zalign - 1)/zalign*zalign;
    L.s1 = row;
    L.s0 = L.s1*L.na[1];
    L.nsd = L.s0*L.na[0] + zalign;       // slack: masked lanes may read one row position past the end
  }

  CU(cudaStreamCreateWithFlags(&h->s_main, cudaStreamNonBlocking));
  {
    // highest priority: the small pack / NCCL / unpack kernels must get SMs while
    // the interior sweep (tens of thousands of blocks) is running on s_main
    int plo = 0, phi_ = 0;
    CU(cudaDeviceGetStreamPriorityRange(&plo, &phi_));
    CU(cudaStreamCreateWithPriority(&h->s_comm, cudaStreamNonBlocking, phi_));
  }
  CU(cudaEventCreateWithFlags(&h->ev_bnd, cudaEventDisableTiming));
  CU(cudaEventCreateWithFlags(&h->ev_halo, cudaEventDisableTiming));
  for (int i = 0; i < 2; i++) {
    CU(cudaStreamCreateWithFlags(&h->s_x[i], cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&h->ev_x[i], cudaEventDisableTiming));
  }
  CU(cudaEventCreate(&h->ev0));
  CU(cudaEventCreate(&h->ev1));

  const size_t pb = plane_bytes(h);
  h->opt_lattices = (desc->lattices == 1) ? 1 : 2;
  for (int i = 0; i < h->opt_lattices; i++) {
    CU(cudaMalloc(&h->lat[i], lat_bytes(h)));
    CU(cudaMemsetAsync(h->lat[i], 0, lat_bytes(h), h->s_main));
  }
  // rho, u, force, grad and delsq are allocated when first needed (need_planes): a
  // run that keeps hydro on demand and never has a per-site force carries only the two
  // lattices (1024^3 on two GPUs: 2 x 82 GB fit in 180 GB that way).
  if (desc->ndist == 2) {
    CU(cudaMalloc(&h->phi, pb));
    CU(cudaMemsetAsync(h->phi, 0, pb, h->s_main));
  }
  // diagnostics scratch: one partial per block of a (128 x 2) tiling
  h->npartial = (i64)((L.n[2] + 127)/128)*((L.n[1] + 1)/2)*L.n[0];
  CU(cudaMalloc(&h->partial, 8*sizeof(double)*h->npartial));
  h->diag_cap = 4*h->npartial + 4096;
  CU(cudaMalloc(&h->diag, 8*sizeof(double)*h->diag_cap));
  CU(cudaMalloc(&h->stats_dev, 8*sizeof(double)));
  CU(cudaMalloc(&h->stats_chain, 2048*sizeof(double)));
  CU(cudaMallocHost(&h->stats_host, 8*sizeof(double)));

  CU(ops->ghost_upload());

  h->force_uniform = true;
  h->opt_fused = 1;
  h->opt_halo = B200LB_HALO_FULL;
  h->opt_hydro_write = 2;        // auto: see run_sweep
  // two sites per thread (128-bit accesses) for D3Q19; one for D3Q27 (27 live populations
  // per site) and D2Q9 (few streams, occupancy matters more): measured, profiles/r1_bench_*
  h->opt_vec = (desc->nvel == 19) ? 2 : 1;
  h->opt_block = 128;
  h->opt_zpack = 1;
  h->opt_halo_lazy = 1;
  h->opt_overlap = 1;
  h->opt_direct = 1;
  h->opt_grad_lazy = 1;
  h->opt_gradient = 7;
  h->opt_p2p = 1;
  h->opt_grad_direct = 1;
  h->opt_pdl = 0;                // measured: no gain, so the plain launch stays the default
  h->launches_at_sync = -1;
  h->opt_ch_fuse = 1;
  h->opt_ch_planes = 8;
  h->opt_ch_exact = 0;
  h->opt_g_moment = 1;
  h->opt_le_fd = 1;
  h->opt_binary_walls = 1;

  b200lb_param_t p0;
  memset(&p0, 0, sizeof(p0));
  p0.rho0 = 1.0; p0.eta_shear = 1.0/6.0; p0.eta_bulk = 1.0/6.0;
  h->prm = p0;
  if (derive_params(h)) return -1;

  CU(cudaStreamSynchronize(h->s_main));
  return 0;
}

extern "C" int b200lb_destroy(b200lb_t * h) {
  if (!h) return 0;
  cudaDeviceSynchronize();
  if (h->comm) g_nccl.CommDestroy(h->comm);
  for (int i = 0; i < h->n_ipc_open; i++) cudaIpcCloseMemHandle(h->ipc_open[i]);
  for (int i = 0; i < h->tev_cap; i++) cudaEventDestroy(h->tev[i]);
  free(h->tev);
  cudaFree(h->p2p.flags);
  cudaFree(h->msg_pack); cudaFree(h->msg_unpack); cudaFree(h->dsend); cudaFree(h->drecv);
  for (int d = 0; d < 3; d++) {
    cudaFree(h->hb[d].to_lo); cudaFree(h->hb[d].to_hi);
    cudaFree(h->hb[d].from_lo); cudaFree(h->hb[d].from_hi);
  }
  for (int i = 0; i < 2; i++) {
    cudaFree(h->lat[i]);
    cudaFree(h->stage[i]);
    cudaStreamDestroy(h->s_x[i]);
    cudaEventDestroy(h->ev_x[i]);
  }
  cudaFree(h->rho); cudaFree(h->u); cudaFree(h->force);
  cudaFree(h->phi2); cudaFree(h->eta); cudaFree(h->jphi); cudaFree(h->le_phibuf); cudaFree(h->le_chbuf);
  cudaFree(h->phi); cudaFree(h->grad); cudaFree(h->delsq); cudaFree(h->noise); cudaFree(h->fnet_dev); cudaFree(h->le_buf);
  cudaFree(h->partial); cudaFree(h->diag); cudaFree(h->stats_dev); cudaFree(h->stats_chain); cudaFreeHost(h->stats_host);
  cudaEventDestroy(h->ev0); cudaEventDestroy(h->ev1);
  cudaStreamDestroy(h->s_main);
  cudaStreamDestroy(h->s_comm);
  cudaEventDestroy(h->ev_bnd); cudaEventDestroy(h->ev_halo);
  free(h);
  return 0;
}

extern "C" long b200lb_nsite(const b200lb_t * h) { return h ? h->nsite_host : 0; }

static int materialise(b200lb_s * h);
static int aa_supported(const b200lb_s * h);
static int join_comm(b200lb_s * h);
static int flush_halo(b200lb_s * h);
static int zhalo_flush(b200lb_s * h, cudaStream_t st);
static int flush_phi_halo(b200lb_s * h);
static int resolve_ch_force(b200lb_s * h);

extern "C" int b200lb_set_option(b200lb_t * h, const char * name, long v) {
  if (!h || !name) return fail("null argument");
  h->jphi_valid = false;   // an option may change how the lattice is addressed (loopback) or swept
  if (!strcmp(name, "fused")) {
    if (!v && h->prop_pending) { if (materialise(h)) return -1; }
    h->opt_fused = (v != 0);
    return 0;
  }
  if (!strcmp(name, "lattices")) {
    if (v != 1 && v != 2) return fail("lattices: 2 (f and fprime, default) or 1 (single-lattice streaming)");
    if (v == h->opt_lattices) return 0;
    if (materialise(h)) return -1;                   // the natural, streamed lattice in lat[cur]
    if (join_comm(h)) return -1;
    if (zhalo_flush(h, h->s_main)) return -1;        // the arrays may change places below
    CU(cudaStreamSynchronize(h->s_main));
    if (v == 1) {
      if (h->p2p_ready && h->cur == 1) {
        // the neighbours have lat[0] and lat[1] mapped by name: keep the current state in lat[0]
        CU(cudaMemcpyAsync(h->lat[0], h->lat[1], plane_bytes(h)*h->nplane, cudaMemcpyDeviceToDevice, h->s_main));
        CU(cudaStreamSynchronize(h->s_main));
        h->cur = 0;
      }
      h->opt_lattices = 1;
      if (aa_supported(h)) { h->opt_lattices = 2; return -1; }
      if (h->cur == 1) { double * t = h->lat[0]; h->lat[0] = h->lat[1]; h->lat[1] = t; h->cur = 0; }
      if (!h->p2p_ready) {
        // (with the peer-memory halo the neighbours have the array mapped: it stays, unused;
        // desc.lattices = 1 never allocates it)
        CU(cudaFree(h->lat[1]));
        h->lat[1] = nullptr;
      }
      h->aa_swapped = false; h->aa_ahead = false; h->fstar_mode = 0;
    }
    else {
      if (h->aa_swapped || h->aa_ahead) return fail("lattices = 2: finish the time step first (lb_propagation)");
      if (h->p2p_ready && !h->lat[1]) return fail("lattices = 2: the neighbours have mapped a single lattice (desc.lattices = 1)");
      if (!h->lat[1]) {
        CU(cudaMalloc(&h->lat[1], lat_bytes(h)));
        CU(cudaMemsetAsync(h->lat[1], 0, lat_bytes(h), h->s_main));
      }
      h->opt_lattices = 2;
    }
    return 0;
  }
  if (!strcmp(name, "halo")) {
    if (v != B200LB_HALO_FULL && v != B200LB_HALO_REDUCED) return fail("bad halo scheme %ld", v);
    h->opt_halo = (int) v; return 0;
  }
  if (!strcmp(name, "hydro_write")) {
    if (v < 0 || v > 2) return fail("hydro_write: 0 (on demand), 1 (every collision, as the reference), 2 (auto)");
    h->opt_hydro_write = (int) v; return 0;
  }
  if (!strcmp(name, "halo_lazy")) { h->opt_halo_lazy = (v != 0); return 0; }
  if (!strcmp(name, "overlap")) { h->opt_overlap = (v != 0); return 0; }
  if (!strcmp(name, "direct")) { h->opt_direct = (v != 0); return 0; }
  if (!strcmp(name, "diag")) { h->opt_diag = (v != 0); return 0; }
  if (!strcmp(name, "grad_lazy")) { h->opt_grad_lazy = (v != 0); return 0; }
  if (!strcmp(name, "ch_fuse")) {
    if (!v) { if (resolve_ch_force(h)) return -1; }
    h->opt_ch_fuse = (v != 0);
    return 0;
  }
  if (!strcmp(name, "ch_exact")) {
    if (resolve_ch_force(h)) return -1;
    h->opt_ch_exact = (v != 0);
    return 0;
  }
  if (!strcmp(name, "g_moment")) { h->opt_g_moment = (v != 0); h->jphi_valid = false; return 0; }
  if (!strcmp(name, "le_fd")) { h->opt_le_fd = (v != 0); return 0; }
  if (!strcmp(name, "binary_walls")) { h->opt_binary_walls = (v != 0); return 0; }
  if (!strcmp(name, "ch_planes")) {
    if (v < 1 || v > 4096) return fail("ch_planes: x-planes marched by one block of the order-parameter kernel, 1..4096");
    h->opt_ch_planes = (int) v;
    return 0;
  }
  if (!strcmp(name, "gradient")) {
    if (v != 7 && v != 27 && v != 28) return fail("gradient: 7 or 27 points (28: the 27-point stencil next to walls, 3d_27pt_solid)");
    if (v >= 27 && h->desc.ndim != 3) return fail("the 27-point gradient is three-dimensional");
    h->opt_gradient = (int) v;
    return 0;
  }
  if (!strcmp(name, "pdl")) { h->opt_pdl = (v != 0); return 0; }
  if (!strcmp(name, "zpack")) {
    if (v < 0 || v > 2) return fail("zpack: 0 (row-end halo words), 1 (packed planes, one word per column), 2 (packed planes, a 32-byte sector per column)");
    h->opt_zpack = (int) v; return 0;
  }
  if (!strcmp(name, "grad_direct")) { h->opt_grad_direct = (v != 0); return 0; }
  if (!strcmp(name, "p2p")) { if (join_comm(h)) return -1; h->opt_p2p = (v != 0); h->pushed_valid = false; return 0; }
  if (!strcmp(name, "loopback")) {
    if (h->desc.nranks > 1 && v) return fail("loopback is for single-rank lattices");
    if (join_comm(h)) return -1;
    if (flush_halo(h)) return -1;
    if (h->phi) { if (flush_phi_halo(h)) return -1; }
    if (v < 0 || v > 15) return fail("loopback: 0, 1 (all periodic directions) or 8 | mask of physical axes");
    h->opt_loopback = (int) v;
    update_selfper(h);
    h->p2p_ready = 0; h->pushed_valid = false;
    // message tables depend on which directions are exchanged
    CU(cudaStreamSynchronize(h->s_main));
    CU(cudaStreamSynchronize(h->s_comm));
    cudaFree(h->msg_pack); cudaFree(h->msg_unpack); cudaFree(h->dsend); cudaFree(h->drecv);
    h->msg_pack = h->msg_unpack = nullptr; h->dsend = h->drecv = nullptr;
    h->direct_state = 0;
    return 0;
  }
  if (!strcmp(name, "vec")) {
    if (v != 1 && v != 2) return fail("vec must be 1 or 2");
    h->opt_vec = (int) v; return 0;
  }
  if (!strcmp(name, "block")) {
    if (v != 32 && v != 64 && v != 128 && v != 256) return fail("block must be 32, 64, 128 or 256");
    h->opt_block = (int) v; return 0;
  }
  if (!strcmp(name, "threads")) {
    if (v != 0 && v != 64 && v != 128 && v != 256) return fail("threads per block of the sweep: 0 (default), 64, 128 or 256");
    h->opt_threads = (int) v; return 0;
  }
  return fail("unknown option '%s'", name);
}

extern "C" long b200lb_get_option(const b200lb_t * h, const char * name) {
  if (!h || !name) return -1;
  if (!strcmp(name, "fused")) return h->opt_fused;
  if (!strcmp(name, "lattices")) return h->opt_lattices;
  if (!strcmp(name, "halo")) return h->opt_halo;
  if (!strcmp(name, "hydro_write")) return h->opt_hydro_write;
  if (!strcmp(name, "halo_lazy")) return h->opt_halo_lazy;
  if (!strcmp(name, "overlap")) return h->opt_overlap;
  if (!strcmp(name, "direct")) return h->opt_direct;
  if (!strcmp(name, "diag")) return h->opt_diag;
  if (!strcmp(name, "grad_lazy")) return h->opt_grad_lazy;
  if (!strcmp(name, "gradient")) return h->opt_gradient;
  if (!strcmp(name, "ch_fuse")) return h->opt_ch_fuse;
  if (!strcmp(name, "ch_planes")) return h->opt_ch_planes;
  if (!strcmp(name, "ch_exact")) return h->opt_ch_exact;
  if (!strcmp(name, "g_moment")) return h->opt_g_moment;
  if (!strcmp(name, "le_fd")) return h->opt_le_fd;
  if (!strcmp(name, "binary_walls")) return h->opt_binary_walls;
  if (!strcmp(name, "jphi_valid")) return h->jphi_valid;
  if (!strcmp(name, "p2p")) return h->opt_p2p;
  if (!strcmp(name, "pdl")) return h->opt_pdl;
  if (!strcmp(name, "zpack")) return h->opt_zpack;
  if (!strcmp(name, "zpacked")) return (h->zpacked[0] ? 1 : 0) | (h->zpacked[1] ? 2 : 0);
  if (!strcmp(name, "grad_direct")) return h->opt_grad_direct;
  if (!strcmp(name, "p2p_ready")) return h->p2p_ready;
  if (!strcmp(name, "loopback")) return h->opt_loopback;
  if (!strcmp(name, "direct_state")) return h->direct_state;
  if (!strcmp(name, "halo_deferred")) return h->halo_deferred;
  if (!strcmp(name, "vec")) return h->opt_vec;
  if (!strcmp(name, "block")) return h->opt_block;
  if (!strcmp(name, "threads")) return h->opt_threads;
  if (!strcmp(name, "prop_pending")) return h->prop_pending;
  if (!strcmp(name, "nsd")) return (long) h->L.nsd;
  if (!strcmp(name, "row_pitch")) return (long) h->L.s1;
  return -1;
}

extern "C" int b200lb_set_param(b200lb_t * h, const b200lb_param_t * prm) {
  if (!h || !prm) return fail("null argument");
  if (h->ch_force_pending && (prm->a != h->prm.a || prm->b != h->prm.b || prm->kappa != h->prm.kappa)) {
    if (resolve_ch_force(h)) return -1;      // the recorded force belongs to the old free energy
  }
  b200lb_param_t keep = h->prm;
  h->prm = *prm;
  if (derive_params(h)) { h->prm = keep; derive_params(h); return -1; }
  return 0;
}

// ---------------------------------------------------------------------------
// sweeps
// ---------------------------------------------------------------------------

// peer-memory halo: how long p2p_sync_kernel waits for a neighbour before it traps
#define P2P_WATCHDOG_NS 30000000000ULL

// s_main waits for an exchange still running on s_comm
static int join_comm(b200lb_s * h) {
  if (h->halo_inflight) {
    CU(cudaStreamWaitEvent(h->s_main, h->ev_halo, 0));
    h->halo_inflight = false;
  }
  return 0;
}

// Sub-boxes for the comm/compute overlap: along every direction decomposed over
// ranks the two outermost slabs (one site thick; sixteen -- one 128-byte line -- along
// the contiguous axis so that slab and interior rows stay line-aligned and coalesced)
// are swept first, then the interior.
// The slabs are disjoint, so the in-place collision touches each site once.
// Returns the number of boxes; the LAST one is the interior. 1 = no split.
static int split_boxes(const b200lb_s * h, int lo[7][3], int hi[7][3]) {
  const Layout & L = h->L;
  int ilo[3] = {0, 0, 0}, ihi[3] = {L.n[0], L.n[1], L.n[2]};
  int nb = 0;
  bool any = false;
  for (int d = 0; d < 3; d++) {
    if (!decomposed(h, d)) continue;
    const int t = (d == 2) ? ((L.n[2] >= 64) ? 16 : 2) : 1;
    if (L.n[d] < 3*t + 1 || (d == 2 && (L.n[2] & 1))) return 1;
    any = true;
  }
  if (!any || !have_comm(h) || !h->opt_overlap) return 1;
  for (int d = 0; d < 3; d++) {
    if (!decomposed(h, d)) continue;
    const int t = (d == 2) ? ((L.n[2] >= 64) ? 16 : 2) : 1;
    for (int side = 0; side < 2; side++) {
      for (int e = 0; e < 3; e++) { lo[nb][e] = ilo[e]; hi[nb][e] = ihi[e]; }
      if (side == 0) { lo[nb][d] = 0; hi[nb][d] = t; }
      else { lo[nb][d] = L.n[d] - t; hi[nb][d] = L.n[d]; }
      nb++;
    }
    ilo[d] = t; ihi[d] = L.n[d] - t;
  }
  for (int e = 0; e < 3; e++) { lo[nb][e] = ilo[e]; hi[nb][e] = ihi[e]; }
  return nb + 1;
}

// Sub-boxes when a bounce-back is folded into the pull: the slabs next to the walls this
// rank holds (one site thick; sixteen along the contiguous axis to keep rows line-aligned)
// run the wall variant of the kernel, the interior -- the LAST box -- the plain one.
// iswall[b] says which. Lattices too thin to split are swept whole by the wall variant.
static int wall_boxes(const b200lb_s * h, const KParams & P, int lo[7][3], int hi[7][3], bool iswall[7]) {
  const Layout & L = h->L;
  int ilo[3] = {0, 0, 0}, ihi[3] = {L.n[0], L.n[1], L.n[2]};
  int nb = 0;
  for (int d = 0; d < 3; d++) {
    if (!P.wall_lo[d] && !P.wall_hi[d]) continue;
    const int t = (d == 2) ? 16 : 1;
    if (L.n[d] < 2*t + 1 || (d == 2 && (L.n[2] % 2))) {
      for (int e = 0; e < 3; e++) { lo[0][e] = 0; hi[0][e] = L.n[e]; }
      iswall[0] = true;
      return 1;
    }
  }
  for (int d = 0; d < 3; d++) {
    const int t = (d == 2) ? 16 : 1;
    for (int side = 0; side < 2; side++) {
      if (!(side == 0 ? P.wall_lo[d] : P.wall_hi[d])) continue;
      for (int e = 0; e < 3; e++) { lo[nb][e] = ilo[e]; hi[nb][e] = ihi[e]; }
      if (side == 0) { lo[nb][d] = 0; hi[nb][d] = t; }
      else { lo[nb][d] = L.n[d] - t; hi[nb][d] = L.n[d]; }
      iswall[nb] = true;
      nb++;
    }
    if (P.wall_lo[d]) ilo[d] = t;
    if (P.wall_hi[d]) ihi[d] = L.n[d] - t;
  }
  for (int e = 0; e < 3; e++) { lo[nb][e] = ilo[e]; hi[nb][e] = ihi[e]; }
  iswall[nb] = false;
  return nb + 1;
}

// ---------------------------------------------------------------------------
// Packed halo of the contiguous axis (KParams::zoff, site_kernels.cuh). Peer-memory sweeps
// of the single fluid on two lattices leave what crosses a z face in the dense planes behind
// the neighbour's lattice, and read their own z-face halo from theirs; zpacked[i] records
// that lat[i] is in that state. Everything else that reads halo sites of a lattice goes
// through zhalo_unpack first (the row-end halo words are then current and the flag clear).
// ---------------------------------------------------------------------------

static bool zpack_usable(const b200lb_s * h) {
  return h->opt_zpack && h->desc.ndist == 1 && h->opt_lattices == 2 && !h->wall_on
    && h->L.h[2] == 1 && decomposed(h, 2);
}

static int zhalo_unpack(b200lb_s * h, int w, cudaStream_t st) {
  if (!h->zpacked[w]) return 0;
  const int mode = h->zpacked[w];
  h->zpacked[w] = 0;
  if (!h->lat[w]) return 0;
  i64 zsx, zsy, zsp; int zvec;
  zpack_geom(h, mode, &zsx, &zsy, &zsp, &zvec);
  const Layout & L = h->L;
  ZPackTab T;
  memset(&T, 0, sizeof(T));
  int n[2] = {0, 0};
  for (int p = 0; p < h->ops->nvel; p++) {
    const int c2 = h->ops->cv[p][h->ops->dax[2]];
    if (c2 ==  1) T.pidx[0][n[0]++] = p;
    if (c2 == -1) T.pidx[1][n[1]++] = p;
  }
  T.np = n[0];
  // the apron of a direction holds pushed values only when that direction is decomposed
  const int x0 = decomposed(h, 0) ? 0 : L.h[0], x1 = L.h[0] + L.n[0] + (decomposed(h, 0) ? L.h[0] : 0);
  const int y0 = decomposed(h, 1) ? 0 : L.h[1], y1 = L.h[1] + L.n[1] + (decomposed(h, 1) ? L.h[1] : 0);
  dim3 block(256, 1, 1);
  dim3 grid((y1 - y0 + 255)/256, x1 - x0, 2*T.np);
  zhalo_unpack_kernel<<<grid, block, 0, st>>>(L, T, h->lat[w], (i64) h->nplane*L.nsd, zsx, zsy, zsp, x0, x1, y0, y1);
  CU(cudaGetLastError());
  h->launches++;
  return 0;
}

static int zhalo_flush(b200lb_s * h, cudaStream_t st) {
  for (int w = 0; w < 2; w++) { if (zhalo_unpack(h, w, st)) return -1; }
  return 0;
}

static int run_sweep(b200lb_s * h, bool pull, bool collide, const double * src,
		     double * dst, int aa = 0) {
  NvtxRange nvtx_(collide ? (pull ? "b200lb: Collision + Propagation (fused sweep)" : "b200lb: Collision (krnl)")
		  : "b200lb: Propagtn (krnl)");
  if (join_comm(h)) return -1;
  SweepArgs a;
  site_kparams(h, a.P);
  a.P.src = src; a.P.dst = dst;
  a.pull = pull; a.collide = collide;
  a.aa = aa;
  a.binary = (h->desc.ndist == 2);
  a.ghost = h->ghost;
  a.noise = collide && h->noise_on;
  a.visc = false;
  // the flux of the order parameter left by phi_from_g stands in for the pulled g
  a.gmom = pull && collide && a.binary && h->prop_pending && h->jphi_valid;
  a.P.jphi = a.gmom ? h->jphi : nullptr;
  h->jphi_valid = false;           // this sweep consumes or replaces the lattice it described
  if (a.noise) {
    if (h->ops->nvel - h->ops->nhydro > B200LB_MAXNOISE)
      return fail("isothermal fluctuations: the site generator yields %d variates per draw, nvel %d has %d ghost modes (src/noise.h:18)",
		  B200LB_MAXNOISE, h->ops->nvel, h->ops->nvel - h->ops->nhydro);
    if (!h->noise) return fail("isothermal fluctuations are on but noise_init was never called");
    if (h->prm.kt < 0.0) return fail("negative temperature");
  }
  // rho,u of this collision: stored by it (32 B/site, src/collision.c:570-579), or left to
  // ensure_hydro. Auto: stored when somebody asked for hydro quantities since the previous
  // collision -- a caller that looks every step gets the reference's behaviour, a time loop
  // that only looks at statistics steps does not pay for the store in between.
  const bool hstore = collide && (h->opt_hydro_write == 1 || (h->opt_hydro_write == 2 && h->hydro_asked));
  a.vec = h->opt_vec;
  a.bx = h->opt_block;
  a.bt = h->opt_threads;
  a.stream = h->s_main;
  if (collide) {
    for (int i = 0; i < 3; i++) a.P.fg[i] = h->prm.force_global[i] + (h->force_uniform ? h->fconst[i] : 0.0);
    if (!h->force_uniform) { if (need_planes(h, &h->force, 3)) return -1; }
    if (hstore) { if (need_hydro(h)) return -1; }
    if (h->desc.ndist == 2 && h->grad_state == 0) {
      if (need_planes(h, &h->grad, 3) || need_planes(h, &h->delsq, 1)) return -1;
    }
    if (h->visc_on) {
      if (a.binary) return fail("lb_collide: the viscosity model belongs to the single-fluid collision (src/collision.c:386-404)");
      if (h->noise_on) return fail("lb_collide: a viscosity model with isothermal fluctuations is not on this path");
      if (need_planes(h, &h->eta, 1)) return -1;
      a.P.eta = h->eta;
      a.visc = true;
      a.P.rho0cs2 = h->prm.rho0*(1.0/3.0);
      a.P.eta_ratio = h->prm.eta_bulk/h->prm.eta_shear;
      a.P.visc_relax = h->prm.nrelax;
    }
    a.P.force = h->force_uniform ? nullptr : h->force;
    h->last_ffield = !h->force_uniform;
    for (int i = 0; i < 3; i++) h->last_fg[i] = a.P.fg[i];
    if (hstore) { a.P.rho = h->rho; a.P.u = h->u; }
    a.P.phi = h->phi;
    if (h->desc.ndist == 2 && h->grad_state == 2) {
      return fail("lb_collide: the order parameter changed after the last field_grad_compute");
    }
    if (h->grad_state == 0) { a.P.grad = h->grad; a.P.delsq = h->delsq; }
  }
  int lo[7][3], hi[7][3];
  bool iswall[7] = {false, false, false, false, false, false, false};
  // Lees-Edwards planes rewrite post-collision populations (also on the y and z faces
  // of the rank) after this sweep: nothing is pushed to the neighbours ahead of that
  const bool push = collide && h->p2p_ready && h->opt_p2p && h->le_nplanes == 0;
  const bool walls = pull && h->prop_wall && h->wall_on;
  if (walls && a.binary && collide) return fail("internal: walls with the binary fluid are folded into the propagation, not into the collision");
  // packed z-face halo: written by this sweep / read by this sweep
  const int wsrc = (src == h->lat[0]) ? 0 : ((src == h->lat[1]) ? 1 : -1);
  const int wdst = (dst == h->lat[0]) ? 0 : ((dst == h->lat[1]) ? 1 : -1);
  const bool zuse = push && aa == 0 && wdst >= 0 && zpack_usable(h);
  if (pull && wsrc >= 0 && h->zpacked[wsrc]) {
    if (zuse && h->zpacked[wsrc] == h->opt_zpack) a.P.zread = 1;
    else if (zhalo_unpack(h, wsrc, h->s_main)) return -1;
  }
  if (zuse) {
    a.P.zoff = (i64) h->nplane*h->L.nsd;
    a.P.zwrite = 1;
    zpack_geom(h, h->opt_zpack, &a.P.zsx, &a.P.zsy, &a.P.zsp, &a.P.zvec);
  }
  const int nbox = walls ? wall_boxes(h, a.P, lo, hi, iswall) : (push ? 1 : split_boxes(h, lo, hi));
  const int vec0 = a.vec;
  h->bnd_valid = false;
  h->pushed_valid = false;
  a.pdl = false;
  if (push) {
    const int which = (dst == h->lat[0]) ? 0 : 1;
    a.P.push = 1;
    // the sweep may start under the synchronisation kernel that precedes it (see
    // lb_sweep_kernel): only when that kernel really is the previous thing in the stream
    a.pdl = h->opt_pdl && h->launches == h->launches_at_sync && !h->tev_on;
    {
      // whole-sector stores of the row-end halo sites (site_kernels.cuh): low halo = last
      // word of an aligned group of four, high halo = first word, everything else in
      // those groups padding of this row
      const Layout & L = h->L;
      const i64 lo = L.z0 + L.h[2] - 1, hi = L.z0 + L.h[2] + L.n[2];
      a.P.zsec = (L.h[2] == 1 && lo >= 3 && (lo + 1) % 4 == 0 && hi % 4 == 0 && hi + 3 < L.s1
		  && L.s1 % 4 == 0 && L.s0 % 4 == 0 && L.nsd % 4 == 0) ? 1 : 0;
    }
    for (int k = 0; k < 27; k++) a.P.peer[k] = h->peer_lat[which][k];
  }
  a.diag = collide && h->opt_diag;
  a.P.diag = h->diag;
  a.P.diag_base = 0;
  if (collide) h->diag_count = 0;
  for (int b = 0; b < nbox; b++) {
    a.wall = walls && iswall[b];
    a.vec = a.wall ? 1 : vec0;
    if (nbox > 1 || walls) {
      for (int e = 0; e < 3; e++) { a.P.lo[e] = lo[b][e]; a.P.hi[e] = hi[b][e]; }
      if (b == nbox - 1 && !walls) {
	// everything a neighbour needs is written: the exchange may start now
	CU(cudaEventRecord(h->ev_bnd, h->s_main));
	h->bnd_valid = true;
      }
    }
    const bool timed = h->tev_on && h->tev_n + 2 <= h->tev_cap;
    if (timed) CU(cudaEventRecord(h->tev[h->tev_n], h->s_main));
    CU(h->ops->sweep(a));
    if (timed) { CU(cudaEventRecord(h->tev[h->tev_n + 1], h->s_main)); h->tev_n += 2; }
    h->launches++;
    if (a.diag) {
      a.P.diag_base += a.nblocks;
      if (a.P.diag_base > h->diag_cap) return fail("internal: diagnostics buffer too small");
      h->diag_count = a.P.diag_base;
    }
  }
  if (push) {
    h->push_step++;
    p2p_sync_kernel<<<1, 32, 0, h->s_main>>>(h->p2p, h->push_step, P2P_WATCHDOG_NS, 1);
    CU(cudaGetLastError());
    h->launches++;
    h->launches_at_sync = h->launches;
    h->pushed_valid = true;
  }
  if (pull) h->prop_wall = false;  // the bounce-back belonged to this propagation
  if (wdst >= 0) h->zpacked[wdst] = zuse ? h->opt_zpack : 0;
  h->halo_deferred = false;        // whatever was written has stale halos until the next lb_halo
  if (collide) {
    if (hstore) { h->u_fill_pending = false; h->hydro_stale = false; }
    else h->hydro_stale = true;
    h->hydro_asked = false;
  }
  return 0;
}

// rho,u that the last collision did not store: recompute them from its
// post-collision populations, which must still be unstreamed in lat[cur]
static int ensure_hydro(b200lb_s * h) {
  h->hydro_asked = true;
  if (!h->hydro_stale) return 0;
  if (join_comm(h)) return -1;
  if (need_hydro(h)) return -1;
  CU(h->ops->hydro(h->L, h->lat[h->cur], h->desc.ndist, h->last_ffield ? h->force : nullptr,
		   h->last_fg, h->rho, h->u, h->fstar_mode, h->selfper, h->s_main));
  h->launches++;
  h->hydro_stale = false;
  h->u_fill_pending = false;
  return 0;
}

// apply a deferred propagation so that lat[cur] is the streamed lattice
static int aa_materialise(b200lb_s * h);

static int materialise(b200lb_s * h) {
  if (h->opt_lattices == 1) return aa_materialise(h);
  if (!h->prop_pending) return 0;
  if (ensure_hydro(h)) return -1;       // last chance: streaming scatters f*
  if (run_sweep(h, true, false, h->lat[h->cur], h->lat[1 - h->cur])) return -1;
  h->cur = 1 - h->cur;
  h->prop_pending = false;
  return 0;
}

// ---------------------------------------------------------------------------
// Single-lattice streaming behind the same API (option lattices = 1). lb_collide
// alternates between the even sweep (collide in place, slots swapped) and the odd sweep
// (pull from the swapped slots, collide, store one propagation ahead); lb_propagation is
// recorded after an even sweep and already done after an odd one. What the reference
// shows at every API boundary is reproduced for the time loop's sequence
// collide - halo - propagate; an observer between an even and an odd sweep gets the
// natural, streamed lattice through a scratch lattice that lives for that call only.
// ---------------------------------------------------------------------------

static int aa_supported(const b200lb_s * h) {
  if (h->desc.ndist != 1) return fail("lattices = 1: the single-fluid distribution (ndist = 1)");
  if (h->wall_on || h->le_nplanes || h->noise_on || h->visc_on || h->open_on || h->ch_on)
    return fail("lattices = 1: walls, Lees-Edwards planes, fluctuations, viscosity model, open boundaries and the finite-difference order parameter keep two lattices");
  for (int d = 0; d < 3; d++) {
    if (h->L.h[d] == 0 || h->selfper[d]) continue;
    // across ranks the odd sweep stores into the neighbours' interior: peer memory only
    if (!decomposed(h, d)) return fail("lattices = 1: every direction periodic");
    if (!(h->p2p_ready && h->opt_p2p)) return fail("lattices = 1 on a decomposed lattice: with the peer-memory halo only (b200lb_p2p_import)");
    if (h->L.n[d] < 2) return fail("lattices = 1: at least two sites per decomposed direction");
  }
  return 0;
}

static int aa_materialise(b200lb_s * h) {
  if (h->aa_ahead) return fail("lattices = 1: the lattice is looked at between lb_collide and lb_propagation of an odd step; finish the step first");
  if (!h->prop_pending) return 0;
  if (!h->aa_swapped) return fail("internal: single-lattice state");
  if (ensure_hydro(h)) return -1;
  // natural order, streamed: pulled out of the swapped slots into a scratch lattice, which
  // then IS the lattice (the array it replaces is released): two lattices exist for the
  // duration of this call only
  double * scratch = nullptr;
  const size_t nb = plane_bytes(h)*h->nplane;
  if (cudaMalloc(&scratch, nb) != cudaSuccess) {
    cudaGetLastError();
    return fail("lattices = 1: looking at the lattice between two sweeps needs room for a second lattice for the duration of the call (%zu bytes)", nb);
  }
  if (run_sweep(h, true, false, h->lat[0], scratch, 3)) { cudaFree(scratch); return -1; }
  // back into the array the neighbours have mapped (halo sites included: they keep what the
  // even sweeps pushed, stale for the streamed lattice as after any propagation)
  {
    const Layout & L = h->L;
    for (int e = 0; e < h->nplane; e++) {
      cudaMemcpy3DParms c;
      memset(&c, 0, sizeof(c));
      const size_t off = (size_t) e*L.nsd + (size_t) L.h[0]*L.s0 + (size_t) L.h[1]*L.s1 + L.z0 + L.h[2];
      c.srcPtr = make_cudaPitchedPtr(scratch + off, L.s1*sizeof(double), L.s1, L.s0/L.s1);
      c.dstPtr = make_cudaPitchedPtr(h->lat[0] + off, L.s1*sizeof(double), L.s1, L.s0/L.s1);
      c.extent = make_cudaExtent((size_t) L.n[2]*sizeof(double), L.n[1], L.n[0]);
      c.kind = cudaMemcpyDeviceToDevice;
      CU(cudaMemcpy3DAsync(&c, h->s_main));
    }
  }
  CU(cudaStreamSynchronize(h->s_main));
  CU(cudaFree(scratch));
  h->prop_pending = false;
  h->aa_swapped = false;
  h->fstar_mode = 0;
  return 0;
}

// An observer of the lattice proper (lb_memcpy D2H, statistics, file records): the states
// the time loop passes through between lb_collide and lb_propagation are not presentable
static int aa_observe(b200lb_s * h) {
  if (h->opt_lattices != 1) return 0;
  if (h->aa_ahead || (h->aa_swapped && !h->prop_pending))
    return fail("lattices = 1: the lattice is looked at between lb_collide and lb_propagation; with a single lattice it is presentable after lb_propagation only");
  return 0;
}

static int aa_collide(b200lb_s * h) {
  if (aa_supported(h)) return -1;
  if (h->aa_ahead) return fail("lattices = 1: lb_collide twice without lb_propagation");
  if (h->prop_pending) {
    // odd sweep: the recorded propagation, the collision, and the next propagation
    if (run_sweep(h, true, true, h->lat[0], h->lat[0], 2)) return -1;
    h->prop_pending = false;
    h->aa_swapped = false;
    h->aa_ahead = true;
    h->fstar_mode = 2;
    return 0;
  }
  if (h->aa_swapped) return fail("lattices = 1: lb_collide twice without lb_propagation");
  if (run_sweep(h, false, true, h->lat[0], h->lat[0], 1)) return -1;
  h->aa_swapped = true;
  h->fstar_mode = 1;
  return 0;
}

extern "C" int b200lb_collide(b200lb_t * h) {
  if (!h) return fail("null handle");
  if (h->opt_lattices == 1) return aa_collide(h);
  if (resolve_ch_force(h)) return -1;     // phi_force_calculation without a Cahn-Hilliard update after it
  if (h->prop_pending && h->prop_wall && h->desc.ndist == 2) { if (materialise(h)) return -1; }
  if (h->prop_pending) {
    // stream-then-collide in one pass over the lattice
    if (run_sweep(h, true, true, h->lat[h->cur], h->lat[1 - h->cur])) return -1;
    h->cur = 1 - h->cur;
    h->prop_pending = false;
    return 0;
  }
  return run_sweep(h, false, true, h->lat[h->cur], h->lat[h->cur]);
}

extern "C" int b200lb_propagate(b200lb_t * h) {
  if (!h) return fail("null handle");
  if (h->opt_lattices == 1) {
    if (h->aa_ahead) { h->aa_ahead = false; return 0; }          // the odd sweep has done it
    if (h->aa_swapped && !h->prop_pending) { h->prop_pending = true; return 0; }
    return fail("lattices = 1: lb_propagation follows lb_collide (a propagation on its own needs the second lattice)");
  }
  if (h->prop_pending) { if (materialise(h)) return -1; }
  h->prop_wall = h->bbl_armed;      // wall_bbl ran since the collision: this streaming bounces
  h->bbl_armed = false;
  if (h->opt_fused) {
    // Lazy: the streaming is folded into the next collision (or done on demand
    // by the first call that observes the lattice).
    h->prop_pending = true;
    h->jphi_valid = false;
    return 0;
  }
  if (ensure_hydro(h)) return -1;
  if (run_sweep(h, true, false, h->lat[h->cur], h->lat[1 - h->cur])) return -1;
  h->cur = 1 - h->cur;
  return 0;
}

// ---------------------------------------------------------------------------
// flat walls: wall_create + wall_commit, wall_bbl, wall_momentum (src/wall.c)
// ---------------------------------------------------------------------------

extern "C" int b200lb_wall_commit(b200lb_t * h, const int isboundary[3], const double ubot[3],
				  const double utop[3]) {
  if (!h || !isboundary) return fail("null argument");
  if (h->desc.ndist != 1 && !h->opt_binary_walls) return fail("walls with the binary fluid: switched off by option binary_walls = 0");
  bool any = false;
  for (int a = 0; a < 3; a++) {
    if (!isboundary[a]) continue;
    any = true;
    if (h->desc.periodic[a]) return fail("boundary_walls: direction %d is periodic", a);
    if (a >= h->desc.ndim) return fail("boundary_walls: direction %d of a %d-d lattice", a, h->desc.ndim);
  }
  // src/wall_rt.c:56-60: a wall that moves along x cannot be an x wall
  if (isboundary[0] && ((ubot && ubot[0] != 0.0) || (utop && utop[0] != 0.0)))
    return fail("cannot have non-zero u_x wall velocity if the wall is in x");
  if (h->prop_pending) { if (materialise(h)) return -1; }
  for (int a = 0; a < 3; a++) {
    h->wall_is[a] = isboundary[a] ? 1 : 0;
    h->wall_ubot[a] = ubot ? ubot[a] : 0.0;
    h->wall_utop[a] = utop ? utop[a] : 0.0;
    h->fnet_host[a] = 0.0;
  }
  h->wall_on = any;
  h->wall_slip = 0;
  for (int a = 0; a < 3; a++) { h->wall_sbot[a] = 0.0; h->wall_stop[a] = 0.0; }
  h->bbl_armed = false;
  h->prop_wall = false;
  if (!h->fnet_dev) CU(cudaMalloc(&h->fnet_dev, 3*sizeof(double)));
  CU(cudaMemsetAsync(h->fnet_dev, 0, 3*sizeof(double), h->s_main));
  return 0;
}

// wall_slip() + param.slip (src/wall.c:284-343, src/wall_rt.c:86-104): fraction of
// specular reflection per wall, bottom and top, after b200lb_wall_commit
extern "C" int b200lb_wall_slip_set(b200lb_t * h, const double sbot[3], const double stop[3],
				    int force_active) {
  if (!h || !sbot || !stop) return fail("null argument");
  if (!h->wall_on) return fail("wall slip: no walls committed");
  bool active = (force_active != 0);
  for (int a = 0; a < 3; a++) {
    if (sbot[a] < 0.0 || sbot[a] > 1.0 || stop[a] < 0.0 || stop[a] > 1.0)
      return fail("wall slip parameters must be 0 <= s <= 1 everywhere");
    if (sbot[a] != 0.0 || stop[a] != 0.0) active = true;
  }
  if (active) {
    for (int a = 0; a < 3; a++) {
      if (h->wall_ubot[a] != 0.0 || h->wall_utop[a] != 0.0)
	return fail("wall slip: the slip bounce-back does not allow a wall velocity (src/wall.c:1114)");
    }
  }
  if (h->prop_pending) { if (materialise(h)) return -1; }
  h->wall_slip = active ? 1 : 0;
  for (int a = 0; a < 3; a++) { h->wall_sbot[a] = sbot[a]; h->wall_stop[a] = stop[a]; }
  return 0;
}

// wall_bbl (src/wall.c:960-991), called between lb_halo and lb_propagation
// (src/ludwig.c:842). The wall's share of the momentum is tallied now, from the
// post-collision populations of the sites on the wall faces; the reversed populations
// themselves are not written into the wall layer: the propagation that follows (fused
// into the next collision, or on its own) takes them from the site they left.
// ---------------------------------------------------------------------------
// open boundaries for the fluid: inflow / outflow "rhou" (open_bc_kernels.cuh)
// ---------------------------------------------------------------------------

extern "C" int b200lb_open_bc_commit(b200lb_t * h, int flow_dir, const double u0[3], double rho0) {
  if (!h) return fail("null handle");
  if (flow_dir < 0) { h->open_on = false; return 0; }
  if (flow_dir > 2 || !u0) return fail("lb_bc_open: flow direction 0, 1 or 2 and an inflow velocity");
  if (h->desc.ndim != 3) return fail("lb_bc_open: three-dimensional models on this path");
  if (h->desc.ndist != 1) return fail("lb_bc_open: the single-fluid distribution (the order parameter has its own open boundaries, src/phi_bc_open.h: not on this path)");
  if (h->desc.periodic[flow_dir]) return fail("lb_bc_open: the flow direction %d is periodic", flow_dir);
  for (int a = 0; a < 3; a++) {
    if (a == flow_dir) continue;
    // src/lb_bc_open_rt.c:65-67: the flow runs along the one direction without walls
    if (!h->wall_on || !h->wall_is[a]) return fail("lb_bc_open: walls in the two directions across the flow (boundary_walls) come first");
    if (h->desc.cartsz[a] > 1) return fail("lb_bc_open: a decomposition across the flow (src/lb_bc_inflow_rhou.c:293, the exchange of the inlet's rho, u) is not on this path");
  }
  if (h->wall_on && h->wall_is[flow_dir]) return fail("lb_bc_open: a wall in the flow direction");
  if (!(rho0 > 0.0)) return fail("lb_bc_outflow_rhou_rho0 must be positive");
  h->open_on = true;
  h->open_id = flow_dir;
  for (int a = 0; a < 3; a++) h->open_u0[a] = u0[a];
  h->open_rho0 = rho0;
  return 0;
}

// inflow->update + impose, outflow->update + impose: src/ludwig.c:820-831
extern "C" int b200lb_open_bc_apply(b200lb_t * h) {
  if (!h) return fail("null handle");
  if (!h->open_on) return 0;
  if (h->prop_pending) return fail("lb_bc_open: called after lb_propagation (it belongs between lb_halo and lb_propagation)");
  if (ensure_hydro(h)) return -1;                  // rho, u of the collision just done
  if (need_hydro(h)) return -1;
  if (join_comm(h)) return -1;
  if (zhalo_flush(h, h->s_main)) return -1;
  OpenTab T;
  memset(&T, 0, sizeof(T));
  T.L = h->L;
  T.nvel = h->ops->nvel;
  T.id = h->open_id;
  T.da = (T.id == 0) ? 1 : 0;
  T.db = (T.id == 2) ? 1 : 2;
  T.has_in = (h->desc.cartcoords[T.id] == 0);
  T.has_out = (h->desc.cartcoords[T.id] == h->desc.cartsz[T.id] - 1);
  for (int a = 0; a < 3; a++) { T.lo[a] = 0; T.hi[a] = h->L.n[a] - 1; T.u0[a] = h->open_u0[a]; }
  T.rho0 = h->open_rho0;
  for (int p = 0; p < T.nvel; p++) {
    for (int a = 0; a < 3; a++) T.cv[p][a] = (signed char) h->ops->cv[p][a];
    T.wv[p] = h->ops->wv[p];
  }
  dim3 block(128, 1, 1);
  dim3 grid((h->L.n[T.db] + 127)/128, h->L.n[T.da], 2);
  open_bc_kernel<<<grid, block, 0, h->s_main>>>(T, h->rho, h->u, h->lat[h->cur]);
  CU(cudaGetLastError());
  h->launches++;
  return 0;
}

extern "C" int b200lb_wall_bbl(b200lb_t * h) {
  if (!h) return fail("null handle");
  if (!h->wall_on) return 0;                       // wall->nlink == 0: nothing to do
  if (h->prop_pending) return fail("wall_bbl: called after lb_propagation (it belongs between lb_halo and lb_propagation)");
  if (join_comm(h)) return -1;
  KParams P;
  site_kparams(h, P);
  WallTab W;
  memset(&W, 0, sizeof(W));
  W.L = h->L;
  W.nvel = h->ops->nvel;
  for (int d = 0; d < 3; d++) { W.wall_lo[d] = P.wall_lo[d]; W.wall_hi[d] = P.wall_hi[d]; }
  for (int p = 0; p < W.nvel; p++) {
    for (int d = 0; d < 3; d++) W.dcv[p][d] = (signed char) h->ops->cv[p][h->ops->dax[d]];
    for (int a = 0; a < 3; a++) W.pcv[p][a] = (signed char) h->ops->cv[p][a];
    W.wv[p] = h->ops->wv[p];
    // x_p of the outgoing link p is minus the correction of the incoming Q - p
    W.xout[p] = (p > 0) ? -P.wall_corr[W.nvel - p] : 0.0;
    for (int mask = 0; mask < 8; mask++) W.qflip[p][mask] = P.wall_q[p][mask];
  }
  W.slip = P.wall_slip;
  for (int d = 0; d < 3; d++) { W.s_lo[d] = P.wall_s_lo[d]; W.s_hi[d] = P.wall_s_hi[d]; W.wrap[d] = h->selfper[d]; }
  int nb = 1, na = 1;
  for (int d = 0; d < 3; d++) {
    for (int side = 0; side < 2; side++) {
      if (!(side == 0 ? W.wall_lo[d] : W.wall_hi[d])) continue;
      const int da = (d == 0) ? 1 : 0, db = (d == 2) ? 1 : 2;
      W.face_axis[W.nface] = d; W.face_hi[W.nface] = side;
      W.nface++;
      if (h->L.n[db] > nb) nb = h->L.n[db];
      if (h->L.n[da] > na) na = h->L.n[da];
    }
  }
  if (W.nface > 0) {
    dim3 block(128, 1, 1);
    dim3 grid((nb + 127)/128, na, W.nface);
    wall_fnet_kernel<<<grid, block, 0, h->s_main>>>(W, h->lat[h->cur], h->fnet_dev);
    CU(cudaGetLastError());
    h->launches++;
  }
  h->bbl_armed = true;
  return 0;
}

// wall_momentum (src/wall.c:1299-1340): this rank's share; the device tally is moved
// to the host total and zeroed, as there. wall_momentum_add: b200lb_wall_momentum_add.
extern "C" int b200lb_wall_momentum(b200lb_t * h, double f[3]) {
  if (!h || !f) return fail("null argument");
  if (h->fnet_dev) {
    double tmp[3];
    if (join_comm(h)) return -1;
    CU(cudaMemcpyAsync(tmp, h->fnet_dev, 3*sizeof(double), cudaMemcpyDeviceToHost, h->s_main));
    CU(cudaMemsetAsync(h->fnet_dev, 0, 3*sizeof(double), h->s_main));
    CU(cudaStreamSynchronize(h->s_main));
    for (int a = 0; a < 3; a++) h->fnet_host[a] += tmp[a];
  }
  for (int a = 0; a < 3; a++) f[a] = h->fnet_host[a];
  return 0;
}

extern "C" int b200lb_wall_momentum_add(b200lb_t * h, const double f[3]) {
  if (!h || !f) return fail("null argument");
  for (int a = 0; a < 3; a++) h->fnet_host[a] += f[a];
  return 0;
}

// ---------------------------------------------------------------------------
// Lees-Edwards planes (src/leesedwards.c, src/model_le.c)
// ---------------------------------------------------------------------------

static int le_location(const b200lb_s * h, int np) {
  // src/leesedwards.c:599-617 with the uniform decomposition's offset
  const int ntotal = h->desc.cartsz[0]*h->desc.nlocal[0];
  const double dx_sep = 1.0*ntotal/h->le_nplanes;
  const double dx_min = 0.5*dx_sep;
  const int offset = h->desc.cartcoords[0]*h->desc.nlocal[0];
  return (int) (dx_min + (np + h->desc.cartcoords[0]*h->le_nlocal)*dx_sep - offset);
}

extern "C" int b200lb_le_commit(b200lb_t * h, int nplanes, double uy, int oscillatory,
				int period, int nt0) {
  if (!h) return fail("null handle");
  if (nplanes < 0) return fail("N_LE_plane must not be negative");
  if (nplanes > 0 && h->ch_on && !h->opt_le_fd) return fail("Lees-Edwards planes for the finite-difference order parameter (src/phi_cahn_hilliard.c:618): switched off by option le_fd = 0");
  if (h->prop_pending) { if (materialise(h)) return -1; }
  h->le_nplanes = 0; h->le_nlocal = 0; h->le_step = 0; h->le_tstart = 0; h->le_field_t = 0.0;
  if (h->le_phibuf) { CU(cudaFree(h->le_phibuf)); h->le_phibuf = nullptr; }
  if (h->le_chbuf) { CU(cudaFree(h->le_chbuf)); h->le_chbuf = nullptr; }
  if (nplanes == 0) return 0;
  const int ntotal = h->desc.cartsz[0]*h->desc.nlocal[0];
  if (ntotal % nplanes) return fail("Number of planes must divide system size");
  if (nplanes % h->desc.cartsz[0]) return fail("Must have a uniform number of planes per process");
  if (oscillatory && period <= 0) return fail("Oscillatory shear must have period > 0");
  if (!h->desc.periodic[0] || !h->desc.periodic[1])
    return fail("Lees-Edwards planes need periodic x and y");
  if (h->desc.cartsz[1] > 1)
    return fail("Lees-Edwards planes with a decomposition in y (src/model_le.c:399) are not on this path");
  if (nplanes/h->desc.cartsz[0] > B200LB_LE_MAXPLANE)
    return fail("more than %d Lees-Edwards planes per rank", B200LB_LE_MAXPLANE);
  h->le_nplanes = nplanes; h->le_nlocal = nplanes/h->desc.cartsz[0];
  h->le_uy = uy; h->le_osc = (oscillatory != 0); h->le_period = period; h->le_nt0 = nt0;
  for (int np = 0; np < h->le_nlocal; np++) {
    const int ic = le_location(h, np);
    if (ic <= h->desc.nhalo || ic > h->desc.nlocal[0] - h->desc.nhalo) {
      h->le_nplanes = 0; h->le_nlocal = 0;
      return fail("Wall at domain boundary");        // sic, src/leesedwards.c:443
    }
  }
  // crossing populations: those with c_x = +1 below each plane, -1 above it
  int nprop = 0;
  for (int p = 1; p < h->ops->nvel; p++) nprop += (h->ops->cv[p][0] == 1);
  const size_t nbuf = (size_t) 2*h->le_nlocal*h->desc.ndist*nprop*h->desc.nlocal[1]*h->desc.nlocal[2];
  if (h->le_buf) { CU(cudaFree(h->le_buf)); h->le_buf = nullptr; }
  CU(cudaMalloc(&h->le_buf, nbuf*sizeof(double)));
  return 0;
}

extern "C" int b200lb_le_plane_location(const b200lb_t * h, int np) {
  if (!h || h->le_nlocal == 0 || np < 0 || np >= h->le_nlocal) { fail("no such Lees-Edwards plane"); return -1; }
  return le_location(h, np);
}

extern "C" int b200lb_le_time_set(b200lb_t * h, int nstep) {
  if (!h) return fail("null handle");
  h->le_step = nstep;
  h->le_tstart = nstep;
  h->le_field_t = 1.0*(h->le_tstart + h->le_step - 1.0);    // src/physics.c:651-659
  return 0;
}

// physics_control_time for the field_leesedwards inside the next field_grad_compute
// (src/field.c:553). b200lb_step keeps it itself.
extern "C" int b200lb_le_field_time_set(b200lb_t * h, double t) {
  if (!h) return fail("null handle");
  h->le_field_t = t;
  return 0;
}

extern "C" int b200lb_le_apply(b200lb_t * h, double t) {
  if (!h) return fail("null handle");
  if (h->le_nlocal == 0) return 0;                 // lees_edw_nplane_local(le) == 0
  if (h->prop_pending) return fail("lb_le_apply_boundary_conditions: called after lb_propagation (it belongs between lb_collide and lb_halo)");
  if (ensure_hydro(h)) return -1;                  // rho, u of the collision, before its populations change
  if (join_comm(h)) return -1;
  if (zhalo_flush(h, h->s_main)) return -1;
  // velocity jump and displacement now: src/leesedwards.c:572-587, 633-655
  const double tle = ((t < 0.0) ? 0.0 : t) - 1.0*h->le_nt0;
  double uy_now = h->le_uy, dy_now = tle*h->le_uy;
  if (h->le_osc) {
    const double omega = 2.0*4.0*atan(1.0)/h->le_period;
    uy_now = h->le_uy*cos(omega*(t - 1.0*h->le_nt0));
    dy_now = h->le_uy*sin(omega*tle)/omega;
  }
  LeTab T;
  memset(&T, 0, sizeof(T));
  T.L = h->L;
  T.nvel = h->ops->nvel; T.ndist = h->desc.ndist;
  for (int p = 0; p < T.nvel; p++) {
    for (int a = 0; a < 3; a++) T.pcv[p][a] = (signed char) h->ops->cv[p][a];
    T.wv[p] = h->ops->wv[p];
    if (p > 0 && h->ops->cv[p][0] == 1) T.nprop++;
  }
  for (int d = 0; d < 3; d++) {
    if (h->ops->dax[d] == 0) T.dx = d;
    if (h->ops->dax[d] == 1) T.dy = d;
  }
  T.da = (T.dx == 0) ? 1 : 0;
  T.db = (T.dx == 2) ? 1 : 2;
  const int ny = h->desc.nlocal[1];
  T.nside = 2*h->le_nlocal;
  for (int np = 0; np < h->le_nlocal; np++) {
    for (int side = 0; side < 2; side++) {
      const int k = 2*np + side;
      T.ic[k] = le_location(h, np) + side - 1;     // 0-based interior coordinate
      T.cx[k] = (side == 0) ? +1 : -1;
      T.duy[k] = (side == 0) ? -1.0*uy_now : uy_now;
      const double dy = fmod((side == 0) ? dy_now : -dy_now, 1.0*ny);
      const int jdy = (int) floor(dy);
      T.jdy[k] = jdy;
      T.fr[k] = dy - jdy;
    }
  }
  dim3 block(128, 1, 1);
  dim3 grid((h->L.n[T.db] + 127)/128, h->L.n[T.da], T.nside);
  le_reproject_kernel<<<grid, block, 0, h->s_main>>>(T, h->lat[h->cur], h->le_buf);
  CU(cudaGetLastError());
  le_interp_kernel<<<grid, block, 0, h->s_main>>>(T, h->le_buf, h->lat[h->cur]);
  CU(cudaGetLastError());
  h->launches += 2;
  // the boundary slabs changed after the sweep recorded them: a halo exchange waits for this
  h->bnd_valid = false;
  h->pushed_valid = false;
  h->halo_deferred = false;
  return 0;
}

// ---------------------------------------------------------------------------
// halo exchange
// ---------------------------------------------------------------------------

// element lists for device axis d: which planes are needed in the low / high halo
static void halo_lists(const b200lb_s * h, int d, int scheme, int nplane_per_dist,
		       bool lattice, HaloParams & H) {
  H.nlo = H.nhi = 0;
  const int nplane = lattice ? h->nplane : nplane_per_dist;
  for (int e = 0; e < nplane; e++) {
    int c = 0;
    if (lattice && scheme == B200LB_HALO_REDUCED) {
      const int p = e % h->desc.nvel;
      c = h->ops->cv[p][h->ops->dax[d]];
      // the low halo is read by sites pulling with c = +1, the high one with c = -1
      if (c == 1)  H.el_lo[H.nlo++] = (unsigned char) e;
      if (c == -1) H.el_hi[H.nhi++] = (unsigned char) e;
    }
    else {
      H.el_lo[H.nlo++] = (unsigned char) e;
      H.el_hi[H.nhi++] = (unsigned char) e;
    }
  }
}

static int halo_launch(b200lb_s * h, int mode, const HaloParams & H, cudaStream_t st) {
  const Layout & L = H.L;
  const int d = H.axis;
  const int fa = (d == 2) ? 1 : 2;
  const int sl = (d == 0) ? 1 : 0;
  dim3 block(fa == 2 ? 128 : 32, fa == 2 ? 2 : 8, 1);
  dim3 grid((L.na[fa] + block.x - 1)/block.x, (L.na[sl] + block.y - 1)/block.y, H.nlo + H.nhi);
  if (grid.z == 0) return 0;
  if (mode == HALO_LOCAL)  halo_kernel<HALO_LOCAL><<<grid, block, 0, st>>>(H);
  if (mode == HALO_PACK)   halo_kernel<HALO_PACK><<<grid, block, 0, st>>>(H);
  if (mode == HALO_UNPACK) halo_kernel<HALO_UNPACK><<<grid, block, 0, st>>>(H);
  CU(cudaGetLastError());
  h->launches++;
  return 0;
}

static int halo_buffers(b200lb_s * h, int d, int nel) {
  HaloBuf & b = h->hb[d];
  const int fa = (d == 2) ? 1 : 2;
  const int sl = (d == 0) ? 1 : 0;
  const i64 nface = (i64) h->L.na[fa]*h->L.na[sl];
  if (b.to_lo && b.nface == nface && h->hb_nel >= nel) return 0;
  const int want = nel > h->nplane ? nel : h->nplane;
  cudaFree(b.to_lo); cudaFree(b.to_hi); cudaFree(b.from_lo); cudaFree(b.from_hi);
  const size_t nb = (size_t) nface*want*sizeof(double);
  CU(cudaMalloc(&b.to_lo, nb)); CU(cudaMalloc(&b.to_hi, nb));
  CU(cudaMalloc(&b.from_lo, nb)); CU(cudaMalloc(&b.from_hi, nb));
  b.nface = nface;
  h->hb_nel = want;
  return 0;
}

// Exchange the halo planes of `data` (nplane planes of nsd doubles).
// which: 0 = every axis, 1 = only axes decomposed over ranks, 2 = only the
// axes that are periodic within this rank (local copies).
static int halo_exchange(b200lb_s * h, double * data, bool lattice, int nplane, int scheme,
			 int which, cudaStream_t st) {
  for (int d = 0; d < 3; d++) {
    if (h->L.h[d] == 0) continue;
    const int a = h->ops->dax[d];
    HaloParams H;
    memset(&H, 0, sizeof(H));
    H.L = h->L; H.data = data; H.axis = d;
    halo_lists(h, d, scheme, nplane, lattice, H);
    if (!decomposed(h, d)) {
      if (!h->desc.periodic[a]) continue;        // physical boundary: nothing to receive
      if (which == 1) continue;
      if (halo_launch(h, HALO_LOCAL, H, st)) return -1;
      continue;
    }
    if (which == 2) continue;
    if (!have_comm(h)) return fail("direction %d is decomposed over %d ranks but b200lb_comm_init was not called", a, h->desc.cartsz[a]);
    if (halo_buffers(h, d, nplane)) return -1;
    HaloBuf & b = h->hb[d];
    H.to_lo = b.to_lo; H.to_hi = b.to_hi; H.from_lo = b.from_lo; H.from_hi = b.from_hi;
    if (halo_launch(h, HALO_PACK, H, st)) return -1;
    const int lo = h->desc.nbr_lo[a], hi = h->desc.nbr_hi[a];
    // Order matters when lo == hi (two ranks, periodic): the peer's first
    // message is its low face, which is my HIGH halo.
    if (h->desc.cartsz[a] == 1) {
      // loopback: this rank is its own neighbour on both sides
      CU(cudaMemcpyAsync(b.from_lo, b.to_hi, sizeof(double)*H.nlo*b.nface, cudaMemcpyDeviceToDevice, st));
      CU(cudaMemcpyAsync(b.from_hi, b.to_lo, sizeof(double)*H.nhi*b.nface, cudaMemcpyDeviceToDevice, st));
    }
    else {
    NC(g_nccl.GroupStart());
    if (lo >= 0) NC(g_nccl.Send(b.to_lo, (size_t) H.nhi*b.nface, ncclDouble, lo, h->comm, st));
    if (hi >= 0) NC(g_nccl.Send(b.to_hi, (size_t) H.nlo*b.nface, ncclDouble, hi, h->comm, st));
    if (hi >= 0) NC(g_nccl.Recv(b.from_hi, (size_t) H.nhi*b.nface, ncclDouble, hi, h->comm, st));
    if (lo >= 0) NC(g_nccl.Recv(b.from_lo, (size_t) H.nlo*b.nface, ncclDouble, lo, h->comm, st));
    NC(g_nccl.GroupEnd());
    }
    // a missing neighbour (non-periodic edge) leaves that halo untouched
    if (lo < 0) H.nlo = 0;
    if (hi < 0) H.nhi = 0;
    if (halo_launch(h, HALO_UNPACK, H, st)) return -1;
  }
  return 0;
}


// ---------------------------------------------------------------------------
// direct reduced exchange (see halo_direct_kernel)
// ---------------------------------------------------------------------------

// rank of the process at Cartesian coordinates coords + dir (physical axes), in
// the row-major order of MPI_Cart_create without reordering (z fastest), or -1
static int cart_rank(const b200lb_s * h, const int dir[3]) {
  int c[3];
  for (int a = 0; a < 3; a++) {
    c[a] = h->desc.cartcoords[a] + dir[a];
    if (c[a] < 0 || c[a] >= h->desc.cartsz[a]) {
      if (!h->desc.periodic[a]) return -1;
      c[a] = (c[a] + h->desc.cartsz[a]) % h->desc.cartsz[a];
    }
  }
  return (c[0]*h->desc.cartsz[1] + c[1])*h->desc.cartsz[2] + c[2];
}

static int direct_setup(b200lb_s * h) {
  if (h->direct_state != 0) return 0;
  h->direct_state = -1;
  const Layout & L = h->L;
  // the neighbour table of the descriptor must agree with row-major ranks
  for (int a = 0; a < 3; a++) {
    if (h->desc.cartsz[a] == 1) continue;
    int dm[3] = {0, 0, 0}, dp[3] = {0, 0, 0};
    dm[a] = -1; dp[a] = 1;
    if (cart_rank(h, dm) != h->desc.nbr_lo[a] || cart_rank(h, dp) != h->desc.nbr_hi[a]) return 0;
  }
  HaloMsg pk[B200LB_MAXMSG], up[B200LB_MAXMSG];
  memset(pk, 0, sizeof(pk)); memset(up, 0, sizeof(up));
  int np = 0, nu = 0;
  i64 poff = 0, uoff = 0;
  h->pack_max = h->unpack_max = 0;
  for (int e0 = -1; e0 <= 1; e0++) {
    for (int e1 = -1; e1 <= 1; e1++) {
      for (int e2 = -1; e2 <= 1; e2++) {
	const int dir[3] = {e0, e1, e2};              // device axes
	if (!e0 && !e1 && !e2) continue;
	bool ok = true;
	int pdir[3] = {0, 0, 0}, mdir[3] = {0, 0, 0};  // physical axes
	for (int d = 0; d < 3; d++) {
	  const int a = h->ops->dax[d];
	  if (dir[d] != 0 && !decomposed(h, d)) ok = false;
	  pdir[a] = dir[d]; mdir[a] = -dir[d];
	}
	if (!ok) continue;
	HaloMsg m;
	memset(&m, 0, sizeof(m));
	for (int n = 0; n < h->desc.ndist; n++) {
	  for (int p = 0; p < h->desc.nvel; p++) {
	    bool match = true;
	    for (int d = 0; d < 3; d++) {
	      if (dir[d] != 0 && h->ops->cv[p][h->ops->dax[d]] != dir[d]) match = false;
	    }
	    if (match) {
	      if (m.npop >= B200LB_MSGPOP) return 0;
	      m.plane[m.npop++] = n*h->desc.nvel + p;
	    }
	  }
	}
	if (m.npop == 0) continue;                     // e.g. corners of D3Q19
	m.nsite = 1;
	for (int d = 0; d < 3; d++) {
	  if (dir[d] == 1)       { m.slo[d] = L.n[d];  m.rlo[d] = 0;          m.dim[d] = 1; }
	  else if (dir[d] == -1) { m.slo[d] = 1;       m.rlo[d] = L.n[d] + 1; m.dim[d] = 1; }
	  else                   { m.slo[d] = L.h[d];  m.rlo[d] = L.h[d];     m.dim[d] = L.n[d]; }
	  m.nsite *= m.dim[d];
	}
	const i64 cnt = m.nsite*m.npop;
	const int to = cart_rank(h, pdir), from = cart_rank(h, mdir);
	if (to >= 0) {
	  m.off = poff; pk[np] = m;
	  h->pack_peer[np] = to; h->pack_off[np] = poff; h->pack_cnt[np] = cnt;
	  poff += cnt; np++;
	  if (cnt > h->pack_max) h->pack_max = cnt;
	}
	if (from >= 0) {
	  m.off = uoff; up[nu] = m;
	  h->unpack_peer[nu] = from; h->unpack_off[nu] = uoff; h->unpack_cnt[nu] = cnt;
	  uoff += cnt; nu++;
	  if (cnt > h->unpack_max) h->unpack_max = cnt;
	}
      }
    }
  }
  if (np == 0 && nu == 0) return 0;
  h->npack = np; h->nunpack = nu;
  CU(cudaMalloc(&h->msg_pack, sizeof(HaloMsg)*B200LB_MAXMSG));
  CU(cudaMalloc(&h->msg_unpack, sizeof(HaloMsg)*B200LB_MAXMSG));
  CU(cudaMemcpy(h->msg_pack, pk, sizeof(HaloMsg)*B200LB_MAXMSG, cudaMemcpyHostToDevice));
  CU(cudaMemcpy(h->msg_unpack, up, sizeof(HaloMsg)*B200LB_MAXMSG, cudaMemcpyHostToDevice));
  CU(cudaMalloc(&h->dsend, sizeof(double)*(poff + 1)));
  CU(cudaMalloc(&h->drecv, sizeof(double)*(uoff + 1)));
  h->direct_state = 1;
  return 0;
}

static int direct_exchange(b200lb_s * h, double * data, cudaStream_t st) {
  if (h->npack > 0) {
    unsigned gx = (unsigned) ((h->pack_max + 255)/256);
    if (gx > 1024) gx = 1024;
    halo_direct_kernel<0><<<dim3(gx, h->npack, 1), 256, 0, st>>>(h->L, h->msg_pack, data, h->dsend);
    CU(cudaGetLastError());
    h->launches++;
  }
  // Sends and receives are listed in the same direction order on every rank, so
  // several messages between the same pair of ranks (2 ranks in a periodic
  // direction) match up in order.
  if (!h->comm) {
    // loopback: every message goes to this rank itself; pack and unpack tables list
    // the same directions in the same order, so one copy moves them all
    i64 total = 0;
    for (int i = 0; i < h->npack; i++) total += h->pack_cnt[i];
    if (h->npack != h->nunpack) return fail("internal: loopback message tables differ");
    CU(cudaMemcpyAsync(h->drecv, h->dsend, sizeof(double)*total, cudaMemcpyDeviceToDevice, st));
  }
  else {
  NC(g_nccl.GroupStart());
  for (int i = 0; i < h->npack; i++) {
    NC(g_nccl.Send(h->dsend + h->pack_off[i], (size_t) h->pack_cnt[i], ncclDouble, h->pack_peer[i], h->comm, st));
  }
  for (int i = 0; i < h->nunpack; i++) {
    NC(g_nccl.Recv(h->drecv + h->unpack_off[i], (size_t) h->unpack_cnt[i], ncclDouble, h->unpack_peer[i], h->comm, st));
  }
  NC(g_nccl.GroupEnd());
  }
  if (h->nunpack > 0) {
    unsigned gx = (unsigned) ((h->unpack_max + 255)/256);
    if (gx > 1024) gx = 1024;
    halo_direct_kernel<1><<<dim3(gx, h->nunpack, 1), 256, 0, st>>>(h->L, h->msg_unpack, data, h->drecv);
    CU(cudaGetLastError());
    h->launches++;
  }
  return 0;
}

// the decomposed directions of the current lattice, by the cheapest valid route
static int lattice_exchange(b200lb_s * h, int scheme, cudaStream_t st) {
  if (scheme == B200LB_HALO_REDUCED && h->pushed_valid) return 0;   // already in the neighbours' halos
  if (zhalo_flush(h, st)) return -1;
  if (scheme == B200LB_HALO_REDUCED && have_comm(h) && h->opt_direct) {
    if (direct_setup(h)) return -1;
    if (h->direct_state == 1) return direct_exchange(h, h->lat[h->cur], st);
  }
  return halo_exchange(h, h->lat[h->cur], true, h->nplane, scheme, 1, st);
}

// lb_halo on the current lattice. Every sweep that streams wraps around the axes
// that are periodic within this rank, so their halo planes are only ever read by
// an observer of the host copy: with halo_lazy those local copies wait for one
// (flush_halo); directions decomposed over ranks are exchanged at once. Doing the
// local axes after the exchanged ones gives the same planes, edges and corners:
// each pass spans the full extent of the other axes and overwrites what an
// earlier pass could not know yet.
static int lattice_halo(b200lb_s * h, int scheme) {
  NvtxRange nvtx_("b200lb: Lattice halos");
  if (h->opt_lattices == 1 && (h->aa_ahead || h->aa_swapped)) return 0;   // see b200lb_halo_swap
  if (!h->opt_halo_lazy) {
    if (join_comm(h)) return -1;
    if (zhalo_flush(h, h->s_main)) return -1;
    return halo_exchange(h, h->lat[h->cur], true, h->nplane, scheme, 0, h->s_main);
  }
  if (have_comm(h) && h->opt_overlap) {
    // The exchange runs on its own stream behind the boundary slabs of the sweep
    // that produced this lattice, next to that sweep's interior launch. s_main
    // joins it again before the next operation on the lattice (join_comm).
    if (join_comm(h)) return -1;
    if (!h->bnd_valid) CU(cudaEventRecord(h->ev_bnd, h->s_main));
    h->bnd_valid = false;
    CU(cudaStreamWaitEvent(h->s_comm, h->ev_bnd, 0));
    if (lattice_exchange(h, scheme, h->s_comm)) return -1;
    CU(cudaEventRecord(h->ev_halo, h->s_comm));
    h->halo_inflight = true;
  }
  else {
    if (lattice_exchange(h, scheme, h->s_main)) return -1;
  }
  h->halo_deferred = true;
  h->halo_deferred_scheme = scheme;
  return 0;
}

static int flush_halo(b200lb_s * h) {
  if (join_comm(h)) return -1;
  if (zhalo_flush(h, h->s_main)) return -1;
  if (!h->halo_deferred) return 0;
  h->halo_deferred = false;
  return halo_exchange(h, h->lat[h->cur], true, h->nplane, h->halo_deferred_scheme, 2, h->s_main);
}

extern "C" int b200lb_halo_swap(b200lb_t * h, int scheme) {
  if (!h) return fail("null handle");
  if (h->opt_lattices == 1 && (h->aa_ahead || h->aa_swapped)) {
    // single lattice, every direction periodic within the rank: the sweeps wrap, and what
    // the odd sweep stored is already where the propagation takes it
    if (aa_supported(h)) return -1;
    return 0;
  }
  if (scheme == 1) scheme = B200LB_HALO_FULL;      // LB_HALO_OPENMP_FULL moves the same data
  if (scheme != B200LB_HALO_FULL && scheme != B200LB_HALO_REDUCED) return fail("bad halo scheme %d", scheme);
  if (h->prop_pending) { if (materialise(h)) return -1; }
  return lattice_halo(h, scheme);
}

extern "C" int b200lb_halo(b200lb_t * h) {
  if (!h) return fail("null handle");
  return b200lb_halo_swap(h, h->opt_halo);
}

// field_grad_compute with Lees-Edwards planes (src/field_grad.c:305-325): after the stencil
// has run over the lattice, the sites next to a plane are recomputed with the displaced,
// interpolated neighbour planes of field_leesedwards (le_field_kernels.cuh)
static int le_grad_fix(b200lb_s * h, int npt) {
  if (h->le_nlocal == 0) return 0;
  const Layout & L = h->L;
  if (h->desc.ndim != 3) return fail("order parameter across Lees-Edwards planes: three-dimensional models");
  if (h->desc.nlocal[0]/h->le_nlocal < 3)
    return fail("order parameter across Lees-Edwards planes: planes closer than 3 sites (the reference's buffer lookup, src/leesedwards.c:1028, breaks down there)");
  LePhiTab T;
  memset(&T, 0, sizeof(T));
  T.L = L;
  T.w0 = h->selfper[0]; T.w1 = h->selfper[1]; T.w2 = h->selfper[2];
  int loc[B200LB_LE_MAXPLANE];
  for (int np = 0; np < h->le_nlocal; np++) loc[np] = le_location(h, np);
  le_phi_tab_fill(T, h->le_nlocal, loc, h->le_uy, h->le_osc, h->le_period, h->le_nt0, h->le_field_t);
  if (!h->le_phibuf) CU(cudaMalloc(&h->le_phibuf, (size_t) T.nbuf*L.s0*sizeof(double)));
  dim3 block(128, 1, 1);
  dim3 gbuf((L.n[2] + 2 + 127)/128, L.n[1] + 2, T.nbuf);
  le_phi_buffer_kernel<<<gbuf, block, 0, h->s_main>>>(T, h->phi, h->le_phibuf);
  CU(cudaGetLastError());
  dim3 gfix((L.n[2] + 127)/128, L.n[1], T.nbuf);
  if (npt == 27) le_grad_fix_kernel<27><<<gfix, block, 0, h->s_main>>>(T, h->phi, h->le_phibuf, h->grad, h->delsq);
  else le_grad_fix_kernel<7><<<gfix, block, 0, h->s_main>>>(T, h->phi, h->le_phibuf, h->grad, h->delsq);
  CU(cudaGetLastError());
  h->launches += 2;
  return 0;
}

static int run_grad7(b200lb_s * h) {
  NvtxRange nvtx_("b200lb: phi gradients (7 point)");
  const Layout & L = h->L;
  if (need_planes(h, &h->grad, 3) || need_planes(h, &h->delsq, 1)) return -1;
  dim3 block(GRAD_TZ, GRAD_TY, 1);
  dim3 grid((L.n[2] + GRAD_TZ - 1)/GRAD_TZ, (L.n[1] + GRAD_TY - 1)/GRAD_TY,
	    (L.n[0] + GRAD_NX - 1)/GRAD_NX);
  grad7_kernel<<<grid, block, 0, h->s_main>>>(L, h->selfper[0], h->selfper[1], h->selfper[2],
					      h->phi, h->grad, h->delsq);
  CU(cudaGetLastError());
  h->launches++;
  if (le_grad_fix(h, 7)) return -1;
  h->last_grad_npt = 7;
  h->grad_state = 0;
  h->ch_grad_valid = true; h->ch_halo_done = false;
  return 0;
}

static int run_grad27(b200lb_s * h) {
  NvtxRange nvtx_("b200lb: phi gradients (27 point)");
  const Layout & L = h->L;
  if (h->desc.ndim != 3) return fail("the 27-point gradient is three-dimensional (ndim = %d)", h->desc.ndim);
  if (need_planes(h, &h->grad, 3) || need_planes(h, &h->delsq, 1)) return -1;
  dim3 block(G27_TZ, G27_TY, 1);
  dim3 grid((L.n[2] + G27_TZ - 1)/G27_TZ, (L.n[1] + G27_TY - 1)/G27_TY,
	    (L.n[0] + G27_NX - 1)/G27_NX);
  if (h->opt_grad_direct) {
    dim3 b2(128, 2, 1);
    dim3 g2((L.n[2] + 127)/128, (L.n[1] + 1)/2, (L.n[0] + G27M_NX - 1)/G27M_NX);
    grad27_direct_kernel<<<g2, b2, 0, h->s_main>>>(L, h->selfper[0], h->selfper[1], h->selfper[2],
						   h->phi, h->grad, h->delsq);
  }
  else {
    grad27_kernel<<<grid, block, 0, h->s_main>>>(L, h->selfper[0], h->selfper[1], h->selfper[2],
						 h->phi, h->grad, h->delsq);
  }
  CU(cudaGetLastError());
  h->launches++;
  if (le_grad_fix(h, 27)) return -1;
  h->last_grad_npt = 27;
  h->grad_state = 0;
  h->ch_grad_valid = true; h->ch_halo_done = false;
  return 0;
}

// grad_3d_27pt_solid_d2 (solid_grad_kernels.cuh): the binary fluid between flat walls
static int run_grad27_solid(b200lb_s * h) {
  const Layout & L = h->L;
  if (h->desc.ndim != 3) return fail("the 27-point gradient is three-dimensional (ndim = %d)", h->desc.ndim);
  if (h->le_nlocal) return fail("3d_27pt_solid with Lees-Edwards planes is not on this path");
  if (need_planes(h, &h->grad, 3) || need_planes(h, &h->delsq, 1)) return -1;
  KParams P;
  site_kparams(h, P);
  SolidGradTab T;
  memset(&T, 0, sizeof(T));
  T.L = L;
  for (int d = 0; d < 3; d++) { T.wrap[d] = h->selfper[d]; T.wall_lo[d] = P.wall_lo[d]; T.wall_hi[d] = P.wall_hi[d]; }
  dim3 block(128, 1, 1);
  dim3 grid((L.n[2] + 127)/128, L.n[1], L.n[0]);
  grad27_solid_kernel<<<grid, block, 0, h->s_main>>>(T, h->phi, h->grad, h->delsq);
  CU(cudaGetLastError());
  h->launches++;
  h->grad_state = 0;
  h->ch_grad_valid = true; h->ch_halo_done = false;
  h->last_grad_npt = 28;
  return 0;
}

extern "C" int b200lb_grad27_solid(b200lb_t * h) {
  if (!h) return fail("null handle");
  if (!h->phi) return fail("no order parameter (ndist = 1)");
  if (h->ch_on) return fail("3d_27pt_solid for the finite-difference order parameter (walls there need the no-flux masks of src/advection_bcs.c) is not on this path");
  return run_grad27_solid(h);
}

extern "C" int b200lb_phi_halo(b200lb_t * h) {
  if (!h) return fail("null handle");
  NvtxRange nvtx_("b200lb: phi halos");
  if (!h->phi) return fail("no order parameter (ndist = 1)");
  if (join_comm(h)) return -1;
  if (!h->opt_halo_lazy) return halo_exchange(h, h->phi, false, 1, B200LB_HALO_FULL, 0, h->s_main);
  // directions decomposed over ranks now; the local periodic copies only when the host
  // looks at phi (the gradient stencil wraps around those directions)
  if (halo_exchange(h, h->phi, false, 1, B200LB_HALO_FULL, 1, h->s_main)) return -1;
  h->phi_halo_deferred = true;
  return 0;
}

static int flush_phi_halo(b200lb_s * h) {
  if (!h->phi_halo_deferred) return 0;
  h->phi_halo_deferred = false;
  return halo_exchange(h, h->phi, false, 1, B200LB_HALO_FULL, 2, h->s_main);
}

// ---------------------------------------------------------------------------
// order parameter, gradient, hydro fills
// ---------------------------------------------------------------------------

extern "C" int b200lb_phi_from_g(b200lb_t * h) {
  if (!h) return fail("null handle");
  NvtxRange nvtx_("b200lb: phi_lb_to_field");
  if (!h->phi || h->desc.ndist != 2) return fail("no order parameter distribution (ndist = 1)");
  // the bounce-back of a pending propagation is part of what phi sums: stream first
  if (h->prop_pending && h->prop_wall) { if (materialise(h)) return -1; }
  if (join_comm(h)) return -1;
  const double * g = h->lat[h->cur] + (i64) h->desc.nvel*h->L.nsd;
  // The collision that follows, if it streams in the same pass, would pull the same g
  // again only to form its first moment: leave that here (24 B/site out, 24 in, against
  // the 8*nvel of g)
  const bool mom = h->opt_g_moment && h->prop_pending && h->opt_fused;
  if (mom) { if (need_planes(h, &h->jphi, 3)) return -1; }
  CU(h->ops->phi_from_g(h->L, h->selfper, g, h->phi, mom ? h->jphi : nullptr, h->prop_pending, h->s_main));
  h->launches++;
  h->jphi_valid = mom;
  h->phi_halo_deferred = false;
  if (h->grad_state == 1) h->grad_state = 2;      // the pending gradient belonged to the old phi
  return 0;
}

extern "C" int b200lb_grad7(b200lb_t * h) {
  if (!h) return fail("null handle");
  if (!h->phi) return fail("no order parameter (ndist = 1)");
  if (resolve_ch_force(h)) return -1;     // a recorded force wants the gradient being replaced
  if (h->opt_grad_lazy && h->desc.ndist == 2 && h->le_nlocal == 0) {
    // The only consumer on the path is the binary collision, which evaluates the same
    // 7-point stencil in registers; the arrays are filled when somebody asks for them.
    h->grad_state = 1;
    return 0;
  }
  return run_grad7(h);
}

extern "C" int b200lb_grad27(b200lb_t * h) {
  if (!h) return fail("null handle");
  if (!h->phi) return fail("no order parameter (ndist = 1)");
  if (resolve_ch_force(h)) return -1;
  // 40 B/site, against ~650 B/site of the binary collision that consumes it: the arrays
  // are always written (the collision reads them; no in-register variant).
  return run_grad27(h);
}

static int fill3(b200lb_s * h, double * data, int nf, const double v[3]) {
  const i64 n = h->L.nsd;
  fill_kernel<<<(unsigned) ((n + 255)/256), 256, 0, h->s_main>>>(data, n, n, nf, v[0], nf > 1 ? v[1] : 0.0, nf > 2 ? v[2] : 0.0);
  CU(cudaGetLastError());
  h->launches++;
  return 0;
}

extern "C" int b200lb_hydro_f_zero(b200lb_t * h, const double f[3]) {
  if (!h) return fail("null handle");
  // The whole field takes one value: remember it instead of sweeping 24 B/site
  // (src/hydro.c:306 fills all nsite every step). Materialised on demand.
  if (h->hydro_stale && h->last_ffield) { if (ensure_hydro(h)) return -1; }
  for (int i = 0; i < 3; i++) h->fconst[i] = f ? f[i] : 0.0;
  h->force_uniform = true;
  h->ch_force_pending = false;            // a recorded phi_force_calculation is wiped with the field
  return 0;
}

extern "C" int b200lb_hydro_u_zero(b200lb_t * h, const double u[3]) {
  if (!h) return fail("null handle");
  for (int i = 0; i < 3; i++) h->uconst[i] = u ? u[i] : 0.0;
  h->u_fill_pending = true;     // the next collision overwrites the interior anyway
  return 0;
}

extern "C" int b200lb_hydro_rho0(b200lb_t * h, double rho0) {
  if (!h) return fail("null handle");
  if (need_hydro(h)) return -1;
  if (ensure_hydro(h)) return -1;           // u of the last collision stays current
  const double v[3] = {rho0, 0.0, 0.0};
  return fill3(h, h->rho, 1, v);
}

extern "C" int b200lb_hydro_correct_momentum(b200lb_t * h) {
  if (!h) return fail("null handle");
  if (resolve_ch_force(h)) return -1;
  if (h->hydro_stale && h->last_ffield) { if (ensure_hydro(h)) return -1; }
  if (need_planes(h, &h->force, 3)) return -1;
  if (h->force_uniform) {                 // make the constant field real first
    if (fill3(h, h->force, 3, h->fconst)) return -1;
    h->force_uniform = false;
  }
  const Layout & L = h->L;
  dim3 block(128, 2, 1);
  dim3 grid((L.n[2] + 127)/128, (L.n[1] + 1)/2, L.n[0]);
  force_sum_kernel<<<grid, block, 0, h->s_main>>>(L, h->force, h->partial);
  CU(stats_final(h->partial, h->npartial, h->stats_dev, h->stats_chain, h->s_main));
  if (h->comm) {
    // MPI_Allreduce(3 doubles, SUM) of src/hydro.c:873, on the device
    NC(g_nccl.AllReduce(h->stats_dev + 4, h->stats_dev + 4, 3, ncclDouble, ncclSum, h->comm, h->s_main));
  }
  double vol = 1.0;
  for (int a = 0; a < 3; a++) vol *= (double) h->desc.nlocal[a]*h->desc.cartsz[a];
  force_correct_kernel<<<grid, block, 0, h->s_main>>>(L, h->force, h->stats_dev, 1.0/vol);
  CU(cudaGetLastError());
  h->launches += 3;
  return 0;
}

// ---------------------------------------------------------------------------
// the order parameter as a finite-difference field ("free_energy symmetric",
// src/ludwig.c:1159-1250): phi_force_calculation + phi_cahn_hilliard in one pass
// (ch_kernels.cu)
// ---------------------------------------------------------------------------

static int field_current(b200lb_s * h, int field);

extern "C" int b200lb_ch_create(b200lb_t * h, int order, int method) {
  if (!h) return fail("null handle");
  if (h->desc.ndist != 1) return fail("phi_ch_create: the finite-difference order parameter goes with ndist = 1 (symmetric_lb carries it as a second distribution)");
  if (h->desc.ndim != 3) return fail("phi_ch_create: two-dimensional models (2d_5pt_fluid gradient) are not on this path");
  if (h->desc.nlocal[2] == 1 && h->desc.cartsz[2] > 1) return fail("phi_ch_create: a lattice one site thick decomposed in z");
  for (int a = 0; a < 3; a++) {
    if (!h->desc.periodic[a]) return fail("phi_ch_create: walls need the solid gradient stencils and the no-flux masks of src/advection_bcs.c: not on this path");
  }
  if (h->le_nplanes && !h->opt_le_fd) return fail("phi_ch_create: Lees-Edwards planes for the order parameter (src/phi_cahn_hilliard.c:618): switched off by option le_fd = 0");
  if (order < 1 || order > 4) return fail("fd_advection_scheme_order %d: orders 1 to 4 are on this path (5 is a host loop in the reference, src/advection.c:471-476)", order);
  if (method != CH_FORCE_NONE && method != CH_FORCE_STRESS_DIVERGENCE && method != CH_FORCE_PHI_GRADMU) {
    return fail("fe_force_method %d: no_force (1), stress_divergence (2) and phi_gradmu (3) are on this path", method);
  }
  if (need_planes(h, &h->phi, 1) || need_planes(h, &h->phi2, 1)) return -1;
  if (need_planes(h, &h->grad, 3) || need_planes(h, &h->delsq, 1)) return -1;
  if (need_hydro(h)) return -1;
  h->ch_on = true;
  h->ch_order = order;
  h->ch_method = method;
  h->ch_force_pending = false;
  h->ch_grad_valid = false;
  h->ch_halo_done = false;
  return 0;
}

extern "C" int b200lb_ch_grad_mu_set(b200lb_t * h, const double gm[3]) {
  if (!h || !gm) return fail("null argument");
  if (gm[0] != h->ch_gmu[0] || gm[1] != h->ch_gmu[1] || gm[2] != h->ch_gmu[2]) {
    if (resolve_ch_force(h)) return -1;     // a recorded force belongs to the old gradient
  }
  for (int a = 0; a < 3; a++) h->ch_gmu[a] = gm[a];
  return 0;
}

// One pass over the lattice: the force on the fluid (do_force), the update of phi
// (do_update), or both.
// The layers of sites around every Lees-Edwards plane again, after ch_kernel has run over
// the lattice (ch_kernels.cuh): buffer planes of phi (4 per plane, time t =
// physics_control_time) and u_x (time t + 1), the gradient in the first buffer plane of
// either side, then the two passes and the plane reduction.
static int ch_le_run(b200lb_s * h, const ChParams & P, bool do_force, bool do_update) {
  const Layout & L = h->L;
  const int np = h->le_nlocal;
  if (np > CH_LE_MAXPLANE) return fail("finite-difference order parameter: more than %d Lees-Edwards planes per rank", CH_LE_MAXPLANE);
  if (h->desc.nlocal[0]/np < 4) return fail("finite-difference order parameter: Lees-Edwards planes closer than 4 sites");
  int loc[B200LB_LE_MAXPLANE];
  for (int p = 0; p < np; p++) {
    loc[p] = le_location(h, p);
    if (loc[p] - 1 < 1 || loc[p] + 2 > L.n[0]) return fail("finite-difference order parameter: a Lees-Edwards plane within two sites of the rank boundary");
  }
  const size_t per = (size_t) 22*np*L.s0;
  if (!h->le_chbuf) {
    CU(cudaMalloc(&h->le_chbuf, (per + 3*np)*sizeof(double)));
    CU(cudaMemsetAsync(h->le_chbuf, 0, (per + 3*np)*sizeof(double), h->s_main));
  }
  double * bphi = h->le_chbuf;
  double * bgrad = bphi + (size_t) 4*np*L.s0;
  double * bdelsq = bgrad + (size_t) 6*np*L.s0;
  double * bux = bdelsq + (size_t) 2*np*L.s0;
  double * scr = bux + (size_t) 2*np*L.s0;
  double * fbar = scr + (size_t) 8*np*L.s0;
  const double t = h->le_field_t;

  LePhiTab T;
  memset(&T, 0, sizeof(T));
  T.L = L;
  T.w0 = h->selfper[0]; T.w1 = h->selfper[1]; T.w2 = h->selfper[2];
  dim3 block(128, 1, 1);
  le_phi_tab_fill4(T, np, loc, h->le_uy, h->le_osc, h->le_period, h->le_nt0, t);
  le_phi_buffer_kernel<<<dim3((L.n[2] + 2 + 127)/128, L.n[1] + 2, T.nbuf), block, 0, h->s_main>>>(T, h->phi, bphi);
  CU(cudaGetLastError());
  LeBufGradTab G;
  memset(&G, 0, sizeof(G));
  G.L = L; G.w1 = h->selfper[1]; G.w2 = h->selfper[2]; G.nplanes = np;
  for (int p = 0; p < np; p++) G.loc[p] = loc[p];
  dim3 gplane((L.n[2] + 127)/128, L.n[1], 2*np);
  if (h->last_grad_npt == 27) le_buffer_grad_kernel<27><<<gplane, block, 0, h->s_main>>>(G, h->phi, bphi, bgrad, bdelsq);
  else le_buffer_grad_kernel<7><<<gplane, block, 0, h->s_main>>>(G, h->phi, bphi, bgrad, bdelsq);
  CU(cudaGetLastError());
  h->launches += 2;
  if (do_update) {
    le_u_tab_fill(T, np, loc, h->le_uy, h->le_osc, h->le_period, h->le_nt0, t);
    le_u_buffer_kernel<<<gplane, block, 0, h->s_main>>>(T, h->u, bux);
    CU(cudaGetLastError());
    h->launches++;
  }
  ChLeParams Q;
  memset(&Q, 0, sizeof(Q));
  Q.nplanes = np;
  for (int p = 0; p < np; p++) Q.loc[p] = loc[p];
  Q.bphi = bphi; Q.bgrad = bgrad; Q.bdelsq = bdelsq; Q.bux = bux; Q.scr = scr; Q.fbar = fbar;
  {
    // lees_edw_plane_dy = t u_y (src/leesedwards.c:923-935), looking up and down
    const int ny = L.n[1];
    double dy = fmod(+(t*h->le_uy), 1.0*ny);
    Q.jdy_up = (int) floor(dy); Q.fr_up = dy - Q.jdy_up;
    dy = fmod(-(t*h->le_uy), 1.0*ny);
    Q.jdy_dn = (int) floor(dy); Q.fr_dn = dy - Q.jdy_dn;
    Q.ra = 0.5/(1.0*h->desc.nlocal[1]*h->desc.cartsz[1]*h->desc.nlocal[2]*h->desc.cartsz[2]);
  }
  if (decomposed(h, 2) && do_force) return fail("finite-difference order parameter with Lees-Edwards planes: the plane average of the force correction across ranks in z is not on this path");
  CU(h->opt_ch_exact ? ch_le_launch_exact(P, Q, do_force, do_update, h->s_main)
     : ch_le_launch_fast(P, Q, do_force, do_update, h->s_main));
  h->launches += do_force ? 3 : 2;
  return 0;
}

static int ch_run(b200lb_s * h, bool do_force, bool do_update) {
  if (!h->ch_on) return fail("no finite-difference order parameter: call b200lb_ch_create first");
  if (do_force && h->ch_method == CH_FORCE_NONE) do_force = false;
  if (do_force) h->ch_force_pending = false;
  if (!do_force && !do_update) return 0;
  if (!h->ch_grad_valid) return fail("phi_force_calculation / phi_cahn_hilliard: field_grad_compute has not run for the current order parameter");
  if (join_comm(h)) return -1;
  bool any_dec = false;
  for (int d = 0; d < 3; d++) {
    if (decomposed(h, d)) any_dec = true;
    if (do_update && h->ch_order >= 3 && !h->selfper[d]) {
      return fail("%s advection reads phi two sites away (nhalo = 2): on this path only around directions that are periodic within the rank",
		  (h->ch_order == 3) ? "third-order" : "fourth-order");
    }
  }
  if (any_dec && !h->ch_halo_done) {
    // The reference evaluates gradient and stress one site into a halo two sites wide;
    // here the neighbours' own values cross instead (same arithmetic on the same
    // operands: identical)
    if (halo_exchange(h, h->grad, false, 3, B200LB_HALO_FULL, 1, h->s_main)) return -1;
    if (halo_exchange(h, h->delsq, false, 1, B200LB_HALO_FULL, 1, h->s_main)) return -1;
    h->ch_halo_done = true;
  }
  if (do_update) {
    // hydro_u_halo, src/phi_cahn_hilliard.c:228: the velocity of the last collision
    if (field_current(h, B200LB_FIELD_U)) return -1;
    if (any_dec) { if (halo_exchange(h, h->u, false, 3, B200LB_HALO_FULL, 1, h->s_main)) return -1; }
  }
  if (do_force) {
    // the force the last collision read is about to change: rho,u it did not store
    // have to be recovered first
    if (h->hydro_stale && h->last_ffield) { if (ensure_hydro(h)) return -1; }
    if (need_planes(h, &h->force, 3)) return -1;
  }
  ChParams P;
  memset(&P, 0, sizeof(P));
  P.L = h->L;
  for (int d = 0; d < 3; d++) P.wrap[d] = h->selfper[d];
  P.phi = h->phi; P.grad = h->grad; P.delsq = h->delsq; P.u = h->u;
  P.force = h->force; P.phi_out = h->phi2;
  P.A = h->prm.a; P.B = h->prm.b; P.kappa = h->prm.kappa; P.mobility = h->prm.mobility;
  for (int a = 0; a < 3; a++) { P.gmu[a] = h->ch_gmu[a]; P.fconst[a] = h->fconst[a]; }
  P.accumulate = h->force_uniform ? 0 : 1;
  P.order = h->ch_order;
  P.method = h->ch_method;
  P.nx = h->opt_ch_planes;
  P.skipz = (h->desc.nlocal[2] == 1) ? 1 : 0;       // wz, src/phi_cahn_hilliard.c:999
  if (h->le_nlocal) {
    // with planes the force is the divergence of the face-interpolated stress whatever
    // the method (src/phi_force.c:89-96)
    if (do_force && P.accumulate) return fail("phi_force_calculation with Lees-Edwards planes: onto a force field that is uniform (hydro_f_zero) only");
    if (P.method != CH_FORCE_NONE) P.method = CH_FORCE_STRESS_DIVERGENCE;
  }
  CU(h->opt_ch_exact ? ch_launch_exact(P, do_force, do_update, h->s_main)
     : ch_launch_fast(P, do_force, do_update, h->s_main));
  h->launches++;
  if (h->le_nlocal) { if (ch_le_run(h, P, do_force, do_update)) return -1; }
  if (do_force) h->force_uniform = false;
  if (do_update) {
    double * t = h->phi; h->phi = h->phi2; h->phi2 = t;
    h->phi_halo_deferred = false;       // halos of the new phi are stale until the next field_halo
    h->ch_grad_valid = false;
    h->ch_halo_done = false;
  }
  return 0;
}

static int resolve_ch_force(b200lb_s * h) {
  if (!h->ch_force_pending) return 0;
  return ch_run(h, true, false);
}

extern "C" int b200lb_phi_force(b200lb_t * h) {
  if (!h) return fail("null handle");
  if (!h->ch_on) return fail("no finite-difference order parameter: call b200lb_ch_create first");
  if (h->ch_method == CH_FORCE_NONE) return 0;
  if (!h->ch_grad_valid) return fail("phi_force_calculation: field_grad_compute has not run for the current order parameter");
  if (h->opt_ch_fuse) {
    // evaluated by the pass of the Cahn-Hilliard update that follows (same inputs), or
    // by whoever looks at the force field first
    h->ch_force_pending = true;
    return 0;
  }
  return ch_run(h, true, false);
}

extern "C" int b200lb_cahn_hilliard(b200lb_t * h) {
  if (!h) return fail("null handle");
  return ch_run(h, h->ch_force_pending, true);
}

extern "C" int b200lb_hydro_u_halo(b200lb_t * h) {
  if (!h) return fail("null handle");
  if (need_hydro(h)) return -1;
  if (field_current(h, B200LB_FIELD_U)) return -1;
  if (join_comm(h)) return -1;
  return halo_exchange(h, h->u, false, 3, B200LB_HALO_FULL, 0, h->s_main);
}

// ---------------------------------------------------------------------------
// viscosity model (src/visc_arrhenius.c; the collision's side is in run_sweep)
// ---------------------------------------------------------------------------

extern "C" int b200lb_visc_model_set(b200lb_t * h, int on) {
  if (!h) return fail("null handle");
  if (on && h->desc.ndist != 1) return fail("viscosity model: the single-fluid collision only (src/collision.c:386-404)");
  if (on) { if (need_planes(h, &h->eta, 1)) return -1; }
  h->visc_on = (on != 0);
  return 0;
}

extern "C" int b200lb_visc_arrhenius_commit(b200lb_t * h, double eta_plus, double eta_minus, double phistar) {
  if (!h) return fail("null handle");
  if (!h->phi) return fail("viscosity_model arrhenius requires a composition (src/ludwig.c:2019): create the order parameter first");
  if (!(eta_plus > 0.0) || !(eta_minus > 0.0) || phistar == 0.0) return fail("viscosity_model arrhenius: eta_plus, eta_minus > 0 and phistar != 0");
  if (b200lb_visc_model_set(h, 1)) return -1;
  h->visc_arrhenius = true;
  h->visc_eta_plus = eta_plus; h->visc_eta_minus = eta_minus; h->visc_phistar = phistar;
  return 0;
}

extern "C" int b200lb_visc_update(b200lb_t * h) {
  if (!h) return fail("null handle");
  if (!h->visc_arrhenius) return fail("no viscosity model to update: call b200lb_visc_arrhenius_commit first");
  if (need_planes(h, &h->eta, 1)) return -1;
  const Layout & L = h->L;
  dim3 block(128, 2, 1);
  dim3 grid((L.n[2] + 127)/128, (L.n[1] + 1)/2, L.n[0]);
  visc_arrhenius_kernel<<<grid, block, 0, h->s_main>>>(L, h->phi, h->eta, h->visc_eta_plus,
						       h->visc_eta_minus, h->visc_phistar);
  CU(cudaGetLastError());
  h->launches++;
  return 0;
}

// ---------------------------------------------------------------------------
// the time loop of src/ludwig.c:530-862 restricted to the hot path
// ---------------------------------------------------------------------------

extern "C" int b200lb_step(b200lb_t * h, int nsteps) {
  if (!h) return fail("null handle");
  const double zero[3] = {0.0, 0.0, 0.0};
  for (int n = 0; n < nsteps; n++) {
    if (b200lb_hydro_f_zero(h, zero)) return -1;
    // physics_control_time of the step about to start (src/physics.c:651-659): what the
    // order parameter is displaced by across the planes; the distributions use the step number
    h->le_field_t = 1.0*(h->le_tstart + h->le_step);
    if (h->desc.ndist == 2) {
      if (b200lb_phi_from_g(h)) return -1;
      if (b200lb_phi_halo(h)) return -1;
      if (h->opt_gradient == 28) { if (b200lb_grad27_solid(h)) return -1; }
      else if (h->opt_gradient == 27) { if (b200lb_grad27(h)) return -1; }
      else if (b200lb_grad7(h)) return -1;
    }
    if (h->ch_on) {                                 // src/ludwig.c:558-575, 727-768
      if (b200lb_phi_halo(h)) return -1;
      if (h->opt_gradient == 27) { if (b200lb_grad27(h)) return -1; }
      else if (b200lb_grad7(h)) return -1;
      if (b200lb_phi_force(h)) return -1;
      if (b200lb_cahn_hilliard(h)) return -1;
    }
    if (b200lb_hydro_u_zero(h, zero)) return -1;
    if (h->visc_on && h->visc_arrhenius) { if (b200lb_visc_update(h)) return -1; }     // src/ludwig.c:794
    h->le_step++;                                   // physics_control_next_step, src/ludwig.c:552
    if (b200lb_collide(h)) return -1;
    if (h->prop_pending) return fail("internal: propagation pending after collide");
    if (h->le_nlocal) { if (b200lb_le_apply(h, 1.0*h->le_step)) return -1; }   // src/ludwig.c:810
    if (lattice_halo(h, h->opt_halo)) return -1;
    if (h->open_on) { if (b200lb_open_bc_apply(h)) return -1; } // src/ludwig.c:820-831
    if (h->wall_on) { if (b200lb_wall_bbl(h)) return -1; }      // src/ludwig.c:837-842
    if (b200lb_propagate(h)) return -1;
  }
  return 0;
}

// ---------------------------------------------------------------------------
// diagnostics
// ---------------------------------------------------------------------------

extern "C" int b200lb_stats(b200lb_t * h, double out[8]) {
  if (!h || !out) return fail("null argument");
  NvtxRange nvtx_("b200lb: stats_distribution");
  if (aa_observe(h)) return -1;
  if (materialise(h)) return -1;
  if (join_comm(h)) return -1;
  const Layout & L = h->L;
  dim3 block(128, 2, 1);
  dim3 grid((L.n[2] + 127)/128, (L.n[1] + 1)/2, L.n[0]);
  CU(h->ops->stats(L, h->lat[h->cur], h->desc.ndist, h->partial, grid, block, h->s_main));
  CU(stats_final(h->partial, h->npartial, h->stats_dev, h->stats_chain, h->s_main));
  h->launches += 3;
  CU(cudaMemcpyAsync(h->stats_host, h->stats_dev, 8*sizeof(double), cudaMemcpyDeviceToHost, h->s_main));
  CU(cudaStreamSynchronize(h->s_main));
  for (int k = 0; k < 8; k++) out[k] = h->stats_host[k];
  return 0;
}

// Conserved sums gathered INSIDE the last collision sweep (option "diag"): the
// state that entered that collision, i.e. what stats_distribution_print /
// stats_distribution_momentum would report before it. No extra pass over the lattice.
extern "C" int b200lb_step_diag(b200lb_t * h, double out[8]) {
  if (!h || !out) return fail("null argument");
  if (!h->opt_diag || h->diag_count == 0) return fail("no in-sweep diagnostics: set option diag = 1 before lb_collide");
  if (join_comm(h)) return -1;
  CU(stats_final(h->diag, h->diag_count, h->stats_dev, h->stats_chain, h->s_main));
  h->launches += 2;
  CU(cudaMemcpyAsync(h->stats_host, h->stats_dev, 8*sizeof(double), cudaMemcpyDeviceToHost, h->s_main));
  CU(cudaStreamSynchronize(h->s_main));
  for (int k = 0; k < 8; k++) out[k] = h->stats_host[k];
  return 0;
}

extern "C" int b200lb_sync(b200lb_t * h) {
  if (!h) return fail("null handle");
  if (join_comm(h)) return -1;
  {
    const cudaError_t e = cudaStreamSynchronize(h->s_main);
    if (e != cudaSuccess) {
      // the watchdog of p2p_sync_kernel traps: the context is gone and stays gone
      if (h->p2p_ready) return fail("cudaStreamSynchronize: %s (peer-memory halo: a neighbour that does not reach step %llu within %d s trips the watchdog, which fails closed; every later call fails too)",
				    cudaGetErrorString(e), h->push_step, (int) (P2P_WATCHDOG_NS/1000000000ULL));
      return fail("cudaStreamSynchronize: %s", cudaGetErrorString(e));
    }
  }
  CU(cudaStreamSynchronize(h->s_x[0]));
  CU(cudaStreamSynchronize(h->s_x[1]));
  return 0;
}

// ---------------------------------------------------------------------------
// transfers: staged, double-buffered on two streams so the copy of one piece
// overlaps the layout conversion of the previous one
// ---------------------------------------------------------------------------

static int ensure_stage(b200lb_s * h) {
  for (int i = 0; i < 2; i++) {
    if (!h->stage[i]) CU(cudaMalloc(&h->stage[i], STAGE_BYTES));
  }
  return 0;
}

// nel planes starting at device plane e0 of `dev`; host array `host` holds them
// in `layout` for all nsite_host sites
// sites [first, first + count) of the host index space; `host` points at the first of them
static int transfer_range(b200lb_s * h, double * dev, int nel, double * host, int layout, bool to_device,
			  i64 first, i64 count);

static int transfer(b200lb_s * h, double * dev, int nel, double * host, int layout, bool to_device) {
  return transfer_range(h, dev, nel, host, layout, to_device, 0, h->nsite_host);
}

static int transfer_range(b200lb_s * h, double * dev, int nel, double * host, int layout, bool to_device,
			  i64 first, i64 count) {
  NvtxRange nvtx_(to_device ? "b200lb: memcpy host to device (layout conversion)" : "b200lb: memcpy device to host (layout conversion)");
  if (first < 0 || count < 0 || first + count > h->nsite_host) return fail("site range [%lld, %lld) outside the %ld host sites", (long long) first, (long long) (first + count), h->nsite_host);
  if (count != h->nsite_host && layout != B200LB_HOST_AOS) return fail("a range of sites is moved in the AOS layout only");
  int vvl = 0;
  if (layout > B200LB_HOST_AOSOA_BASE) {
    // -DADDR_AOSOA -DVVL=V (src/memory.h:260-284): AOS over blocks of V sites
    vvl = layout - B200LB_HOST_AOSOA_BASE;
    if (vvl < 2 || vvl > 64) return fail("AoSoA host layout: block length %d (VVL 2..64)", vvl);
    if (h->nsite_host % vvl) return fail("AoSoA host layout: %ld sites are not a multiple of the block length %d", h->nsite_host, vvl);
    layout = B200LB_HOST_AOS;
  }
  if (ensure_stage(h)) return -1;
  if (join_comm(h)) return -1;
  CU(cudaStreamSynchronize(h->s_main));
  XferParams X;
  memset(&X, 0, sizeof(X));
  X.L = h->L;
  for (int a = 0; a < 3; a++) { X.hostdim[a] = h->hostdim[a]; X.dax[a] = h->ops->dax[a]; }
  X.dev = dev;
  const i64 ns = h->nsite_host;
  int turn = 0;
  if (layout == B200LB_HOST_AOS) {
    i64 chunk = (i64) (STAGE_BYTES/sizeof(double))/nel;
    if (vvl > 1) chunk = chunk/vvl*vvl;       // pieces start on a block boundary
    X.aos = 1; X.nel = nel; X.e0 = 0; X.vvl = vvl;
    const i64 last = first + count;
    for (i64 s0 = first; s0 < last; s0 += chunk, turn ^= 1) {
      const i64 s1 = (s0 + chunk < last) ? s0 + chunk : last;
      cudaStream_t st = h->s_x[turn];
      X.stage = h->stage[turn]; X.hs0 = s0; X.hs1 = s1;
      const size_t nb = (size_t) (s1 - s0)*nel*sizeof(double);
      const unsigned nblk = (unsigned) ((s1 - s0 + 255)/256);
      if (to_device) {
	CU(cudaMemcpyAsync(X.stage, host + (s0 - first)*nel, nb, cudaMemcpyHostToDevice, st));
	xfer_kernel<1><<<nblk, 256, 0, st>>>(X);
      }
      else {
	xfer_kernel<0><<<nblk, 256, 0, st>>>(X);
	CU(cudaMemcpyAsync(host + (s0 - first)*nel, X.stage, nb, cudaMemcpyDeviceToHost, st));
      }
      CU(cudaGetLastError());
      h->launches++;
    }
  }
  else if (layout == B200LB_HOST_SOA) {
    const i64 chunk = (i64) (STAGE_BYTES/sizeof(double));
    X.aos = 0; X.nel = 1;
    for (int e = 0; e < nel; e++) {
      for (i64 s0 = 0; s0 < ns; s0 += chunk, turn ^= 1) {
	const i64 s1 = (s0 + chunk < ns) ? s0 + chunk : ns;
	cudaStream_t st = h->s_x[turn];
	X.stage = h->stage[turn]; X.hs0 = s0; X.hs1 = s1; X.e0 = e;
	const size_t nb = (size_t) (s1 - s0)*sizeof(double);
	const unsigned nblk = (unsigned) ((s1 - s0 + 255)/256);
	if (to_device) {
	  CU(cudaMemcpyAsync(X.stage, host + e*ns + s0, nb, cudaMemcpyHostToDevice, st));
	  xfer_kernel<1><<<nblk, 256, 0, st>>>(X);
	}
	else {
	  xfer_kernel<0><<<nblk, 256, 0, st>>>(X);
	  CU(cudaMemcpyAsync(host + e*ns + s0, X.stage, nb, cudaMemcpyDeviceToHost, st));
	}
	CU(cudaGetLastError());
	h->launches++;
      }
    }
  }
  else {
    return fail("unknown host layout %d", layout);
  }
  CU(cudaStreamSynchronize(h->s_x[0]));
  CU(cudaStreamSynchronize(h->s_x[1]));
  return 0;
}

extern "C" int b200lb_f_upload(b200lb_t * h, const double * f, int layout) {
  if (!h || !f) return fail("null argument");
  if (ensure_hydro(h)) return -1;   // hydro of the state being replaced stays available
  h->prop_pending = false;          // the host copy replaces whatever was in flight
  h->aa_swapped = false; h->aa_ahead = false; h->fstar_mode = 0;
  h->jphi_valid = false;
  h->bbl_armed = false;
  h->prop_wall = false;
  h->halo_deferred = false;
  h->zpacked[h->cur] = 0;           // the host copy brings its halo sites along
  if (transfer(h, h->lat[h->cur], h->nplane, const_cast<double *>(f), layout, true)) return -1;
  if (h->p2p_ready && h->opt_p2p) {
    // The upload wrote this rank's halo sites too. Neighbours may push into them only
    // after that, so the upload takes part in the signal/wait protocol like a sweep:
    // nobody's next sweep starts before every neighbour has finished uploading.
    h->push_step++;
    p2p_sync_kernel<<<1, 32, 0, h->s_main>>>(h->p2p, h->push_step, P2P_WATCHDOG_NS, 1);
    CU(cudaGetLastError());
    h->launches++;
    h->launches_at_sync = h->launches;
    h->pushed_valid = false;
  }
  return 0;
}

// lb_io_read in pieces (src/model.c:1568-1649 reads the lattice through an aggregator buffer,
// never as one array): host sites [first, first + count) of the AOS layout, f pointing at the
// first of them. The first piece (first = 0) resets the lattice state like b200lb_f_upload;
// the caller uploads every site before it steps.
extern "C" int b200lb_f_upload_range(b200lb_t * h, const double * f, long first, long count) {
  if (!h || !f) return fail("null argument");
  if (first == 0) {
    if (ensure_hydro(h)) return -1;
    h->aa_swapped = false; h->aa_ahead = false; h->fstar_mode = 0;
    h->prop_pending = false;
    h->jphi_valid = false;
    h->bbl_armed = false;
    h->prop_wall = false;
    h->halo_deferred = false;
    h->zpacked[h->cur] = 0;
  }
  if (transfer_range(h, h->lat[h->cur], h->nplane, const_cast<double *>(f), B200LB_HOST_AOS, true, first, count)) return -1;
  if (first + count == h->nsite_host && h->p2p_ready && h->opt_p2p) {
    // the last piece: like b200lb_f_upload, the upload takes part in the peer protocol
    h->push_step++;
    p2p_sync_kernel<<<1, 32, 0, h->s_main>>>(h->p2p, h->push_step, P2P_WATCHDOG_NS, 1);
    CU(cudaGetLastError());
    h->launches++;
    h->launches_at_sync = h->launches;
    h->pushed_valid = false;
  }
  return 0;
}

// ---------------------------------------------------------------------------
// isothermal fluctuations: generator state (noise_t, src/noise.c)
// ---------------------------------------------------------------------------

extern "C" int b200lb_noise_init(b200lb_t * h, int master_seed) {
  if (!h) return fail("null handle");
  if (need_planes(h, &h->noise, 2)) return -1;
  NoiseInit N;
  memset(&N, 0, sizeof(N));
  N.L = h->L;
  for (int d = 0; d < 3; d++) N.dax[d] = h->ops->dax[d];
  for (int a = 0; a < 3; a++) {
    // uniform decomposition: cs_nlocal_offset = coords*nlocal (src/coords.c)
    N.noffset[a] = h->desc.cartcoords[a]*h->desc.nlocal[a];
    N.ntotal[a]  = h->desc.cartsz[a]*h->desc.nlocal[a];
  }
  const unsigned int seed0[4] = {13u, 12953u, 712357u, 22383979u};     // src/noise.c:147
  for (int k = 0; k < 4; k++) N.seed0[k] = seed0[k];
  if (master_seed > 0) N.seed0[0] = (unsigned int) master_seed;
  dim3 block(128, 1, 1);
  dim3 grid((h->L.na[2] + 127)/128, h->L.na[1], h->L.na[0]);
  noise_init_kernel<<<grid, block, 0, h->s_main>>>(N, h->noise);
  CU(cudaGetLastError());
  h->launches++;
  return 0;
}

extern "C" int b200lb_noise_present_set(b200lb_t * h, int on) {
  if (!h) return fail("null handle");
  if (on && h->ops->nvel - h->ops->nhydro > B200LB_MAXNOISE)
    return fail("isothermal fluctuations: nvel %d has %d ghost modes, the site generator yields %d variates per draw (src/noise.h:18)",
		h->ops->nvel, h->ops->nvel - h->ops->nhydro, B200LB_MAXNOISE);
  h->noise_on = (on != 0);
  return 0;
}

static int transfer(b200lb_s * h, double * dev, int nel, double * host, int layout, bool to_device);

// noise_memcpy (src/noise.c:253-284). AOS: state[index*4 + ia] is two 64-bit elements
// per site, which is the device's own pairing; SOA: four planes of 32-bit words.
static int noise_transfer(b200lb_s * h, unsigned int * host, int layout, bool to_device) {
  if (!h->noise) return fail("noise_init was never called");
  if (layout == B200LB_HOST_AOS) {
    return transfer(h, h->noise, 2, reinterpret_cast<double *>(host), layout, to_device);
  }
  if (layout != B200LB_HOST_SOA) return fail("unknown host layout %d", layout);
  if (ensure_stage(h)) return -1;
  if (join_comm(h)) return -1;
  CU(cudaStreamSynchronize(h->s_main));
  XferParams X;
  memset(&X, 0, sizeof(X));
  X.L = h->L;
  for (int a = 0; a < 3; a++) { X.hostdim[a] = h->hostdim[a]; X.dax[a] = h->ops->dax[a]; }
  X.dev = h->noise;
  X.aos = 0; X.nel = 1;
  const i64 ns = h->nsite_host;
  const i64 chunk = (i64) (STAGE_BYTES/sizeof(unsigned int));
  int turn = 0;
  for (int ia = 0; ia < 4; ia++) {
    for (i64 s0 = 0; s0 < ns; s0 += chunk, turn ^= 1) {
      const i64 s1 = (s0 + chunk < ns) ? s0 + chunk : ns;
      cudaStream_t st = h->s_x[turn];
      X.stage = h->stage[turn]; X.hs0 = s0; X.hs1 = s1; X.e0 = ia/2; X.half = 1 + (ia & 1);
      const size_t nb = (size_t) (s1 - s0)*sizeof(unsigned int);
      const unsigned nblk = (unsigned) ((s1 - s0 + 255)/256);
      if (to_device) {
	CU(cudaMemcpyAsync(X.stage, host + ia*ns + s0, nb, cudaMemcpyHostToDevice, st));
	xfer_kernel<1><<<nblk, 256, 0, st>>>(X);
      }
      else {
	xfer_kernel<0><<<nblk, 256, 0, st>>>(X);
	CU(cudaMemcpyAsync(host + ia*ns + s0, X.stage, nb, cudaMemcpyDeviceToHost, st));
      }
      CU(cudaGetLastError());
      h->launches++;
    }
  }
  CU(cudaStreamSynchronize(h->s_x[0]));
  CU(cudaStreamSynchronize(h->s_x[1]));
  return 0;
}

extern "C" int b200lb_noise_state_upload(b200lb_t * h, const unsigned int * state, int layout) {
  if (!h || !state) return fail("null argument");
  if (need_planes(h, &h->noise, 2)) return -1;
  return noise_transfer(h, const_cast<unsigned int *>(state), layout, true);
}

extern "C" int b200lb_noise_state_download(b200lb_t * h, unsigned int * state, int layout) {
  if (!h || !state) return fail("null argument");
  return noise_transfer(h, state, layout, false);
}

extern "C" int b200lb_f_download(b200lb_t * h, double * f, int layout) {
  if (!h || !f) return fail("null argument");
  if (aa_observe(h)) return -1;
  if (h->prop_pending) {
    if (materialise(h)) return -1;      // streamed lattice: its halos are stale, as in the reference
  }
  else {
    if (flush_halo(h)) return -1;       // lb_halo was the last thing done: make its planes real
  }
  return transfer(h, h->lat[h->cur], h->nplane, f, layout, false);
}

static int field_info(b200lb_s * h, int field, double ** dev, int * nf) {
  if (field >= B200LB_FIELD_PHI && field <= B200LB_FIELD_DELSQ && h->desc.ndist != 2 && !h->ch_on) {
    return fail("field %d needs an order parameter (ndist = 2, or b200lb_ch_create)", field);
  }
  switch (field) {
  case B200LB_FIELD_RHO:   if (need_hydro(h)) return -1; *dev = h->rho; *nf = 1; break;
  case B200LB_FIELD_U:     if (need_hydro(h)) return -1; *dev = h->u; *nf = 3; break;
  case B200LB_FIELD_FORCE: if (need_planes(h, &h->force, 3)) return -1; *dev = h->force; *nf = 3; break;
  case B200LB_FIELD_PHI:   *dev = h->phi; *nf = 1; break;
  case B200LB_FIELD_GRAD:  if (need_planes(h, &h->grad, 3)) return -1; *dev = h->grad; *nf = 3; break;
  case B200LB_FIELD_DELSQ: if (need_planes(h, &h->delsq, 1)) return -1; *dev = h->delsq; *nf = 1; break;
  case B200LB_FIELD_ETA:   if (need_planes(h, &h->eta, 1)) return -1; *dev = h->eta; *nf = 1; break;
  default: return fail("unknown field %d", field);
  }
  return 0;
}

extern "C" int b200lb_field_upload(b200lb_t * h, int field, const double * host, int layout) {
  if (!h || !host) return fail("null argument");
  double * dev; int nf;
  if (field_info(h, field, &dev, &nf)) return -1;
  if (field == B200LB_FIELD_FORCE) {
    if (h->hydro_stale && h->last_ffield) { if (ensure_hydro(h)) return -1; }
    h->force_uniform = false;
    h->ch_force_pending = false;          // overwritten by this upload
  }
  if (field >= B200LB_FIELD_PHI && field <= B200LB_FIELD_DELSQ) { if (resolve_ch_force(h)) return -1; }
  if (field == B200LB_FIELD_U || field == B200LB_FIELD_RHO) { if (ensure_hydro(h)) return -1; }   // the other of the two stays current
  if (field == B200LB_FIELD_PHI) {
    h->phi_halo_deferred = false;
    if (h->grad_state == 1) h->grad_state = 2;
    h->ch_grad_valid = false;
  }
  if (field == B200LB_FIELD_GRAD || field == B200LB_FIELD_DELSQ) {
    h->grad_state = 0;
    h->ch_grad_valid = true; h->ch_halo_done = false;    // the caller's gradient (field_grad_memcpy H2D)
  }
  if (field == B200LB_FIELD_U) h->u_fill_pending = false;
  return transfer(h, dev, nf, const_cast<double *>(host), layout, true);
}

// bring the device arrays of `field` up to date with everything recorded lazily
static int field_current(b200lb_s * h, int field) {
  if (field == B200LB_FIELD_FORCE) { if (resolve_ch_force(h)) return -1; }
  if (field == B200LB_FIELD_FORCE && h->force_uniform) {
    if (fill3(h, h->force, 3, h->fconst)) return -1;
  }
  if (field == B200LB_FIELD_U && h->u_fill_pending) {
    if (fill3(h, h->u, 3, h->uconst)) return -1;
    h->u_fill_pending = false;
  }
  if (field == B200LB_FIELD_U || field == B200LB_FIELD_RHO) {
    if (ensure_hydro(h)) return -1;
  }
  if (field == B200LB_FIELD_PHI) { if (flush_phi_halo(h)) return -1; }
  if (field == B200LB_FIELD_GRAD || field == B200LB_FIELD_DELSQ) {
    if (h->grad_state == 2) return fail("the gradient of the previous order parameter was never materialised (option grad_lazy = 0 keeps the arrays current)");
    if (h->grad_state == 1) { if (run_grad7(h)) return -1; }
  }
  return 0;
}

extern "C" int b200lb_field_download(b200lb_t * h, int field, double * host, int layout) {
  if (!h || !host) return fail("null argument");
  double * dev; int nf;
  if (field_info(h, field, &dev, &nf)) return -1;
  if (field_current(h, field)) return -1;
  return transfer(h, dev, nf, host, layout, false);
}

// Device reductions over the interior sites of this rank, per component:
// out[4*c + 0..3] = sum, sum of squares, min, max. What stats_velocity_minmax
// (src/stats_velocity.c:43-107) and stats_field_info (src/phi_stats.c:89-132,145-230)
// compute on the host after a full-field D2H; the caller finishes with the same
// MPI_Reduce over ranks as the reference.
extern "C" int b200lb_field_stats(b200lb_t * h, int field, double * out) {
  if (!h || !out) return fail("null argument");
  double * dev; int nf;
  if (field_info(h, field, &dev, &nf)) return -1;
  if (field_current(h, field)) return -1;
  if (join_comm(h)) return -1;
  const Layout & L = h->L;
  dim3 block(128, 2, 1);
  dim3 grid((L.n[2] + 127)/128, (L.n[1] + 1)/2, L.n[0]);
  for (int c = 0; c < nf; c++) {
    field_stats_kernel<<<grid, block, 0, h->s_main>>>(L, dev + (i64) c*L.nsd, h->partial);
    CU(stats_final(h->partial, h->npartial, h->stats_dev, h->stats_chain, h->s_main));
    h->launches += 3;
    CU(cudaMemcpyAsync(h->stats_host, h->stats_dev, 8*sizeof(double), cudaMemcpyDeviceToHost, h->s_main));
    CU(cudaStreamSynchronize(h->s_main));
    for (int k = 0; k < 4; k++) out[4*c + k] = h->stats_host[k];
  }
  return 0;
}

// ---------------------------------------------------------------------------
// b200lb_field_stats without any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph
// capture, tight serving loops, harnesses that forbid it).
//
// Same kernels as b200lb_field_stats; the per-block partials and the folded result live
// in a file-scope device array instead of h->partial / h->stats_dev. Every GPU that
// loads the library holds the array (B200LB_NOALLOC_BYTES) until exit. ALL calls on a
// device share it and it keeps the last call's contents: calls that overlap on the GPU
// (two handles, two streams) race. field_stats_kernel writes every partial slot the fold
// reads, so only the result slots are cleared per call.
// ---------------------------------------------------------------------------

// B200LB_NOALLOC_MAXBLOCK (include/b200lb.h): 65536 blocks of 128 x 2 sites, up to 256^3
#define B200LB_NOALLOC_DOUBLES  (8*B200LB_NOALLOC_MAXBLOCK + 8 + 2048)   // partials, result, the 256 x 8 chain sums of stats_final
#define B200LB_NOALLOC_MAXDEV   64

static __device__ __align__(256) double g_noalloc_scratch[B200LB_NOALLOC_DOUBLES];

// In host code g_noalloc_scratch is only a stand-in for the symbol: the device address
// comes from cudaGetSymbolAddress, once per device, with that device current.
static double * g_noalloc_ptr[B200LB_NOALLOC_MAXDEV];

static int noalloc_scratch(double ** out) {
  int dev = -1;
  CU(cudaGetDevice(&dev));
  if (dev < 0 || dev >= B200LB_NOALLOC_MAXDEV) return fail("no-alloc scratch: device index %d out of range", dev);
  if (!g_noalloc_ptr[dev]) {
    double * ptr = nullptr;
    CU(cudaGetSymbolAddress((void **) &ptr, g_noalloc_scratch));
    if (!ptr) return fail("no-alloc scratch: cudaGetSymbolAddress returned a null address");
    g_noalloc_ptr[dev] = ptr;
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}

// out[4*c + 0..3] = sum, sum of squares, min, max per component over the interior sites,
// as b200lb_field_stats. Calls no cudaMalloc / cudaFree: a field whose device planes do
// not exist yet, or whose values are still recorded lazily (materialising them may
// allocate), is refused with a message instead.
extern "C" int b200lb_field_stats_noalloc(b200lb_t * h, int field, double * out) {
  if (!h || !out) return fail("null argument");
  double * dev = nullptr; int nf = 0;
  bool lazy = false;
  switch (field) {
  case B200LB_FIELD_RHO:   dev = h->rho;   nf = 1; lazy = h->hydro_stale; break;
  case B200LB_FIELD_U:     dev = h->u;     nf = 3; lazy = h->hydro_stale || h->u_fill_pending; break;
  case B200LB_FIELD_FORCE: dev = h->force; nf = 3; lazy = h->force_uniform || h->ch_force_pending; break;
  case B200LB_FIELD_PHI:   dev = h->phi;   nf = 1; break;      // interior sites only: deferred halo copies do not matter
  case B200LB_FIELD_GRAD:  dev = h->grad;  nf = 3; lazy = (h->grad_state != 0); break;
  case B200LB_FIELD_DELSQ: dev = h->delsq; nf = 1; lazy = (h->grad_state != 0); break;
  case B200LB_FIELD_ETA:   dev = h->eta;   nf = 1; break;
  default: return fail("unknown field %d", field);
  }
  if (!dev) return fail("b200lb_field_stats_noalloc: field %d has no device planes yet and this entry point allocates nothing (upload it, or call b200lb_field_stats once)", field);
  if (lazy) return fail("b200lb_field_stats_noalloc: field %d is recorded lazily (its values would have to be materialised, which may allocate): call b200lb_field_download or b200lb_field_stats first", field);
  const Layout & L = h->L;
  dim3 block(128, 2, 1);
  dim3 grid((L.n[2] + 127)/128, (L.n[1] + 1)/2, L.n[0]);
  const i64 nblock = (i64) grid.x*grid.y*grid.z;
  if (nblock > B200LB_NOALLOC_MAXBLOCK)
    return fail("b200lb_field_stats_noalloc: %lld blocks of 128 x 2 sites exceed the static scratch (B200LB_NOALLOC_MAXBLOCK = %d, i.e. up to 256^3 interior sites): use b200lb_field_stats", (long long) nblock, B200LB_NOALLOC_MAXBLOCK);
  double * scratch = nullptr;
  if (noalloc_scratch(&scratch)) return -1;
  double * result = scratch + (size_t) 8*B200LB_NOALLOC_MAXBLOCK;
  if (join_comm(h)) return -1;
  for (int c = 0; c < nf; c++) {
    CU(cudaMemsetAsync(result, 0, 8*sizeof(double), h->s_main));
    field_stats_kernel<<<grid, block, 0, h->s_main>>>(L, dev + (i64) c*L.nsd, scratch);
    CU(stats_final(scratch, nblock, result, result + 8, h->s_main));
    h->launches += 3;
    CU(cudaMemcpyAsync(h->stats_host, result, 8*sizeof(double), cudaMemcpyDeviceToHost, h->s_main));
    CU(cudaStreamSynchronize(h->s_main));
    for (int k = 0; k < 4; k++) out[4*c + k] = h->stats_host[k];
  }
  return 0;
}

// lb_io_aggr_pack / lb_io_aggr_unpack with binary records (src/model.c:1498-1560,
// lb_write_buf :1385): the interior sites in (ic,jc,kc) order, kc fastest, each record
// [n][p] doubles -- packed straight from the device layout, so the halo shell never
// crosses PCIe and the host does no repacking pass.
static int io_transfer(b200lb_s * h, double * buf, bool to_device) {
  if (ensure_stage(h)) return -1;
  if (join_comm(h)) return -1;
  CU(cudaStreamSynchronize(h->s_main));
  XferParams X;
  memset(&X, 0, sizeof(X));
  X.L = h->L;
  for (int a = 0; a < 3; a++) { X.hostdim[a] = h->desc.nlocal[a]; X.dax[a] = h->ops->dax[a]; X.hoff[a] = 1; }
  X.dev = h->lat[h->cur];
  const int nel = h->nplane;
  const i64 ns = (i64) h->desc.nlocal[0]*h->desc.nlocal[1]*h->desc.nlocal[2];
  const i64 chunk = (i64) (STAGE_BYTES/sizeof(double))/nel;
  X.aos = 1; X.nel = nel; X.e0 = 0;
  int turn = 0;
  for (i64 s0 = 0; s0 < ns; s0 += chunk, turn ^= 1) {
    const i64 s1 = (s0 + chunk < ns) ? s0 + chunk : ns;
    cudaStream_t st = h->s_x[turn];
    X.stage = h->stage[turn]; X.hs0 = s0; X.hs1 = s1;
    const size_t nb = (size_t) (s1 - s0)*nel*sizeof(double);
    const unsigned nblk = (unsigned) ((s1 - s0 + 255)/256);
    if (to_device) {
      CU(cudaMemcpyAsync(X.stage, buf + s0*nel, nb, cudaMemcpyHostToDevice, st));
      xfer_kernel<1><<<nblk, 256, 0, st>>>(X);
    }
    else {
      xfer_kernel<0><<<nblk, 256, 0, st>>>(X);
      CU(cudaMemcpyAsync(buf + s0*nel, X.stage, nb, cudaMemcpyDeviceToHost, st));
    }
    CU(cudaGetLastError());
    h->launches++;
  }
  CU(cudaStreamSynchronize(h->s_x[0]));
  CU(cudaStreamSynchronize(h->s_x[1]));
  return 0;
}

extern "C" long b200lb_io_record_doubles(const b200lb_t * h) {
  return h ? (long) h->nplane : 0;
}

extern "C" int b200lb_io_pack(b200lb_t * h, double * buf) {
  if (!h || !buf) return fail("null argument");
  if (aa_observe(h)) return -1;
  if (h->prop_pending) { if (materialise(h)) return -1; }
  return io_transfer(h, buf, false);
}

extern "C" int b200lb_io_unpack(b200lb_t * h, const double * buf) {
  if (!h || !buf) return fail("null argument");
  if (ensure_hydro(h)) return -1;
  h->prop_pending = false;
  h->aa_swapped = false; h->aa_ahead = false; h->fstar_mode = 0;
  h->jphi_valid = false;
  h->bbl_armed = false;
  h->prop_wall = false;
  h->halo_deferred = false;
  h->zpacked[h->cur] = 0;
  if (io_transfer(h, const_cast<double *>(buf), true)) return -1;
  if (h->p2p_ready && h->opt_p2p) {
    // like lb_memcpy H2D: the restart read takes part in the signal/wait protocol
    h->push_step++;
    p2p_sync_kernel<<<1, 32, 0, h->s_main>>>(h->p2p, h->push_step, P2P_WATCHDOG_NS, 1);
    CU(cudaGetLastError());
    h->launches++;
    h->launches_at_sync = h->launches;
    h->pushed_valid = false;
  }
  return 0;
}

// ---------------------------------------------------------------------------
// NCCL communicator
// ---------------------------------------------------------------------------

extern "C" int b200lb_nccl_unique_id(char id[128]) {
  if (nccl_load()) return -1;
  ncclUniqueId uid;
  NC(g_nccl.GetUniqueId(&uid));
  static_assert(sizeof(uid) == 128, "ncclUniqueId is 128 bytes");
  memcpy(id, &uid, 128);
  return 0;
}

extern "C" int b200lb_comm_init(b200lb_t * h, const char id[128]) {
  if (!h || !id) return fail("null argument");
  if (nccl_load()) return -1;
  ncclUniqueId uid;
  memcpy(&uid, id, 128);
  NC(g_nccl.CommInitRank(&h->comm, h->desc.nranks, uid, h->desc.rank));
  return 0;
}

// ---------------------------------------------------------------------------
// peer-memory halo: CUDA IPC handles of the two lattices and the flag array
// ---------------------------------------------------------------------------

static int p2p_flags_alloc(b200lb_s * h) {
  if (h->p2p.flags) return 0;
  CU(cudaMalloc(&h->p2p.flags, 32*sizeof(unsigned long long)));
  CU(cudaMemset(h->p2p.flags, 0, 32*sizeof(unsigned long long)));
  return 0;
}

extern "C" int b200lb_p2p_export(b200lb_t * h, char out[192]) {
  if (!h || !out) return fail("null argument");
  if (h->opt_lattices == 1) {
    // one array: its handle twice (the importer maps it once)
    if (p2p_flags_alloc(h)) return -1;
    cudaIpcMemHandle_t hd1[3];
    CU(cudaIpcGetMemHandle(&hd1[0], h->lat[0]));
    hd1[1] = hd1[0];
    CU(cudaIpcGetMemHandle(&hd1[2], h->p2p.flags));
    memcpy(out, hd1, 192);
    return 0;
  }
  if (p2p_flags_alloc(h)) return -1;
  cudaIpcMemHandle_t hd[3];
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
  CU(cudaIpcGetMemHandle(&hd[0], h->lat[0]));
  CU(cudaIpcGetMemHandle(&hd[1], h->lat[1]));
  CU(cudaIpcGetMemHandle(&hd[2], h->p2p.flags));
  memcpy(out, hd, 192);
  return 0;
}

// all = nranks x 192 bytes (every rank's export, in rank order). With nranks = 1 and
// option loopback the rank is its own neighbour and no IPC is involved.
extern "C" int b200lb_p2p_import(b200lb_t * h, const char * all, int nranks) {
  if (!h) return fail("null handle");
  if (nranks != h->desc.nranks) return fail("p2p_import: nranks %d != %d", nranks, h->desc.nranks);
  if (p2p_flags_alloc(h)) return -1;
  const Layout & L = h->L;
  for (int d = 0; d < 3; d++) {
    if (decomposed(h, d) && L.n[d] < 2) return fail("peer-memory halo needs at least 2 sites per decomposed direction");
  }
  double * rlat[2][64]; unsigned long long * rflags[64];
  bool opened[64];
  if (nranks > 64) return fail("p2p_import: at most 64 ranks");
  for (int r = 0; r < nranks; r++) opened[r] = false;
  for (int k = 0; k < 27; k++) { h->peer_lat[0][k] = h->peer_lat[1][k] = nullptr; h->p2p.peer_flags[k] = nullptr; }
  int npeer = 0;
  for (int k = 0; k < 27; k++) {
    const int dir[3] = {k/9 - 1, (k/3)%3 - 1, k%3 - 1};        // device axes
    if (k == 13) continue;
    bool ok = true;
    int pdir[3] = {0, 0, 0};
    for (int d = 0; d < 3; d++) {
      if (dir[d] != 0 && !decomposed(h, d)) ok = false;
      pdir[h->ops->dax[d]] = dir[d];
    }
    if (!ok) continue;
    const int r = cart_rank(h, pdir);
    if (r < 0) continue;
    if (r == h->desc.rank) {
      h->peer_lat[0][k] = h->lat[0]; h->peer_lat[1][k] = h->lat[1]; h->p2p.peer_flags[k] = h->p2p.flags;
    }
    else {
      if (!all) return fail("p2p_import: handles missing");
      if (!opened[r]) {
        if (h->n_ipc_open + 3 > (int) (sizeof(h->ipc_open)/sizeof(h->ipc_open[0]))) return fail("p2p_import: too many peer mappings");
        cudaIpcMemHandle_t hd[3];
        memcpy(hd, all + (size_t) r*192, 192);
        void * p0 = nullptr, * p1 = nullptr, * p2 = nullptr;
        CU(cudaIpcOpenMemHandle(&p0, hd[0], cudaIpcMemLazyEnablePeerAccess));
        if (memcmp(&hd[0], &hd[1], sizeof(hd[0])) == 0) p1 = p0;          // a single-lattice neighbour
        else CU(cudaIpcOpenMemHandle(&p1, hd[1], cudaIpcMemLazyEnablePeerAccess));
        CU(cudaIpcOpenMemHandle(&p2, hd[2], cudaIpcMemLazyEnablePeerAccess));
        rlat[0][r] = (double *) p0; rlat[1][r] = (double *) p1; rflags[r] = (unsigned long long *) p2;
        h->ipc_open[h->n_ipc_open++] = p0; if (p1 != p0) h->ipc_open[h->n_ipc_open++] = p1; h->ipc_open[h->n_ipc_open++] = p2;
        opened[r] = true;
      }
      h->peer_lat[0][k] = rlat[0][r]; h->peer_lat[1][k] = rlat[1][r]; h->p2p.peer_flags[k] = rflags[r];
    }
    npeer++;
  }
  h->p2p_ready = (npeer > 0) ? 1 : 0;
  h->push_step = 0;
  h->pushed_valid = false;
  return 0;
}

// ---------------------------------------------------------------------------
// timing, pinned memory
// ---------------------------------------------------------------------------

extern "C" int b200lb_timer_start(b200lb_t * h) {
  if (!h) return fail("null handle");
  if (join_comm(h)) return -1;
  CU(cudaStreamSynchronize(h->s_main));
  CU(cudaEventRecord(h->ev0, h->s_main));
  return 0;
}

extern "C" int b200lb_timer_stop(b200lb_t * h, float * ms) {
  if (!h || !ms) return fail("null argument");
  if (join_comm(h)) return -1;
  CU(cudaEventRecord(h->ev1, h->s_main));
  CU(cudaEventSynchronize(h->ev1));
  CU(cudaEventElapsedTime(ms, h->ev0, h->ev1));
  return 0;
}

// TIMER_COLLIDE_KERNEL / TIMER_PROP_KERNEL (src/collision.c:198-207, src/propagation.c:84-91):
// the reference brackets the kernel launch with host timers behind a device synchronise.
// Here CUDA events are recorded on the launching stream right before and after every sweep
// launch, so the sum is the time of the sweep kernels alone whatever else a step launches
// (halo kernels, peer synchronisation, wall slabs count as sweeps of their own).
extern "C" int b200lb_kernel_timer_start(b200lb_t * h, int max_launches) {
  if (!h) return fail("null handle");
  if (max_launches < 1 || max_launches > 65536) return fail("kernel timer: 1..65536 launches");
  if (2*max_launches > h->tev_cap) {
    cudaEvent_t * ev = (cudaEvent_t *) realloc(h->tev, sizeof(cudaEvent_t)*2*max_launches);
    if (!ev) return fail("out of host memory");
    h->tev = ev;
    while (h->tev_cap < 2*max_launches) { CU(cudaEventCreate(&h->tev[h->tev_cap])); h->tev_cap++; }
  }
  h->tev_n = 0;
  h->tev_on = true;
  return 0;
}

// ms = total time of the sweep launches recorded since b200lb_kernel_timer_start,
// nlaunch = how many there were; stops the timer
extern "C" int b200lb_kernel_timer_stop(b200lb_t * h, double * ms, long * nlaunch) {
  if (!h || !ms || !nlaunch) return fail("null argument");
  h->tev_on = false;
  if (join_comm(h)) return -1;
  CU(cudaStreamSynchronize(h->s_main));
  double total = 0.0;
  for (int i = 0; i + 1 < h->tev_n; i += 2) {
    float t = 0.0f;
    CU(cudaEventElapsedTime(&t, h->tev[i], h->tev[i + 1]));
    total += t;
  }
  *ms = total;
  *nlaunch = h->tev_n/2;
  h->tev_n = 0;
  return 0;
}

extern "C" long b200lb_launch_count(const b200lb_t * h) { return h ? h->launches : 0; }

extern "C" int b200lb_host_alloc(void ** ptr, long nbytes) {
  if (!ptr) return fail("null argument");
  CU(cudaMallocHost(ptr, (size_t) nbytes));
  return 0;
}

extern "C" int b200lb_host_free(void * ptr) {
  if (ptr) CU(cudaFreeHost(ptr));
  return 0;
}

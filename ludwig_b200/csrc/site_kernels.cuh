// site_kernels.cuh -- the lattice sweep: pull streaming and/or collision.
//
// ONE kernel template covers the three sweeps of the reference:
//   PULL=0 COLLIDE=1   lb_collision_mrt1 / lb_collision_mrt2, in place
//                      (src/collision.c:224-599, :660-1027)
//   PULL=1 COLLIDE=0   lb_propagation_kernel (src/propagation.c:162-212)
//   PULL=1 COLLIDE=1   fused stream-then-collide used by the lazy-propagation
//                      mode: f'(r) = C( f(r - c_p) ), one read + one write of
//                      every population per step.
// It is not a port: the reference's generic Q x Q mode products become a
// pair-symmetric projection onto the 10 (6 in 2-d) hydrodynamic moments with
// constant-folded coefficients, ghost modes are touched only when a scheme
// relaxes them partially, and every access is a coalesced 64/128-bit access of
// the site-contiguous SoA layout (layout.cuh).

#pragma once

#include "layout.cuh"
#include "models.cuh"
#include "aux_kernels.cuh"

// Each model lives in its own translation unit (model_*.cu), which defines its
// ghost-mode projector table before including this file.
template <class M> __device__ __forceinline__ const GhostTab & ghost_tab() { return c_ghost; }

template <bool NC> __device__ __forceinline__ double ld1(const double * p) {
  if (NC) return __ldg(p);
  return *p;
}
template <bool NC> __device__ __forceinline__ double2 ld2(const double * p) {
  if (NC) return __ldg(reinterpret_cast<const double2 *>(p));
  return *reinterpret_cast<const double2 *>(p);
}

// ---------------------------------------------------------------------------
// Multiple-relaxation-time collision of one site (src/collision.c:339-558).
//   f      in: pre-collision populations, out: post-collision
//   F      total force on the site (global + field), physical components
//   sth    thermodynamic stress added to the equilibrium stress (binary fluid,
//          src/collision.c:833-848), symmetric storage, or nullptr
// Writes rho and u = (rho u + F/2)/rho (src/collision.c:355-382).
//   GM     bit 2 = viscosity model: relaxation times from the site's viscosity P.eta
//          (src/collision.c:393-404); ghost-mode handling: bit 0 = the scheme keeps a fraction of the incoming
//          ghost modes (BGK, TRT), bit 1 = fluctuations: shat (symmetric storage) is
//          added to the relaxed stress and ghat[k] to ghost mode k
//          (src/collision.c:526-544); bit 3 (sweep, binary fluid, PULL): the first
//          moment of the pulled g comes from P.jphi and g itself is not read
// ---------------------------------------------------------------------------

//   rts, rtb   inverse relaxation times of shear and bulk stress at this site
//   gk2    nullptr: ghost mode k keeps the fraction P.gkeep[k]; with a viscosity model
//          the fractions follow the site's viscosity: gk2[0] for the modes that relax
//          with the shear rate (all of them for BGK; k = 0, 4, 8 for TRT), gk2[1] for
//          the others (lb_relaxation_time_ghosts_v, src/collision.c:1443-1538)
template <class M, int GM, bool STH>
__device__ __forceinline__ void
mrt_site(double (&f)[M::NVEL], const double (&F)[3], const double * sth,
	 const double * shat, const double * ghat,
	 const KParams & P, const double rts, const double rtb, const double * gk2,
	 double & rho_out, double (&u)[3]) {

  constexpr bool GHOST = (GM & 1) != 0;
  constexpr bool NOISE = (GM & 2) != 0;

  constexpr int Q  = M::NVEL;
  constexpr int ND = M::NDIM;
  constexpr int NS = NSym<M>::value;
  constexpr int H  = (Q - 1)/2;
  constexpr int NG = Q - M::NHYDRO;
  constexpr double rdim = 1.0/ND;

  // ghost modes of the incoming state (only when some are kept)
  double mg[((GM & 3) != 0) ? NG : 1];
  if (GHOST) {
    const GhostTab & T = ghost_tab<M>();
    #pragma unroll
    for (int k = 0; k < NG; k++) {
      double s = 0.0;
      #pragma unroll
      for (int p = 0; p < Q; p++) s = fma(T.fwd[k][p], f[p], s);
      const double keep = (gk2 == nullptr) ? P.gkeep[k] : gk2[(k % 4 == 0) ? 0 : 1];
      mg[k] = keep*s;
    }
  }

  // hydrodynamic moments from (p, Q-p) pairs
  double rho = f[0];
  double j[ND];
  double S[NS];
  #pragma unroll
  for (int a = 0; a < ND; a++) j[a] = 0.0;
  #pragma unroll
  for (int k = 0; k < NS; k++) S[k] = 0.0;

  #pragma unroll
  for (int p = 1; p <= H; p++) {
    const double s = f[p] + f[Q - p];
    const double d = f[p] - f[Q - p];
    rho += s;
    #pragma unroll
    for (int a = 0; a < ND; a++) {
      if (M::cv(p, a) ==  1) j[a] += d;
      if (M::cv(p, a) == -1) j[a] -= d;
      #pragma unroll
      for (int b = a; b < ND; b++) {
	if (M::cv(p, a)*M::cv(p, b) ==  1) S[sidx<M>(a, b)] += s;
	if (M::cv(p, a)*M::cv(p, b) == -1) S[sidx<M>(a, b)] -= s;
      }
    }
  }
  // S_ab = sum_p (c_a c_b - cs2 delta_ab) f_p
  #pragma unroll
  for (int a = 0; a < ND; a++) S[sidx<M>(a, a)] -= (1.0/3.0)*rho;

  const double rrho = 1.0/rho;
  #pragma unroll
  for (int a = 0; a < 3; a++) u[a] = 0.0;
  #pragma unroll
  for (int a = 0; a < ND; a++) u[a] = rrho*(j[a] + 0.5*F[a]);

  // relax the stress towards rho u u (+ sth): trace with rtau_bulk, traceless
  // part with rtau_shear, then the force correction (src/collision.c:413-474)
  double seq[NS];
  double tr_s = 0.0, tr_seq = 0.0;
  #pragma unroll
  for (int a = 0; a < ND; a++) {
    #pragma unroll
    for (int b = a; b < ND; b++) {
      double e = rho*u[a]*u[b];
      if (STH) e += sth[sidx<M>(a, b)];
      seq[sidx<M>(a, b)] = e;
    }
    tr_s   += S[sidx<M>(a, a)];
    tr_seq += seq[sidx<M>(a, a)];
  }
  #pragma unroll
  for (int a = 0; a < ND; a++) {
    S[sidx<M>(a, a)]   -= rdim*tr_s;
    seq[sidx<M>(a, a)] -= rdim*tr_seq;
  }
  tr_s = tr_s - rtb*(tr_s - tr_seq);
  const double fcorr = 2.0 - rts;
  #pragma unroll
  for (int a = 0; a < ND; a++) {
    #pragma unroll
    for (int b = a; b < ND; b++) {
      const int k = sidx<M>(a, b);
      double s = S[k] - rts*(S[k] - seq[k]);
      if (a == b) s += rdim*tr_s;
      s += fcorr*(u[a]*F[b] + F[a]*u[b]);
      if (NOISE) s += shat[k];
      S[k] = s;
    }
  }
  #pragma unroll
  for (int a = 0; a < ND; a++) j[a] += F[a];

  // back-projection: f_p = w_p (rho + 3 c.j + 9/2 (c_a c_a - 1/3) S_aa + 9 c_a c_b S_ab)
  double A0 = rho;
  #pragma unroll
  for (int a = 0; a < ND; a++) {
    A0 -= 1.5*S[sidx<M>(a, a)];
    j[a] *= 3.0;
    #pragma unroll
    for (int b = a; b < ND; b++) S[sidx<M>(a, b)] *= (a == b) ? 4.5 : 9.0;
  }
  f[0] = M::wv(0)*A0;
  #pragma unroll
  for (int p = 1; p <= H; p++) {
    double ev = A0;
    double od = 0.0;
    #pragma unroll
    for (int a = 0; a < ND; a++) {
      if (M::cv(p, a) ==  1) od += j[a];
      if (M::cv(p, a) == -1) od -= j[a];
      #pragma unroll
      for (int b = a; b < ND; b++) {
	if (M::cv(p, a)*M::cv(p, b) ==  1) ev += S[sidx<M>(a, b)];
	if (M::cv(p, a)*M::cv(p, b) == -1) ev -= S[sidx<M>(a, b)];
      }
    }
    f[p]     = M::wv(p)*(ev + od);
    f[Q - p] = M::wv(p)*(ev - od);
  }

  if (NOISE) {
    #pragma unroll
    for (int k = 0; k < NG; k++) mg[k] = GHOST ? mg[k] + ghat[k] : ghat[k];
  }
  if ((GM & 3) != 0) {
    const GhostTab & T = ghost_tab<M>();
    #pragma unroll
    for (int p = 0; p < Q; p++) {
      double s = f[p];
      #pragma unroll
      for (int k = 0; k < NG; k++) s = fma(T.bwd[p][k], mg[k], s);
      f[p] = s;
    }
  }

  rho_out = rho;
}

// ---------------------------------------------------------------------------
// Order-parameter distribution of the binary fluid (src/collision.c:956-1023):
// g is rebuilt from phi, the relaxed flux j_phi and phi u u + mu delta.
// ---------------------------------------------------------------------------

// jin: sum_p c_p g_p of the incoming populations; every g_p is written, none is read
template <class M>
__device__ __forceinline__ void
phi_site_moment(double (&g)[M::NVEL], const double (&jin)[M::NDIM], double phi, double mu,
		const double (&u)[3], double rtau2) {

  constexpr int Q  = M::NVEL;
  constexpr int ND = M::NDIM;
  constexpr int NS = NSym<M>::value;
  constexpr int H  = (Q - 1)/2;

  double j[ND];
  #pragma unroll
  for (int a = 0; a < ND; a++) j[a] = jin[a];
  double S[NS];
  double B0 = 0.0;
  #pragma unroll
  for (int a = 0; a < ND; a++) {
    j[a] = 3.0*(j[a] - rtau2*(j[a] - phi*u[a]));
    #pragma unroll
    for (int b = a; b < ND; b++) {
      double s = phi*u[a]*u[b];
      if (a == b) { s += mu; B0 -= 1.5*s; s *= 4.5; } else { s *= 9.0; }
      S[sidx<M>(a, b)] = s;
    }
  }
  g[0] = M::wv(0)*B0 + phi;
  #pragma unroll
  for (int p = 1; p <= H; p++) {
    double ev = B0;
    double od = 0.0;
    #pragma unroll
    for (int a = 0; a < ND; a++) {
      if (M::cv(p, a) ==  1) od += j[a];
      if (M::cv(p, a) == -1) od -= j[a];
      #pragma unroll
      for (int b = a; b < ND; b++) {
	if (M::cv(p, a)*M::cv(p, b) ==  1) ev += S[sidx<M>(a, b)];
	if (M::cv(p, a)*M::cv(p, b) == -1) ev -= S[sidx<M>(a, b)];
      }
    }
    g[p]     = M::wv(p)*(ev + od);
    g[Q - p] = M::wv(p)*(ev - od);
  }
}

template <class M>
__device__ __forceinline__ void
phi_site(double (&g)[M::NVEL], double phi, double mu, const double (&u)[3],
	 double rtau2) {
  double j[M::NDIM];
  g_first_moment<M>(g, j);
  phi_site_moment<M>(g, j, phi, mu, u, rtau2);
}

// ---------------------------------------------------------------------------
// Isothermal fluctuations. Every site owns a KISS99 generator (Marsaglia 1999; four
// 32-bit words: a linear congruential word, a three-shift xorshift word and two 16-bit
// multiply-with-carry words, src/noise.c:456-472); one 32-bit output, with its two
// low bits dropped, yields up to ten three-bit indices into Ladd's eight-level
// discrete unit-variance distribution (src/noise.c:70-77,398-425).
// ---------------------------------------------------------------------------

__device__ __forceinline__ unsigned int kiss99(unsigned int (&w)[4]) {
  w[0] = 69069u*w[0] + 1234567u;
  unsigned int x = w[1];
  x ^= x << 17; x ^= x >> 13; x ^= x << 5;
  w[1] = x;
  w[2] = 36969u*(w[2] & 65535u) + (w[2] >> 16);
  w[3] = 18000u*(w[3] & 65535u) + (w[3] >> 16);
  return w[1] + (w[0] ^ ((w[2] << 16) + w[3]));
}

// level b of {-big, -small, 0, 0, 0, 0, +small, +big}
__device__ __forceinline__ double noise_level(unsigned int b, double big, double small) {
  b &= 7u;
  const double mag = (b == 0u || b == 7u) ? big : ((b == 1u || b == 6u) ? small : 0.0);
  return (b < 2u) ? -mag : mag;
}

// Random stress (symmetric storage) and ghost terms for one site: two draws of the
// site's generator, six variates for the stress (always six, also in two dimensions)
// and NVEL - NHYDRO for the ghosts (src/collision.c:1826-1920). w is advanced.
template <class M>
__device__ __forceinline__ void
fluctuations(unsigned int (&w)[4], const KParams & P,
	     double (&shat)[NSym<M>::value], double (&ghat)[M::NVEL - M::NHYDRO]) {

  constexpr int ND = M::NDIM;
  constexpr int NG = M::NVEL - M::NHYDRO;
  static_assert(NG <= B200LB_MAXNOISE, "one draw yields at most ten variates");

  unsigned int bits = kiss99(w) >> 2;
  double r[6];
  #pragma unroll
  for (int k = 0; k < 6; k++) { r[k] = noise_level(bits, P.nlv_big, P.nlv_small); bits >>= 3; }
  // r = xx xy xz yy yz zz
  const double dg[3] = {r[0], r[3], r[5]};
  const double od[3] = {r[1], r[2], r[4]};        // xy xz yz
  const double tr = (1.0/ND)*(dg[0] + dg[1] + (ND - 2.0)*dg[2]);
  #pragma unroll
  for (int a = 0; a < ND; a++) {
    shat[sidx<M>(a, a)] = (dg[a] - tr)*P.nv_diag + tr*P.nv_bulk;
    #pragma unroll
    for (int b = a + 1; b < ND; b++) shat[sidx<M>(a, b)] = od[a + b - 1]*P.nv_off;
  }

  if (P.nghost_draw) {          // uniform branch
    bits = kiss99(w) >> 2;
    #pragma unroll
    for (int k = 0; k < NG; k++) {
      ghat[k] = P.nv_ghost[k]*noise_level(bits, P.nlv_big, P.nlv_small);
      bits >>= 3;
    }
  }
  else {
    #pragma unroll
    for (int k = 0; k < NG; k++) ghat[k] = 0.0;
  }
}

// ---------------------------------------------------------------------------
// Packed halo of the contiguous axis (KParams::zoff). Populations with a component along the
// contiguous device axis, counted in index order per sign: D3Q19 has 5 per face, D3Q27 9,
// D2Q9 3.
// ---------------------------------------------------------------------------

template <class M> HD constexpr int zface_count() {
  int n = 0;
  for (int p = 0; p < M::NVEL; p++) if (dcv<M>(p, 2) == 1) n++;
  return n;
}
// plane of population p in the packed block: face 0 (c_z = +1) first, then face 1; a table
// the front end evaluates, so that unrolled loops fold the entries into immediates
struct ZPlaneTab { int pl[27]; };
template <class M> HD constexpr ZPlaneTab make_zplane_tab() {
  ZPlaneTab t{};
  int nlo = 0, nhi = 0;
  for (int p = 0; p < M::NVEL; p++) {
    t.pl[p] = 0;
    if (dcv<M>(p, 2) ==  1) t.pl[p] = nlo++;
    if (dcv<M>(p, 2) == -1) t.pl[p] = zface_count<M>() + nhi++;
  }
  return t;
}
template <class M> HD constexpr int zplane(int p) {
  constexpr ZPlaneTab t = make_zplane_tab<M>();
  return t.pl[p];
}
// slot of population p (c_z != 0) of the halo site above/below halo-inclusive column (xa, ya),
// relative to the start of the packed planes
template <class M> __device__ __forceinline__ i64 zslot(const KParams & P, const int p, const int xa, const int ya) {
  return (i64) zplane<M>(p)*P.zsp + (i64) xa*P.zsx + (i64) ya*P.zsy;
}
__device__ __forceinline__ void zstore(double * a, const double val, const int zvec) {
  if (zvec) {
#ifdef __CUDA_ARCH__
    asm volatile("st.global.v4.f64 [%0], {%2, %1, %1, %1};" :: "l"(a), "d"(0.0), "d"(val) : "memory");
#else
    a[0] = val; a[1] = 0.0; a[2] = 0.0; a[3] = 0.0;
#endif
  }
  else {
    *a = val;
  }
}

// ---------------------------------------------------------------------------
// The sweep. Thread (x) covers VEC consecutive sites of the contiguous axis,
// thread (y) a row; blockIdx.z is the slowest axis. With VEC = 2 populations
// moving along the contiguous axis are read as two 64-bit loads (they straddle
// the aligned pair), all others and every store are 128-bit.
// ---------------------------------------------------------------------------

// PUSH: the peer-memory halo rides on this sweep (its own instantiations: the kernels of a
// single-GPU run carry none of that code and keep their register count)
// AA: single-lattice streaming (the "AA pattern", Bailey et al. 2009) -- src == dst, every
// location is read and written by exactly one site per sweep:
//   1  even step: collide in place; population p is stored in the slot of its opposite,
//      Q - p, at the same site (post-collision, unstreamed, "swapped")
//   2  odd step: population p is pulled from the opposite slot of the site it streams from,
//      collided, and stored in its own slot at the site it streams TO: the lattice comes
//      out one propagation ahead, in the natural order
//   3  no collision: pull from the swapped slots into another lattice in natural order
//      (what an observer of the state between an even and an odd step needs)
template <class M, int VEC, bool PULL, bool COLLIDE, bool BINARY, int GM, bool DIAG, bool WALL = false, bool PUSH = false, int AA = 0>
__global__ void __launch_bounds__(256, PUSH ? 2 : 0)
lb_sweep_kernel(const KParams P) {

  constexpr int Q  = M::NVEL;
  constexpr int ND = M::NDIM;
  constexpr int NDIST = BINARY ? 2 : 1;
  constexpr bool NC = PULL && AA != 2;   // source is read-only when it is not the destination
  static_assert(AA == 0 || (!BINARY && !WALL && (GM & ~1) == 0), "single-lattice streaming: plain single-fluid sweeps");
  static_assert(AA != 3 || !PUSH, "the materialising sweep stores locally");
  static_assert(AA != 1 || (!PULL && COLLIDE), "AA even step is an in-place collision");
  static_assert(AA != 2 || (PULL && COLLIDE), "AA odd step pulls and collides");
  static_assert(AA != 3 || (PULL && !COLLIDE), "AA materialisation only streams");
  // GM bit 3: phi_from_g_kernel<PULL, MOM> has left the first moment of the pulled g in
  // P.jphi, which is all the binary collision takes from it: g is written, not read
  constexpr bool GMOM = BINARY && PULL && COLLIDE && ((GM & 8) != 0);
  constexpr int NLOAD = GMOM ? 1 : NDIST;
  static_assert(((GM & 8) == 0) || GMOM, "GM bit 3 belongs to the fused binary sweep");

  const Layout & L = P.L;

  int i2 = P.lo[2] + VEC*(blockIdx.x*blockDim.x + threadIdx.x);
  int i1 = P.lo[1] + blockIdx.y*blockDim.y + threadIdx.y;
  const int i0 = P.lo[0] + blockIdx.z;
  int nv;
  if (i2 >= P.hi[2] || i1 >= P.hi[1]) {
    // Outside the box. With in-kernel diagnostics every thread of the block must
    // reach the reduction: such threads redo the box's first site and store nothing.
    if (!DIAG) return;
    i2 = P.lo[2]; i1 = P.lo[1]; nv = 0;
  }
  else {
    nv = (VEC == 2 && i2 + 1 < P.hi[2]) ? 2 : 1;
  }
  double dq[8] = {0.0, 0.0, 1.0e300, -1.0e300, 0.0, 0.0, 0.0, 0.0};

  if constexpr (PUSH) {
    // Programmatic dependent launch: this grid may have started while the kernel before it
    // in the stream -- p2p_sync_kernel, still waiting for the neighbours' flags -- is
    // running. Only a thread whose sites touch a decomposed face reads halo sites or stores
    // into a neighbour, and only such a thread has to wait for that kernel to complete;
    // everybody else streams and collides meanwhile. (Without the launch attribute the
    // instruction returns at once.)
    bool edge = false;
    if (!P.wrap[0] && L.h[0]) edge = edge || i0 == 0 || i0 == L.n[0] - 1;
    if (!P.wrap[1] && L.h[1]) edge = edge || i1 == 0 || i1 == L.n[1] - 1;
    if (!P.wrap[2] && L.h[2]) edge = edge || i2 == 0 || i2 + VEC - 1 >= L.n[2] - 1;
#ifdef __CUDA_ARCH__
    if (edge) asm volatile("griddepcontrol.wait;" ::: "memory");
#endif
  }

  const i64 site = (i64)(i0 + L.h[0])*L.s0 + (i64)(i1 + L.h[1])*L.s1
                 + L.z0 + L.h[2] + i2;

  // Offsets that reach coordinate -1 / +1 along each axis. An axis that is
  // periodic within this rank wraps around here, so its halo planes are never
  // read (and need not be filled): SURVEY 7.4 "self-exchange fast path".
  i64 om[2], op[2];        // axes 0 and 1: block-uniform
  int zm[VEC], zp[VEC];    // contiguous axis: differs only for the first / last site of a row
  if (PULL || BINARY) {
    om[0] = -L.s0; op[0] = L.s0; om[1] = -L.s1; op[1] = L.s1;
    if (P.wrap[0]) {
      if (i0 == 0) om[0] = (i64) (L.n[0] - 1)*L.s0;
      if (i0 == L.n[0] - 1) op[0] = -(i64) (L.n[0] - 1)*L.s0;
    }
    if (P.wrap[1]) {
      if (i1 == 0) om[1] = (i64) (L.n[1] - 1)*L.s1;
      if (i1 == L.n[1] - 1) op[1] = -(i64) (L.n[1] - 1)*L.s1;
    }
    #pragma unroll
    for (int v = 0; v < VEC; v++) {
      zm[v] = (P.wrap[2] && (i2 + v == 0)) ? L.n[2] - 1 : -1;
      zp[v] = (P.wrap[2] && (i2 + v == L.n[2] - 1)) ? -(L.n[2] - 1) : 1;
    }
  }

  double f[NDIST][VEC][Q];

  // Packed z-face planes (KParams::zoff) of the source lattice: which of this thread's sites
  // pull from beyond a z face, and the column part of the slot for c = -1, 0, +1 along x and
  // y (32-bit: the planes of one lattice hold fewer than 2^31 words)
  bool zlo[VEC], zhi[VEC];
  int zcx[3], zcy[3], zpl = 0;
  if constexpr (PUSH && AA == 0 && !BINARY && PULL) {
    const bool zr = P.zread && !P.wrap[2];
    #pragma unroll
    for (int v = 0; v < VEC; v++) { zlo[v] = zr && (i2 + v == 0); zhi[v] = zr && (i2 + v == L.n[2] - 1); }
    zpl = (int) P.zsp;
    #pragma unroll
    for (int c = -1; c <= 1; c++) {
      int xa = i0 + L.h[0] - c, ya = i1 + L.h[1] - c;
      if (P.wrap[0]) { if (xa < L.h[0]) xa += L.n[0]; else if (xa >= L.h[0] + L.n[0]) xa -= L.n[0]; }
      if (P.wrap[1]) { if (ya < L.h[1]) ya += L.n[1]; else if (ya >= L.h[1] + L.n[1]) ya -= L.n[1]; }
      zcx[c + 1] = xa*(int) P.zsx;
      zcy[c + 1] = ya*(int) P.zsy;
    }
  }
  else {
    #pragma unroll
    for (int v = 0; v < VEC; v++) { zlo[v] = false; zhi[v] = false; }
    #pragma unroll
    for (int c = 0; c < 3; c++) { zcx[c] = 0; zcy[c] = 0; }
  }

  // Walls: is this site next to a wall layer, per device axis and side? (axes 0 and 1
  // are uniform over a row of threads; the contiguous axis differs per site)
  bool w0lo = false, w0hi = false, w1lo = false, w1hi = false;
  bool w2lo[VEC], w2hi[VEC];
  #pragma unroll
  for (int v = 0; v < VEC; v++) { w2lo[v] = false; w2hi[v] = false; }
  if (WALL && PULL) {
    w0lo = P.wall_lo[0] && (i0 == 0);  w0hi = P.wall_hi[0] && (i0 == L.n[0] - 1);
    w1lo = P.wall_lo[1] && (i1 == 0);  w1hi = P.wall_hi[1] && (i1 == L.n[1] - 1);
    #pragma unroll
    for (int v = 0; v < VEC; v++) {
      w2lo[v] = P.wall_lo[2] && (i2 + v == 0);
      w2hi[v] = P.wall_hi[2] && (i2 + v == L.n[2] - 1);
    }
  }

  if constexpr (WALL && PULL) {
    // Wall slabs (one site per thread). A population whose pull would reach into a wall
    // layer along some axes (the wall normal of that link) is instead
    //   no slip:  the one that left this site in the opposite direction (+ moving wall)
    //   slip s:   (1 - s) of that + s of its mirror image in the wall, taken from the
    //             site reached by the tangential part of the step alone
    // (wall_bbl_kernel / wall_bbl_slip_kernel, src/wall.c:997-1210, in pull form).
    static_assert(!WALL || VEC == 1, "wall slabs are swept one site per thread");
    #pragma unroll
    for (int n = 0; n < NDIST; n++) {
      #pragma unroll
      for (int p = 0; p < Q; p++) {
	const int c0 = dcv<M>(p, 0), c1 = dcv<M>(p, 1), c2 = dcv<M>(p, 2);
	const i64 o0 = (c0 == 1) ? om[0] : ((c0 == -1) ? op[0] : 0);
	const i64 o1 = (c1 == 1) ? om[1] : ((c1 == -1) ? op[1] : 0);
	const i64 o2 = (c2 == 1) ? zm[0] : ((c2 == -1) ? zp[0] : 0);
	const bool n0 = (c0 == 1 && w0lo) || (c0 == -1 && w0hi);
	const bool n1 = (c1 == 1 && w1lo) || (c1 == -1 && w1hi);
	const bool n2 = (c2 == 1 && w2lo[0]) || (c2 == -1 && w2hi[0]);
	const double * plane = P.src + (i64)(n*Q)*L.nsd + site;
	double val;
	if (p == 0 || !(n0 || n1 || n2)) {
	  val = ld1<NC>(plane + (i64) p*L.nsd + o0 + o1 + o2);
	}
	else {
	  val = ld1<NC>(plane + (i64)(Q - p)*L.nsd);
	  if (!P.wall_slip) {
	    val += P.wall_corr[p];
	  }
	  else {
	    const int nn = (n0 ? 1 : 0) + (n1 ? 1 : 0) + (n2 ? 1 : 0);
	    const bool tang = (c0 != 0 && !n0) || (c1 != 0 && !n1) || (c2 != 0 && !n2);
	    if (tang && nn < 3) {
	      // a face takes its wall's fraction, an edge the mean of its two walls
	      double sl = (n0 ? ((c0 == 1) ? P.wall_s_lo[0] : P.wall_s_hi[0]) : 0.0)
		        + (n1 ? ((c1 == 1) ? P.wall_s_lo[1] : P.wall_s_hi[1]) : 0.0)
		        + (n2 ? ((c2 == 1) ? P.wall_s_lo[2] : P.wall_s_hi[2]) : 0.0);
	      if (nn == 2) sl *= 0.5;
	      if (sl != 0.0) {
		const int q = P.wall_q[p][(n0 ? 1 : 0) | (n1 ? 2 : 0) | (n2 ? 4 : 0)];
		const double fk = ld1<NC>(plane + (i64) q*L.nsd + (n0 ? 0 : o0) + (n1 ? 0 : o1) + (n2 ? 0 : o2));
		// separately rounded, as the reference's (1 - s) fi + s fk without contraction
		val = __dadd_rn(__dmul_rn(1.0 - sl, val), __dmul_rn(sl, fk));
	      }
	    }
	  }
	}
	f[n][0][p] = val;
      }
    }
  }
  else {
  #pragma unroll
  for (int n = 0; n < NLOAD; n++) {
    #pragma unroll
    for (int p = 0; p < Q; p++) {
      // AA 2, 3: population p waits in the slot of its opposite
      const int ps = ((AA == 2 || AA == 3) && p > 0) ? Q - p : p;
      const double * src = P.src + (i64)(n*Q + ps)*L.nsd + site;
      int c2 = 0;
      if (PULL) {
	if (dcv<M>(p, 0) ==  1) src += om[0];
	if (dcv<M>(p, 0) == -1) src += op[0];
	if (dcv<M>(p, 1) ==  1) src += om[1];
	if (dcv<M>(p, 1) == -1) src += op[1];
	c2 = dcv<M>(p, 2);
      }
      if (c2 == 0) {
	if (VEC == 2) {
	  const double2 v = ld2<NC>(src);
	  f[n][0][p] = v.x; f[n][VEC - 1][p] = v.y;
	}
	else {
	  f[n][0][p] = ld1<NC>(src);
	}
      }
      else {
	// moving along the contiguous axis: the pair straddles the aligned one
	#pragma unroll
	for (int v = 0; v < VEC; v++) {
	  const double * at = src + v + (c2 == 1 ? zm[v] : zp[v]);
	  if constexpr (PUSH && AA == 0 && !BINARY) {
	    // first / last site of a row whose halo site the neighbours' sweeps left in the packed
	    // planes: another address for the same load. A select, not a branch -- a branch
	    // per population keeps the loads of a thread from being issued back to back, and
	    // the sweep then waits for memory one population at a time (0.81 -> 1.00 ms at
	    // 256^3, profiles/r2b_zpack.md)
	    const bool pk = (c2 == 1) ? zlo[v] : zhi[v];
	    const double * zat = P.src + P.zoff
	      + (i64) (zplane<M>(p)*zpl + zcx[dcv<M>(p, 0) + 1] + zcy[dcv<M>(p, 1) + 1]);
	    at = pk ? zat : at;
	  }
	  f[n][v][p] = ld1<NC>(at);
	}
      }
    }
  }
  }

  if (COLLIDE) {
    #pragma unroll
    for (int v = 0; v < VEC; v++) {
      const i64 s = site + v;
      double F[3] = {P.fg[0], P.fg[1], P.fg[2]};
      if (P.force != nullptr) {          // uniform branch: per-site force field present
	#pragma unroll
	for (int a = 0; a < ND; a++) F[a] += P.force[a*L.nsd + s];
      }
      double rho, u[3];
      constexpr int NGH = (Q > M::NHYDRO) ? Q - M::NHYDRO : 1;
      double shat[NSym<M>::value], ghat[NGH];
      if constexpr ((GM & 2) != 0) {
	// the site's generator: words (s0,s1) and (s2,s3) travel as two 64-bit values
	double * ns = P.noise + s;
	const double w01 = ns[0], w23 = ns[L.nsd];
	unsigned int w[4] = {(unsigned int) __double2loint(w01), (unsigned int) __double2hiint(w01),
			     (unsigned int) __double2loint(w23), (unsigned int) __double2hiint(w23)};
	fluctuations<M>(w, P, shat, ghat);
	if (v < nv) {
	  ns[0]     = __hiloint2double((int) w[1], (int) w[0]);
	  ns[L.nsd] = __hiloint2double((int) w[3], (int) w[2]);
	}
      }
      if (BINARY) {
	const double phi = P.phi[s];
	double dsq;
	double gr[3];
	if (P.grad != nullptr) {
	  dsq = P.delsq[s];
	  #pragma unroll
	  for (int a = 0; a < 3; a++) gr[a] = P.grad[a*L.nsd + s];
	}
	else {
	  // gradient not materialised: the 7-point stencil of
	  // src/gradient_3d_7pt_fluid.c:232-313 straight from phi (same operations in
	  // the same order, so the values are those grad7_kernel would store)
	  const double * ph = P.phi + s;
	  const double xp = __ldg(ph + op[0]), xm = __ldg(ph + om[0]);
	  const double yp = __ldg(ph + op[1]), ym = __ldg(ph + om[1]);
	  const double zq = __ldg(ph + zp[v]), zn = __ldg(ph + zm[v]);
	  gr[0] = 0.5*(xp - xm); gr[1] = 0.5*(yp - ym); gr[2] = 0.5*(zq - zn);
	  dsq = xp + xm + yp + ym + zq + zn - __dmul_rn(6.0, phi);
	}
	// fe_symm_str_v / fe_symm_mu (src/symmetric.c:371-420, :307-320)
	const double p0 = 0.5*P.A*phi*phi + 0.75*P.B*phi*phi*phi*phi
	  - P.kappa*phi*dsq
	  - 0.5*P.kappa*(gr[0]*gr[0] + gr[1]*gr[1] + gr[2]*gr[2]);
	double sth[NSym<M>::value];
	#pragma unroll
	for (int a = 0; a < ND; a++) {
	  #pragma unroll
	  for (int b = a; b < ND; b++) {
	    sth[sidx<M>(a, b)] = ((a == b) ? p0 : 0.0) + P.kappa*gr[a]*gr[b];
	  }
	}
	const double mu = P.A*phi + P.B*phi*phi*phi - P.kappa*dsq;
	mrt_site<M, GM, true>(f[0][v], F, sth, shat, ghat, P, P.rtau_shear, P.rtau_bulk, nullptr, rho, u);
	if constexpr (GMOM) {
	  double jin[ND];
	  #pragma unroll
	  for (int a = 0; a < ND; a++) jin[a] = __ldg(P.jphi + a*L.nsd + s);
	  phi_site_moment<M>(f[NDIST - 1][v], jin, phi, mu, u, P.rtau2);
	}
	else {
	  phi_site<M>(f[NDIST - 1][v], phi, mu, u, P.rtau2);
	}
      }
      else {
	if constexpr ((GM & 4) != 0) {   // viscosity model: its own instantiations (GM bit 2)
	  // lb_relaxation_time_shear_v / _bulk_v / _ghosts_v with the site's viscosity
	  const double eta = P.eta[s];
	  const double rts = 1.0/(0.5 + eta/P.rho0cs2);
	  const double rtb = (P.visc_relax == 1) ? rts : 1.0/(0.5 + (P.eta_ratio*eta)/P.rho0cs2);
	  double gk2[2] = {1.0 - rts, 1.0 - rts};
	  if (P.visc_relax == 2) {
	    const double tau = eta/P.rho0cs2;
	    gk2[1] = 1.0 - fmin(0.5 + 2.0*tau/(tau + 3.0/8.0), 2.0);
	  }
	  mrt_site<M, GM, false>(f[0][v], F, nullptr, shat, ghat, P, rts, rtb, gk2, rho, u);
	}
	else {
	  mrt_site<M, GM, false>(f[0][v], F, nullptr, shat, ghat, P, P.rtau_shear, P.rtau_bulk, nullptr, rho, u);
	}
      }
      if (DIAG && (v < nv)) {
	// conserved sums of the state that entered this collision: rho, rho^2, min,
	// max, momentum sum_p c_p f_p = rho u - F/2, and phi (binary)
	dq[0] += rho; dq[1] += rho*rho; dq[2] = fmin(dq[2], rho); dq[3] = fmax(dq[3], rho);
	#pragma unroll
	for (int a = 0; a < ND; a++) dq[4 + a] += rho*u[a] - 0.5*F[a];
	if (BINARY) dq[7] += P.phi[s];
      }
      if (P.u != nullptr && (v < nv)) {  // uniform branch: hydro stored this step
	if (!BINARY) P.rho[s] = rho;     // the binary collision leaves rho alone
	#pragma unroll
	for (int a = 0; a < 3; a++) P.u[a*L.nsd + s] = u[a];
      }
    }
  }

  if constexpr (AA == 2) {
    // odd step of the single-lattice scheme: every population goes to the site it streams
    // to (the mirror image of the pull above; same wrap-aware offsets with the roles of
    // the two directions exchanged)
    // decomposed directions (PUSH): a population that leaves the local box belongs to an
    // interior site of the neighbour -- stored there through the peer mapping -- and is
    // also left in the local halo site, where rho,u on demand look for it
    int ex[3] = {0, 0, 0};
    if constexpr (PUSH) {
      if (!P.wrap[0]) ex[0] = (i0 == 0) ? -1 : ((i0 == L.n[0] - 1) ? 1 : 0);
      if (!P.wrap[1]) ex[1] = (i1 == 0) ? -1 : ((i1 == L.n[1] - 1) ? 1 : 0);
    }
    #pragma unroll
    for (int p = 0; p < Q; p++) {
      double * dst = P.dst + (i64) p*L.nsd + site;
      if (dcv<M>(p, 0) ==  1) dst += op[0];
      if (dcv<M>(p, 0) == -1) dst += om[0];
      if (dcv<M>(p, 1) ==  1) dst += op[1];
      if (dcv<M>(p, 1) == -1) dst += om[1];
      const int c2 = dcv<M>(p, 2);
      if (c2 == 0) {
	if (VEC == 2 && nv == 2) *reinterpret_cast<double2 *>(dst) = make_double2(f[0][0][p], f[0][VEC - 1][p]);
	else if (nv >= 1) *dst = f[0][0][p];
      }
      else {
	#pragma unroll
	for (int v = 0; v < VEC; v++) {
	  if (v < nv) dst[v + (c2 == 1 ? zp[v] : zm[v])] = f[0][v][p];
	}
      }
      if constexpr (PUSH) {
	if (p > 0) {
	  const int s0 = (dcv<M>(p, 0) != 0 && ex[0] == dcv<M>(p, 0)) ? dcv<M>(p, 0) : 0;
	  const int s1 = (dcv<M>(p, 1) != 0 && ex[1] == dcv<M>(p, 1)) ? dcv<M>(p, 1) : 0;
	  #pragma unroll
	  for (int v = 0; v < VEC; v++) {
	    if (v < nv) {
	      int s2 = 0;
	      if (c2 != 0 && !P.wrap[2]) {
		const int e2 = (i2 + v == 0) ? -1 : ((i2 + v == L.n[2] - 1) ? 1 : 0);
		if (e2 == c2) s2 = c2;
	      }
	      if (s0 | s1 | s2) {
		double * peer = P.peer[(s0 + 1)*9 + (s1 + 1)*3 + (s2 + 1)];
		if (peer != nullptr) {
		  // the same site in the neighbour's frame: one box extent back along each
		  // crossed direction
		  const i64 at = (dst - P.dst) + v + (c2 == 0 ? 0 : (c2 == 1 ? zp[v] : zm[v]))
		    - (i64) s0*L.n[0]*L.s0 - (i64) s1*L.n[1]*L.s1 - (i64) s2*L.n[2];
		  peer[at] = f[0][v][p];
		}
	      }
	    }
	  }
	}
      }
    }
  }
  else {
  #pragma unroll
  for (int n = 0; n < NDIST; n++) {
    #pragma unroll
    for (int p = 0; p < Q; p++) {
      // AA 1 (even step): post-collision population p into the slot of its opposite
      const int pd = (AA == 1 && p > 0) ? Q - p : p;
      double * dst = P.dst + (i64)(n*Q + pd)*L.nsd + site;
      if (VEC == 2 && nv == 2) {
	*reinterpret_cast<double2 *>(dst) = make_double2(f[n][0][p], f[n][VEC - 1][p]);
      }
      else if (nv >= 1) {
	*dst = f[n][0][p];
      }
    }
  }
  }

  // (single-lattice streaming, even sweep: the same, into the swapped slot Q - p of the
  // neighbour's halo site, where its odd sweep pulls from; the odd sweep stores into the
  // neighbours' interior itself, above)
  // Peer-memory halo: a site on the face of the local box also stores the
  // populations that leave through that face directly into the halo site of the
  // neighbouring GPU's lattice (NVLink peer pointers, mapped with CUDA IPC), for
  // faces, edges and corners alike. This replaces pack kernel + NCCL + unpack kernel:
  // the transfer rides on the sweep itself. Same population set as the reduced
  // scheme (src/model.c:1192-1219): p goes to direction s iff s picks components of
  // c_p and the site lies on the boundary in each picked component.
  if constexpr (COLLIDE && PUSH && AA != 2) {
    // Two sites per thread, both on the same x or y face and neither at a row end: their
    // populations leave as ONE 128-bit store each, so a warp writes whole contiguous lines
    // into the neighbour. Site by site (below) the same data are 8-byte stores with a
    // 16-byte stride across the warp -- half-filled, byte-masked lines over NVLink.
    bool pair_done = false;
    if constexpr (VEC == 2) {
      if (nv == 2 && i2 > 0 && i2 + 1 < L.n[2] - 1) {
	const int e0 = (i0 == 0) ? -1 : ((i0 == L.n[0] - 1) ? 1 : 0);
	const int e1 = (i1 == 0) ? -1 : ((i1 == L.n[1] - 1) ? 1 : 0);
	if ((e0 != 0) != (e1 != 0)) {
	  pair_done = true;
	  #pragma unroll
	  for (int d = 0; d < 2; d++) {
	    const int sd = (d == 0) ? e0 : e1;
	    if (sd != 0) {
	      double * peer = P.peer[13 + sd*((d == 0) ? 9 : 3)];
	      if (peer != nullptr) {
		const i64 ext = (d == 0) ? (i64) L.n[0]*L.s0 : (i64) L.n[1]*L.s1;
		double * base = peer + site - sd*ext;
		#pragma unroll
		for (int n = 0; n < NDIST; n++) {
		  #pragma unroll
		  for (int p = 1; p < Q; p++) {
		    if (dcv<M>(p, d) == 1 || dcv<M>(p, d) == -1) {
		      if (dcv<M>(p, d) == sd) {
			*reinterpret_cast<double2 *>(base + (i64)(n*Q + ((AA == 1) ? Q - p : p))*L.nsd)
			  = make_double2(f[n][0][p], f[n][VEC - 1][p]);
		      }
		    }
		  }
		}
	      }
	    }
	  }
	}
      }
    }
    #pragma unroll
    for (int v = 0; v < VEC; v++) {
      if (v < nv && !pair_done) {
	int e[3];
	e[0] = (i0 == 0) ? -1 : ((i0 == L.n[0] - 1) ? 1 : 0);
	e[1] = (i1 == 0) ? -1 : ((i1 == L.n[1] - 1) ? 1 : 0);
	e[2] = (i2 + v == 0) ? -1 : ((i2 + v == L.n[2] - 1) ? 1 : 0);
	const int ne = (e[0] != 0) + (e[1] != 0) + (e[2] != 0);
	if (ne == 1) {
	  // A site on one face only (all but the edges of the box): the populations with
	  // that component go to ONE neighbour, at one offset. Six short branches with
	  // compile-time population lists instead of the general 7-way loop below --
	  // along the contiguous axis every row ends in such a site, so half of all warps
	  // come through here (profiles/r2_push_axes_before.log: 45 us per 256^3 step
	  // through the general loop).
	  #pragma unroll
	  for (int d = 0; d < 3; d++) {
	    if (e[d] != 0) {
	      const int sd = e[d];
	      const int kdir = 13 + sd*((d == 0) ? 9 : ((d == 1) ? 3 : 1));
	      double * peer = P.peer[kdir];
	      if (peer != nullptr) {
		const i64 ext = (d == 0) ? (i64) L.n[0]*L.s0 : ((d == 1) ? (i64) L.n[1]*L.s1 : (i64) L.n[2]);
		double * base = peer + site + v - sd*ext;
		if (d == 2 && AA == 0 && !BINARY && P.zwrite) {
		  // the neighbour's packed planes: consecutive rows are consecutive words
		  #pragma unroll
		  for (int p = 1; p < Q; p++) {
		    if (dcv<M>(p, 2) == 1 || dcv<M>(p, 2) == -1) {
		      if (dcv<M>(p, 2) == sd) {
			zstore(peer + P.zoff + zslot<M>(P, p, i0 + L.h[0], i1 + L.h[1]), f[0][v][p], P.zvec);
		      }
		    }
		  }
		}
		else if (d == 2 && P.zsec) {
		  // The contiguous axis: each row ends in ONE halo site, whose 32-byte sector
		  // holds nothing else but row padding. An 8-byte store leaves a partially
		  // written sector, which HBM (ECC) turns into a read-modify-write: 2 n_x n_y
		  // x 5 of them per step cost 39 us at 256^3 (profiles/r2_push_axes_after.log).
		  // Store the whole sector instead -- the halo value and three padding words
		  // nobody reads -- as one 256-bit store (also one NVLink packet, not a
		  // quarter-filled one).
		  #pragma unroll
		  for (int n = 0; n < NDIST; n++) {
		    #pragma unroll
		    for (int p = 1; p < Q; p++) {
		      if (dcv<M>(p, 2) == 1 || dcv<M>(p, 2) == -1) {
			if (dcv<M>(p, 2) == sd) {
			  double * a = base + (i64)(n*Q + ((AA == 1) ? Q - p : p))*L.nsd;
			  const double val = f[n][v][p];
			  if (sd == 1) {   // lands in the low halo: last word of its sector
#ifdef __CUDA_ARCH__
			    asm volatile("st.global.v4.f64 [%0], {%1, %1, %1, %2};" :: "l"(a - 3), "d"(0.0), "d"(val) : "memory");
#else
			    a[-3] = 0.0; a[-2] = 0.0; a[-1] = 0.0; a[0] = val;      // tools/push_emulate.cpp steps through this on the host
#endif
			  }
			  else {           // lands in the high halo: first word of its sector
#ifdef __CUDA_ARCH__
			    asm volatile("st.global.v4.f64 [%0], {%2, %1, %1, %1};" :: "l"(a), "d"(0.0), "d"(val) : "memory");
#else
			    a[0] = val; a[1] = 0.0; a[2] = 0.0; a[3] = 0.0;
#endif
			  }
			}
		      }
		    }
		  }
		}
		else if (sd == 1) {
		  #pragma unroll
		  for (int n = 0; n < NDIST; n++) {
		    #pragma unroll
		    for (int p = 1; p < Q; p++) {
		      if (dcv<M>(p, d) == 1) base[(i64)(n*Q + ((AA == 1) ? Q - p : p))*L.nsd] = f[n][v][p];
		    }
		  }
		}
		else {
		  #pragma unroll
		  for (int n = 0; n < NDIST; n++) {
		    #pragma unroll
		    for (int p = 1; p < Q; p++) {
		      if (dcv<M>(p, d) == -1) base[(i64)(n*Q + ((AA == 1) ? Q - p : p))*L.nsd] = f[n][v][p];
		    }
		  }
		}
	      }
	    }
	  }
	}
	else if (ne > 1) {
	  #pragma unroll
	  for (int n = 0; n < NDIST; n++) {
	    #pragma unroll
	    for (int p = 1; p < Q; p++) {
	      #pragma unroll
	      for (int m = 1; m < 8; m++) {
		// sub-direction m of c_p (compile time); skip picks of a zero component
		const int s0 = (m & 1) ? dcv<M>(p, 0) : 0;
		const int s1 = (m & 2) ? dcv<M>(p, 1) : 0;
		const int s2 = (m & 4) ? dcv<M>(p, 2) : 0;
		const bool valid = (!(m & 1) || dcv<M>(p, 0) != 0) && (!(m & 2) || dcv<M>(p, 1) != 0)
		  && (!(m & 4) || dcv<M>(p, 2) != 0);
		if (valid) {
		  const bool on = (s0 == 0 || e[0] == s0) && (s1 == 0 || e[1] == s1) && (s2 == 0 || e[2] == s2);
		  double * peer = P.peer[(s0 + 1)*9 + (s1 + 1)*3 + (s2 + 1)];
		  if (on && peer != nullptr) {
		    if (s2 != 0 && AA == 0 && !BINARY && P.zwrite) {
		      // a halo site beyond a z face of that neighbour (an edge or corner of its box)
		      zstore(peer + P.zoff + zslot<M>(P, p, i0 + L.h[0] - s0*L.n[0], i1 + L.h[1] - s1*L.n[1]), f[n][v][p], P.zvec);
		    }
		    else {
		      const i64 rs = site + v - (i64) s0*L.n[0]*L.s0 - (i64) s1*L.n[1]*L.s1 - (i64) s2*L.n[2];
		      peer[(i64)(n*Q + ((AA == 1) ? Q - p : p))*L.nsd + rs] = f[n][v][p];
		    }
		  }
		}
	      }
	    }
	  }
	}
      }
    }
  }

  if (DIAG) {
    const i64 b = ((i64) blockIdx.z*gridDim.y + blockIdx.y)*gridDim.x + blockIdx.x;
    block_reduce8(dq, P.diag + 8*(P.diag_base + b));
  }
}

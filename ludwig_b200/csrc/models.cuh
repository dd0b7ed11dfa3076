// models.cuh -- velocity sets as compile-time tables.
//
// Ordering and weights are interface facts of the reference (host arrays are
// indexed by p): D3Q19 src/lb_d3q19.h:27-40, D3Q27 src/lb_d3q27.h:27-43,
// D2Q9 src/lb_d2q9.h:26-35. All sets satisfy cv[nvel-p] = -cv[p]
// (tests/unit/test_lb_model.c:103-377), which the kernels exploit by working
// on (p, nvel-p) pairs.
//
// Everything is a constexpr FUNCTION with a function-local table, so that fully
// unrolled device loops fold the entries into immediates (namespace-scope
// constexpr arrays are not addressable from device code).

#pragma once

#define HD __host__ __device__ __forceinline__

struct D3Q19 {
  static constexpr int NVEL = 19;
  static constexpr int NDIM = 3;
  static constexpr int NHYDRO = 10;
  HD static constexpr int cv(int p, int a) {
    constexpr int t[19][3] = {
      { 0, 0, 0},
      { 1, 1, 0}, { 1, 0, 1}, { 1, 0, 0}, { 1, 0,-1}, { 1,-1, 0}, { 0, 1, 1},
      { 0, 1, 0}, { 0, 1,-1}, { 0, 0, 1}, { 0, 0,-1}, { 0,-1, 1}, { 0,-1, 0},
      { 0,-1,-1}, {-1, 1, 0}, {-1, 0, 1}, {-1, 0, 0}, {-1, 0,-1}, {-1,-1, 0}};
    return t[p][a];
  }
  HD static constexpr double wv(int p) {
    const int c2 = cv(p,0)*cv(p,0) + cv(p,1)*cv(p,1) + cv(p,2)*cv(p,2);
    return c2 == 0 ? 12.0/36.0 : (c2 == 1 ? 2.0/36.0 : 1.0/36.0);
  }
  // device axis d (0 slowest .. 2 contiguous) -> physical axis
  HD static constexpr int dax(int d) { return d; }
};

struct D3Q27 {
  static constexpr int NVEL = 27;
  static constexpr int NDIM = 3;
  static constexpr int NHYDRO = 10;
  // p = 0 rest, then lexicographic (x,y,z) over {-1,0,1}^3 without the origin
  HD static constexpr int cv(int p, int a) {
    if (p == 0) return 0;
    const int i = p - 1;
    const int q = (i < 13) ? i : i + 1;
    return (a == 0) ? q/9 - 1 : ((a == 1) ? (q/3)%3 - 1 : q%3 - 1);
  }
  HD static constexpr double wv(int p) {
    const int c2 = cv(p,0)*cv(p,0) + cv(p,1)*cv(p,1) + cv(p,2)*cv(p,2);
    return c2 == 0 ? 64.0/216.0
         : (c2 == 1 ? 16.0/216.0 : (c2 == 2 ? 4.0/216.0 : 1.0/216.0));
  }
  HD static constexpr int dax(int d) { return d; }
};

struct D2Q9 {
  static constexpr int NVEL = 9;
  static constexpr int NDIM = 2;
  static constexpr int NHYDRO = 6;
  HD static constexpr int cv(int p, int a) {
    constexpr int t[9][3] = {
      { 0, 0, 0},
      { 1, 1, 0}, { 1, 0, 0}, { 1,-1, 0}, { 0, 1, 0},
      { 0,-1, 0}, {-1, 1, 0}, {-1, 0, 0}, {-1,-1, 0}};
    return t[p][a];
  }
  HD static constexpr double wv(int p) {
    const int c2 = cv(p,0)*cv(p,0) + cv(p,1)*cv(p,1);
    return c2 == 0 ? 16.0/36.0 : (c2 == 1 ? 4.0/36.0 : 1.0/36.0);
  }
  // A 2-d lattice (nlocal[Z] = 1) is held with y contiguous: device axes
  // (slow, mid, fast) = physical (z, x, y); the z extent is 1 with no halo.
  HD static constexpr int dax(int d) { return d == 0 ? 2 : d - 1; }
};

// velocity component along device axis d
template <class M> HD constexpr int dcv(int p, int d) { return M::cv(p, M::dax(d)); }

// index of (a,b), a <= b, in the symmetric-stress storage used by the kernels:
// 3-d (xx,xy,xz,yy,yz,zz), 2-d (xx,xy,yy) -- the order of the reference's
// hydrodynamic modes (src/lb_d3q19.c:112-150, src/lb_d2q9.c:137-168).
template <class M> HD constexpr int sidx(int a, int b) {
  return (M::NDIM == 3) ? (a == 0 ? b : (a == 1 ? 2 + b : 5)) : (a == 0 ? b : 2);
}
template <class M> struct NSym { static constexpr int value = M::NDIM*(M::NDIM + 1)/2; };

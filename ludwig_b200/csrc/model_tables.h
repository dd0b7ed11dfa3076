// model_tables.h -- host-side construction of the ghost-mode projectors.
//
// The reference builds ma (modes from f) per model in src/lb_d3q19.c:112-150,
// src/lb_d3q27.c:152-195, src/lb_d2q9.c:137-168, normalisers
// na[m] = 1/sum_p w_p ma[m][p]^2, and mi[p][m] = w_p na[m] ma[m][p]
// (src/model.c:358-390). The kernels project the hydrodynamic modes in closed
// form; only the ghost (kinetic) rows are tabulated, for the relaxation schemes
// that keep a fraction of them (BGK, TRT).

#pragma once

#include "layout.cuh"

// polynomial rows of the mode matrix for ghost index k (0-based after the
// hydrodynamic modes) at velocity c
static inline double ghost_row(int nvel, int k, const int c[3]) {
  const double cs2 = 1.0/3.0;
  const double cx = c[0], cy = c[1], cz = c[2];
  const double qxx = cx*cx - cs2, qyy = cy*cy - cs2, qzz = cz*cz - cs2;
  if (nvel == 9) {
    switch (k) {
    case 0: return 9.0*qxx*qyy;
    case 1: return 6.0*qyy*cx;
    default: return 6.0*qxx*cy;
    }
  }
  if (nvel == 19) {
    const double c2 = cx*cx + cy*cy + cz*cz;
    const double chi1 = (2.0*c2 - 3.0)*(3.0*cz*cz - c2);
    const double chi2 = (2.0*c2 - 3.0)*(cy*cy - cx*cx);
    const double chi3 = 3.0*c2*c2 - 6.0*c2 + 1.0;
    switch (k) {
    case 0: return chi1;
    case 1: return chi1*cx;
    case 2: return chi1*cy;
    case 3: return chi1*cz;
    case 4: return chi2;
    case 5: return chi2*cx;
    case 6: return chi2*cy;
    case 7: return chi2*cz;
    default: return chi3;
    }
  }
  switch (k) {      /* D3Q27: third and higher order Hermite polynomials */
  case 0: return 3.0*qxx*cy;
  case 1: return 3.0*qxx*cz;
  case 2: return 3.0*qyy*cz;
  case 3: return 3.0*qyy*cx;
  case 4: return 3.0*qzz*cx;
  case 5: return 3.0*qzz*cy;
  case 6: return cx*cy*cz;
  case 7: return 9.0*qxx*qyy;
  case 8: return 9.0*qyy*qzz;
  case 9: return 9.0*qzz*qxx;
  case 10: return 9.0*qxx*cy*cz;
  case 11: return 9.0*qyy*cz*cx;
  case 12: return 9.0*qzz*cx*cy;
  case 13: return 9.0*qxx*qyy*cz;
  case 14: return 9.0*qyy*qzz*cx;
  case 15: return 9.0*qzz*qxx*cy;
  default: return 27.0*qxx*qyy*qzz;
  }
}

template <class M>
static inline void build_ghost_tab(GhostTab & T) {
  const int Q = M::NVEL, NG = Q - M::NHYDRO;
  for (int k = 0; k < B200LB_MAXGHOST; k++) {
    for (int p = 0; p < 27; p++) { T.fwd[k][p] = 0.0; T.bwd[p][k] = 0.0; }
  }
  for (int k = 0; k < NG; k++) {
    double wip = 0.0;
    double row[27];
    for (int p = 0; p < Q; p++) {
      const int c[3] = {M::cv(p, 0), M::cv(p, 1), M::cv(p, 2)};
      row[p] = ghost_row(Q, k, c);
      wip += M::wv(p)*row[p]*row[p];
    }
    const double na = 1.0/wip;
    for (int p = 0; p < Q; p++) {
      T.fwd[k][p] = row[p];
      T.bwd[p][k] = M::wv(p)*na*row[p];
    }
  }
  // The reference's unrolled D3Q19 forward transform multiplies f[4] by zero in
  // mode 13 where the matrix entry is -1 (src/collision.c:2300). It is part of
  // the reference's BGK/TRT results, so it is reproduced for parity.
  if (Q == 19) T.fwd[13 - M::NHYDRO][4] = 0.0;
}

// layout.cuh -- the private device layout and the kernel parameter blocks.
//
// Device layout ("site-contiguous SoA"): one plane of doubles per population,
//   f[(n*nvel + p)*nsd + site],  site = i0*s0 + i1*s1 + z0 + i2
// with (i0,i1,i2) halo-inclusive 0-based device coordinates, i2 contiguous.
// Rows are padded so that the first interior site of every row sits on a
// multiple of `zalign` doubles (z0 = zalign - h2): 128-bit accesses of site
// pairs are then always aligned and rows start on 128-byte lines.
// For 3-d models device axes = physical (x,y,z); for D2Q9 they are (z,x,y) with
// the unit z extent first (see models.cuh), so the contiguous axis is always long.
//
// The reference's layouts are src/memory.h:124-130 (AOS) and :187 (SOA) with int
// indices; here every index is 64-bit.

#pragma once

typedef long long i64;

struct Layout {
  int n[3];     // interior extent per device axis
  int h[3];     // halo width per device axis (1, or 0 for the unit axis of a 2-d lattice)
  int na[3];    // n + 2h
  int z0;       // padding in front of each row
  i64 s0, s1;   // strides of device axes 0 and 1 (axis 2 has stride 1)
  i64 nsd;      // doubles per population plane
};

#define B200LB_MAXGHOST 17
#define B200LB_MAXNOISE 10     // variates per draw of the site generator (NNOISE_MAX, src/noise.h:18)

// Per-launch parameters (passed by value: they live in the constant bank).
struct KParams {
  Layout L;
  const double * src;     // lattice read
  double * dst;           // lattice written (== src for the in-place collision)
  double * rho;           // [nsd]
  double * u;             // [3][nsd]
  const double * force;   // [3][nsd] or nullptr when the force field is uniformly fzero
  const double * phi;     // [nsd]
  const double * grad;    // [3][nsd]
  const double * delsq;   // [nsd]
  int push;               // write outgoing populations straight into the neighbours' halo sites
  int zsec;               // the halo site at either end of a row has a 32-byte sector of its own (rest: padding)
  // Packed halo of the contiguous axis (peer-memory sweeps, single fluid, two lattices): the
  // populations that cross a z face do not go to the row-end halo words -- one isolated
  // 32-byte sector per row and population on either side -- but to dense planes behind the
  // lattice's last population plane, [face][k][x][y] over the halo-inclusive x-y extent,
  // face 0 = low halo (populations with c_z = +1, k-th of them in index order), face 1 = high
  i64 zoff;               // doubles from a lattice's base to its packed planes; 0 = not in use
  int zread;              // the source lattice's z-face halo is in its packed planes
  int zwrite;             // this sweep pushes across z faces into the neighbours' packed planes
  i64 zsx, zsy, zsp;      // strides (doubles) of the halo-inclusive x and y column index and of a plane
  int zvec;               // a column owns a whole 32-byte sector (its value + three padding words), stored as one
  double * peer[27];      // neighbour's destination lattice per direction (dx+1)*9+(dy+1)*3+(dz+1), or null
  double * diag;          // per-block partial sums of the in-kernel diagnostics, or unused
  i64 diag_base;          // first partial slot of this launch
  int lo[3], hi[3];       // sub-box of the interior, 0-based interior coordinates, [lo,hi)
  int wrap[3];            // device axis is periodic within this rank: pulls wrap instead of reading halos
  double rtau_shear, rtau_bulk;
  // viscosity model (src/collision.c:386-404): per-site shear viscosity, or nullptr
  const double * jphi;    // [ndim][nsd] first moment of the pulled g (sweep with GM bit 3)
  const double * eta;     // [nsd] hydro->eta
  double rho0cs2;         // rho0 cs^2: rtau = 1/(1/2 + eta/(rho0 cs^2))
  double eta_ratio;       // (eta_bulk/eta_shear) of the Newtonian fluid
  int visc_relax;         // relaxation scheme, for the ghost rates that follow eta (BGK, TRT)
  double fg[3];           // uniform part of the force: force_global + constant field value
  double rtau2, A, B, kappa;
  double gkeep[B200LB_MAXGHOST];   // 1 - rtau_k for each ghost mode
  // flat no-slip walls (src/wall.c): the wall occupies the halo layer of a non-periodic
  // direction; populations a pull would fetch from there are this site's own reversed
  // post-collision populations (+ the moving-wall term): wall_bbl folded into the pull
  int wall_lo[3], wall_hi[3];   // per device axis: this rank holds the low / high wall
  double wall_corr[27];         // per incoming population q: -2 rcs2 w rho0 (c_{Q-q} . u_wall)
  // partial slip (wall_bbl_slip_kernel, src/wall.c:1118-1210): fraction of specular
  // reflection per wall; no wall velocity then
  int wall_slip;
  double wall_s_lo[3], wall_s_hi[3];   // per device axis
  signed char wall_q[27][8];    // velocity index of c_p with the device-axis components in
                                // the bit mask reversed (mirror image in those walls)
  // isothermal fluctuations (src/noise.c, src/collision.c:476-515,1663-1745)
  double * noise;         // [2][nsd]: KISS99 words (s0,s1) and (s2,s3) of every site as 64-bit pairs
  double nv_diag, nv_off; // stress variance: var_eta sqrt(2) on the diagonal, var_eta off it
  double nv_bulk;         // variance of the trace
  double nv_ghost[B200LB_MAXNOISE];  // per ghost mode
  int nghost_draw;        // 0 = ghost modes off: no second draw, no ghost noise
  double nlv_big, nlv_small;         // sqrt(2 + sqrt 2), sqrt(2 - sqrt 2): the discrete levels
};

// Ghost-mode projectors (host-built, one __constant__ copy per model).
struct GhostTab {
  double fwd[B200LB_MAXGHOST][27];   // m_k = sum_p fwd[k][p] f_p
  double bwd[27][B200LB_MAXGHOST];   // f_p += sum_k bwd[p][k] m_k
};

"""ludwig_b200 -- B200 (sm_100a) lattice-Boltzmann time-step engine behind Ludwig's
call surface. The compute path is libb200lb.so (hand-written CUDA, C ABI in
include/b200lb.h); this package only loads it and mirrors the reference's function
names for tests and benchmarks. There is no CPU fallback: importing works anywhere,
creating a lattice needs a CUDA device and the built library.
"""

from .lb import (Lattice, Param, B200LBError, build, library_path, load,  # noqa: F401
                 pinned_empty, pinned_free,
                 HOST_AOS, HOST_SOA, HOST_AOSOA, RELAX_M10, RELAX_BGK, RELAX_TRT,
                 HALO_FULL, HALO_REDUCED, FIELD_RHO, FIELD_U, FIELD_FORCE,
                 FIELD_PHI, FIELD_GRAD, FIELD_DELSQ, FIELD_ETA, ABI_SYMBOLS,
                 FORCE_NONE, FORCE_STRESS_DIVERGENCE, FORCE_PHI_GRADMU)
from .decomp import Decomposition  # noqa: F401

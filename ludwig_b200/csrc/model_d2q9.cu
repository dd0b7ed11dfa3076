// model_d2q9.cu -- D2Q9 instantiation of the sweep kernels (see model_impl.cuh)
#define B200LB_MODEL D2Q9
#define B200LB_OPS_FN model_ops_d2q9
#include "model_impl.cuh"

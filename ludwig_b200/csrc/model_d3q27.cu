// model_d3q27.cu -- D3Q27 instantiation of the sweep kernels (see model_impl.cuh)
#define B200LB_MODEL D3Q27
#define B200LB_OPS_FN model_ops_d3q27
#include "model_impl.cuh"

// model_d3q19.cu -- D3Q19 instantiation of the sweep kernels (see model_impl.cuh)
#define B200LB_MODEL D3Q19
#define B200LB_OPS_FN model_ops_d3q19
#include "model_impl.cuh"

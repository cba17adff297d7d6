// pipelined NLL kernels (smoothing-parameter grid search), float
#include "eksb_nll_pipe.cuh"
template int eksb::kf_nll_pipe<float>(const float*, const float*, const float*, const float*, const float*, const float*, int, long, long, int, int, int, const float*, int, double*, int*, const float*, cudaStream_t);

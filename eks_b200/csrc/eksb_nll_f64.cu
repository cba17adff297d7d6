// pipelined NLL kernels (smoothing-parameter grid search), double
#include "eksb_nll_pipe.cuh"
template int eksb::kf_nll_pipe<double>(const double*, const double*, const double*, const double*, const double*, const double*, int, long, long, int, int, int, const double*, int, double*, int*, const double*, cudaStream_t);

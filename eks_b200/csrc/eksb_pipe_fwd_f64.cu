// warp-specialised fwd kernels, double
#define EKSB_PIPE_FWD 1
#include "eksb_pipe_impl.cuh"
template int eksb::kf_forward_pipe<double>(const double*, const double*, const double*, const double*, const double*, const double*, int, long, long, int, int, int, int, double*, int*, const double*, cudaStream_t);

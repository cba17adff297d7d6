// warp-specialised bwd kernels, double
#define EKSB_PIPE_BWD 1
#include "eksb_pipe_impl.cuh"
template int eksb::rts_backward_pipe<double>(const double*, const double*, const double*, const double*, int, long, long, int, int, int, double*, double*, double*, int*, cudaStream_t);
template int eksb::rts_backward_pipe_vt<double>(const double*, const double*, const double*, const double*, int, long, long, int, int, int, double*, double*, int*, eksb::VTiles, cudaStream_t);

// warp-specialised bwd kernels, float
#define EKSB_PIPE_BWD 1
#include "eksb_pipe_impl.cuh"
template int eksb::rts_backward_pipe<float>(const float*, const float*, const float*, const float*, int, long, long, int, int, int, float*, float*, float*, int*, cudaStream_t);
template int eksb::rts_backward_pipe_vt<float>(const float*, const float*, const float*, const float*, int, long, long, int, int, int, float*, float*, int*, eksb::VTiles, cudaStream_t);

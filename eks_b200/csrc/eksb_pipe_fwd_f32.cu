// warp-specialised fwd kernels, float
#define EKSB_PIPE_FWD 1
#include "eksb_pipe_impl.cuh"
template int eksb::kf_forward_pipe<float>(const float*, const float*, const float*, const float*, const float*, const float*, int, long, long, int, int, int, int, float*, int*, const float*, cudaStream_t);
template int eksb::kf_forward_pipe_vt<float>(const float*, const float*, const float*, const float*, const float*, const float*, int, long, long, int, int, int, int, float*, int*, const float*, eksb::VTiles, cudaStream_t);

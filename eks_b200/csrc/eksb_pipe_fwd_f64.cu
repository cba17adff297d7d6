// warp-specialised fwd kernels, double
#define EKSB_PIPE_FWD 1
#include "eksb_pipe_impl.cuh"
template int eksb::kf_forward_pipe<double>(const double*, const double*, const double*, const double*, const double*, const double*, int, long, long, int, int, int, int, double*, int*, const double*, cudaStream_t);
template int eksb::kf_forward_pipe_vt<double>(const double*, const double*, const double*, const double*, const double*, const double*, int, long, long, int, int, int, int, double*, int*, const double*, eksb::VTiles, cudaStream_t);
size_t eksb::pipe_vt_target_tiles(int n, int p, int elem_size) { return eksb::pipe_vt_target_tiles_impl(n, p, elem_size); }

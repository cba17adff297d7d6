// one dtype x one pass per translation unit so `make -j` compiles them in parallel
#define EKSB_ONLY_BWD 1
#include "eksb_kf.cuh"
using namespace eksb;

extern "C" int eksb_rts_backward_f32(const float* filt, const float* C, const float* A, const float* Q, int shared_mask,
                                     long B, long T, int n, int p, int a_mode, float* smth, float* ym, float* yv,
                                     float* cv, int* status, void* stream) {
  return rts_backward_api<float>(filt, C, A, Q, shared_mask, B, T, n, p, a_mode, smth, ym, yv, cv, status, stream);
}

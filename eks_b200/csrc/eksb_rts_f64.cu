// one dtype x one pass per translation unit so `make -j` compiles them in parallel
#define EKSB_ONLY_BWD 1
#include "eksb_kf.cuh"
using namespace eksb;

extern "C" int eksb_rts_backward_f64(const double* filt, const double* C, const double* A, const double* Q,
                                     int shared_mask, long B, long T, int n, int p, int a_mode, double* smth,
                                     double* ym, double* yv, double* cv, int* status, void* stream) {
  return rts_backward_api<double>(filt, C, A, Q, shared_mask, B, T, n, p, a_mode, smth, ym, yv, cv, status, stream);
}

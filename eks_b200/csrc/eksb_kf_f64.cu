// one dtype x one pass per translation unit so `make -j` compiles them in parallel
#define EKSB_ONLY_FWD 1
#include "eksb_kf.cuh"
using namespace eksb;

extern "C" int eksb_kf_forward_f64(const double* obs, const double* m0, const double* S0, const double* C,
                                   const double* A, const double* Q, int shared_mask, long B, long T, int n, int p,
                                   int a_mode, int first, const double* q_scale, int n_scale, double* filt,
                                   double* nll, int* status, void* stream) {
  return kf_forward_api<double>(obs, m0, S0, C, A, Q, shared_mask, B, T, n, p, a_mode, first, q_scale, n_scale, filt, nll, status, stream);
}

extern "C" int eksb_kf_forward_offset_f64(const double* obs, const double* y_offset, const double* m0, const double* S0, const double* C,
                                          const double* A, const double* Q, int shared_mask, long B, long T, int n, int p,
                                          int a_mode, int first, double* filt, double* nll, int* status, void* stream) {
  return kf_forward_api<double>(obs, m0, S0, C, A, Q, shared_mask, B, T, n, p, a_mode, first, nullptr, 0, filt, nll, status, stream, y_offset);
}

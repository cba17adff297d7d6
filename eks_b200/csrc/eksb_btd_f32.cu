// A7 block-tridiagonal solve, float
#define EKSB_PIPE_BTD 1
#include "eksb_pipe_impl.cuh"
using namespace eksb;

extern "C" int eksb_btd_solve_f32(const float* obs, const float* m0, const float* S0, const float* A, const float* Bmat, const float* E,
                                  int shared_mask, long B, long T, int n, int p, int a_mode, float* work, float* sol, float* ym,
                                  int* status, void* stream) {
  if (!obs || !m0 || !S0 || !A || !Bmat || !E || !work || !sol || B <= 0 || T <= 0) { set_error("btd_solve: null pointer or empty batch"); return EKSB_EARG; }
  if (a_mode != EKSB_A_DIAG && a_mode != EKSB_A_DENSE) { set_error("btd_solve: bad a_mode %d", a_mode); return EKSB_EARG; }
  return btd_solve_grouped<float>(obs, m0, S0, Bmat, A, E, shared_mask, B, T, n, p, a_mode, work, sol, ym, status, (cudaStream_t)stream);
}

// A7 block-tridiagonal solve, double
#define EKSB_PIPE_BTD 1
#include "eksb_pipe_impl.cuh"
using namespace eksb;

extern "C" int eksb_btd_solve_f64(const double* obs, const double* m0, const double* S0, const double* A, const double* Bmat, const double* E,
                                  int shared_mask, long B, long T, int n, int p, int a_mode, double* work, double* sol, double* ym,
                                  int* status, void* stream) {
  if (!obs || !m0 || !S0 || !A || !Bmat || !E || !work || !sol || B <= 0 || T <= 0) { set_error("btd_solve: null pointer or empty batch"); return EKSB_EARG; }
  if (a_mode != EKSB_A_DIAG && a_mode != EKSB_A_DENSE) { set_error("btd_solve: bad a_mode %d", a_mode); return EKSB_EARG; }
  return btd_solve_grouped<double>(obs, m0, S0, Bmat, A, E, shared_mask, B, T, n, p, a_mode, work, sol, ym, status, (cudaStream_t)stream);
}

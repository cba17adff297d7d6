// Host build of the per-thread programs in eks_b200/csrc/eksb_math.cuh.
// TEST INFRASTRUCTURE ONLY: lets the CPU-only build box check the exact code the
// CUDA kernels run (same templates, lane stride 1) against the numpy oracle.
// Never loaded by the eks_b200 package.
#include <vector>
#include <cstring>
#include "../../eks_b200/csrc/eksb_math.cuh"
#include "../../eks_b200/csrc/eksb_dims.h"

using namespace eksb;

template <typename T, int N, int P, int AM>
static int run_case(long Tn, const T* y, const T* ev, const T* m0, const T* S0, const T* C, const T* A, const T* Q,
                    int first, T* filt, T* smth, T* ym, T* yv, T* cv, double* nll) {
  constexpr int NS = Dim<N>::NS;
  Params<T, N, P, AM> prm;
  load_params<T, N, P, AM>(prm, C, A, Q);
  std::vector<T> obs(Tn * 2 * P);
  for (long t = 0; t < Tn; ++t)
    for (int j = 0; j < P; ++j) { obs[t * 2 * P + j] = y[t * P + j]; obs[t * 2 * P + P + j] = ev[t * P + j]; }
  T m[N], V[NS];
  for (int i = 0; i < N; ++i) m[i] = m0[i];
  load_sym<T, N>(V, S0);
  double acc = 0.0;
  bool ok = kf_forward_range<T, N, P, AM, 1, 3, true>(prm, obs.data(), filt, Tn, first, m, V, acc);
  *nll = acc;
  if (smth) {
    T ms[N], Vs[NS];
    ok = rts_backward_range<T, N, P, AM, 1, 3>(prm, filt, smth, ym, yv, cv, Tn, true, ms, Vs) && ok;
  }
  return ok ? 0 : 1;
}

template <typename T>
static int dispatch(int n, int p, int am, long Tn, const T* y, const T* ev, const T* m0, const T* S0, const T* C,
                    const T* A, const T* Q, int first, T* filt, T* smth, T* ym, T* yv, T* cv, double* nll) {
#define X(N_, P_)                                                                                             \
  if (n == N_ && p == P_)                                                                                     \
    return am ? run_case<T, N_, P_, 1>(Tn, y, ev, m0, S0, C, A, Q, first, filt, smth, ym, yv, cv, nll)        \
              : run_case<T, N_, P_, 0>(Tn, y, ev, m0, S0, C, A, Q, first, filt, smth, ym, yv, cv, nll);
  EKSB_FOR_EACH_DIM(X)
#undef X
  return -1;
}

extern "C" int hs_kf_rts_f64(int n, int p, int am, long Tn, const double* y, const double* ev, const double* m0,
                             const double* S0, const double* C, const double* A, const double* Q, int first,
                             double* filt, double* smth, double* ym, double* yv, double* cv, double* nll) {
  return dispatch<double>(n, p, am, Tn, y, ev, m0, S0, C, A, Q, first, filt, smth, ym, yv, cv, nll);
}
extern "C" int hs_kf_rts_f32(int n, int p, int am, long Tn, const float* y, const float* ev, const float* m0,
                             const float* S0, const float* C, const float* A, const float* Q, int first,
                             float* filt, float* smth, float* ym, float* yv, float* cv, double* nll) {
  return dispatch<float>(n, p, am, Tn, y, ev, m0, S0, C, A, Q, first, filt, smth, ym, yv, cv, nll);
}

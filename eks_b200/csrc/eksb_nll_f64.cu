// pipelined NLL kernels (smoothing-parameter grid search), double
#include "eksb_nll_pipe.cuh"
template int eksb::kf_nll_pipe<double>(const double*, const double*, const double*, const double*, const double*, const double*, int, long, long, int, int, int, const double*, int, double*, int*, const double*, cudaStream_t);

using namespace eksb;

extern "C" size_t eksb_nll_workspace_bytes(long B, long T, int n, int p, int n_scale, int elem_size) {
  if (elem_size != 4 && elem_size != 8) return 0;
  return nll_workspace_bytes(B, T, n, p, n_scale, elem_size);
}

extern "C" int eksb_kf_nll_ws_f64(const double* obs, const double* y_offset, const double* m0, const double* S0, const double* C, const double* A,
                                  const double* Q, int shared_mask, long B, long T, int n, int p, int a_mode, int first,
                                  const double* q_scale, int n_scale, double* nll, int* status, void* workspace,
                                  size_t workspace_bytes, void* stream) {
  if (!obs || !m0 || !S0 || !C || !A || !Q || !nll || B <= 0 || T <= 0) { set_error("kf_nll: null pointer or empty batch"); return EKSB_EARG; }
  if (first != kPriorUpdate && first != kPriorOnly) { set_error("kf_nll: bad first-frame mode %d", first); return EKSB_EARG; }
  if (q_scale && n_scale < 1) { set_error("kf_nll: n_scale < 1"); return EKSB_EARG; }
  if (a_mode == EKSB_A_DIAG && nll_split_dims_ok(n, p) && use_pipe(B)) {
    const char* e = getenv("EKSB_NLL_PIPE");
    if (!(e && e[0] == '0'))
      return kf_nll_split<double>(obs, y_offset, m0, S0, C, A, Q, shared_mask, B, T, n, p, first, q_scale, q_scale ? n_scale : 0, nll,
                             status, workspace, workspace_bytes, (cudaStream_t)stream);
  }
  // dense A, other shapes, EKSB_PIPE=0 / EKSB_NLL_PIPE=0: the sequential entry points
  if (!y_offset) return eksb_kf_forward_f64(obs, m0, S0, C, A, Q, shared_mask, B, T, n, p, a_mode, first, q_scale, q_scale ? n_scale : 0, nullptr, nll, status, stream);
  if (!q_scale) return eksb_kf_forward_offset_f64(obs, y_offset, m0, S0, C, A, Q, shared_mask, B, T, n, p, a_mode, first, nullptr, nll, status, stream);
  set_error("kf_nll: y_offset with q_scale needs the pipelined kernel (diagonal A, (n,p) in (2,4),(3,4),(3,6),(3,8))");
  return EKSB_EDIM;
}

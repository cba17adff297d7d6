// K4 parallel-in-time scan, double
#include "eksb_scan_impl.cuh"
using namespace eksb;

extern "C" int eksb_kf_rts_scan_f64(const double* obs, const double* m0, const double* S0, const double* C, const double* A, const double* Q,
                                   int shared_mask, long B, long T, int n, int p, int a_mode, long chunk, double* workspace,
                                   size_t workspace_bytes, double* filt, double* smth, double* ym, int* status, void* stream) {
  return scan_api<double>(obs, m0, S0, C, A, Q, shared_mask, B, T, n, p, a_mode, chunk, workspace, workspace_bytes, filt, smth, ym, status, stream);
}

extern "C" size_t eksb_scan_workspace_bytes(long B, long T, int n, long chunk, int elem_size) {
  if (B <= 0 || T <= 0 || n <= 0) return 0;
  return scan_workspace_elems(B, T, n, chunk > 0 ? chunk : scan_auto_chunk(B, T, n, elem_size)) * (size_t)elem_size;
}
extern "C" long eksb_scan_chunk_len(long B, long T, int n, long chunk, int elem_size) { return chunk > 0 ? chunk : scan_auto_chunk(B, T, n, elem_size); }

extern "C" int eksb_kf_forward_ws_f64(const double* obs, const double* y_offset, const double* m0, const double* S0, const double* C,
                                     const double* A, const double* Q, int shared_mask, long B, long T, int n, int p, int a_mode,
                                     int first, double* filt, double* nll, int* status, double* workspace, size_t workspace_bytes,
                                     void* stream) {
  return forward_ws_api<double>(obs, y_offset, m0, S0, C, A, Q, shared_mask, B, T, n, p, a_mode, first, filt, nll, status, workspace,
                              workspace_bytes, stream);
}

extern "C" int eksb_rts_backward_ws_f64(const double* filt, const double* C, const double* A, const double* Q, int shared_mask, long B,
                                       long T, int n, int p, int a_mode, double* smth, double* ym, double* yv, double* cv, int* status,
                                       double* workspace, size_t workspace_bytes, void* stream) {
  return backward_ws_api<double>(filt, C, A, Q, shared_mask, B, T, n, p, a_mode, smth, ym, yv, cv, status, workspace, workspace_bytes,
                               stream);
}

extern "C" size_t eksb_route_workspace_bytes(long B, long T, int n, int p, int elem_size) { return route_workspace_bytes(B, T, n, p, elem_size); }

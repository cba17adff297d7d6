// K4 parallel-in-time scan, float
#include "eksb_scan_impl.cuh"
using namespace eksb;

extern "C" int eksb_kf_rts_scan_f32(const float* obs, const float* m0, const float* S0, const float* C, const float* A, const float* Q,
                                   int shared_mask, long B, long T, int n, int p, int a_mode, long chunk, float* workspace,
                                   size_t workspace_bytes, float* filt, float* smth, float* ym, int* status, void* stream) {
  return scan_api<float>(obs, m0, S0, C, A, Q, shared_mask, B, T, n, p, a_mode, chunk, workspace, workspace_bytes, filt, smth, ym, status, stream);
}

extern "C" int eksb_kf_forward_ws_f32(const float* obs, const float* y_offset, const float* m0, const float* S0, const float* C,
                                     const float* A, const float* Q, int shared_mask, long B, long T, int n, int p, int a_mode,
                                     int first, float* filt, double* nll, int* status, float* workspace, size_t workspace_bytes,
                                     void* stream) {
  return forward_ws_api<float>(obs, y_offset, m0, S0, C, A, Q, shared_mask, B, T, n, p, a_mode, first, filt, nll, status, workspace,
                              workspace_bytes, stream);
}

extern "C" int eksb_rts_backward_ws_f32(const float* filt, const float* C, const float* A, const float* Q, int shared_mask, long B,
                                       long T, int n, int p, int a_mode, float* smth, float* ym, float* yv, float* cv, int* status,
                                       float* workspace, size_t workspace_bytes, void* stream) {
  return backward_ws_api<float>(filt, C, A, Q, shared_mask, B, T, n, p, a_mode, smth, ym, yv, cv, status, workspace, workspace_bytes,
                               stream);
}

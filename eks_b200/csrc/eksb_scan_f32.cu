// K4 parallel-in-time scan, float
#include "eksb_scan_impl.cuh"
using namespace eksb;

extern "C" int eksb_kf_rts_scan_f32(const float* obs, const float* m0, const float* S0, const float* C, const float* A, const float* Q,
                                   int shared_mask, long B, long T, int n, int p, int a_mode, long chunk, float* workspace,
                                   size_t workspace_bytes, float* filt, float* smth, float* ym, int* status, void* stream) {
  return scan_api<float>(obs, m0, S0, C, A, Q, shared_mask, B, T, n, p, a_mode, chunk, workspace, workspace_bytes, filt, smth, ym, status, stream);
}

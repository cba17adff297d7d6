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

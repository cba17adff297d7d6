// K4 parallel-in-time scan, double
#include <mutex>
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

// ------------------------------------------------------------------------------------------------------------
// eksb_kf_rts_scan_static_f64: eksb_kf_rts_scan_f64 for callers that must not allocate (code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The scan's workspace (chunk summaries, boundary states, cut positions) lives in
// this file-scope device array instead of a caller-allocated buffer, so every GPU that loads the library holds
// EKSB_SCAN_STATIC_SCRATCH_BYTES of device memory until the process exits.  In host code the array's name is only a
// host-side stand-in for the symbol, never a device address: the address is looked up once per device with
// cudaGetSymbolAddress (that device current) and cached.  All calls share the array and it keeps the previous
// call's contents: the only words that must start at zero (the two "fold must run" flags) are cleared on the
// stream by every call (scan_run_fwd_off / scan_run_bwd), everything else is written before it is read.  Two calls
// that overlap on the GPU (different streams, or different host threads without ordering) race on it.
// ------------------------------------------------------------------------------------------------------------
static_assert(EKSB_SCAN_STATIC_SCRATCH_BYTES % 256 == 0, "scratch size");
static __device__ __align__(256) unsigned char g_scan_static_scratch[EKSB_SCAN_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kMaxDevices = 64;
double* g_scan_static_ptr[kMaxDevices];      // device address of g_scan_static_scratch on device i (nullptr: not looked up yet)
std::mutex g_scan_static_mutex;

// set-up, once per device: the CURRENT device's copy of the array
int scan_static_scratch(double** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    set_error("kf_rts_scan_static: cudaGetDevice failed or device index >= %d (%s)", kMaxDevices, cudaGetErrorString(e));
    return EKSB_ECUDA;
  }
  std::lock_guard<std::mutex> lock(g_scan_static_mutex);
  if (!g_scan_static_ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_scan_static_scratch);
    if (e != cudaSuccess || !p) {
      set_error("kf_rts_scan_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return EKSB_ECUDA;
    }
    g_scan_static_ptr[dev] = static_cast<double*>(p);
  }
  *out = g_scan_static_ptr[dev];
  return EKSB_OK;
}
}  // namespace

extern "C" int eksb_kf_rts_scan_static_f64(const double* obs, const double* m0, const double* S0, const double* C, const double* A,
                                          const double* Q, int shared_mask, long B, long T, int n, int p, int a_mode, long chunk,
                                          double* filt, double* smth, double* ym, int* status, void* stream) {
  if (B <= 0 || T <= 0 || n <= 0) { set_error("kf_rts_scan_static: empty batch"); return EKSB_EARG; }
  const long L = chunk > 0 ? chunk : scan_auto_chunk(B, T, n, (int)sizeof(double));
  const size_t need = scan_workspace_elems(B, T, n, L) * sizeof(double);
  if (need > (size_t)EKSB_SCAN_STATIC_SCRATCH_BYTES) {
    set_error("kf_rts_scan_static: B=%ld, T=%ld, chunk=%ld needs %zu bytes of scan workspace, more than the static scratch of "
              "%zu bytes (EKSB_SCAN_STATIC_SCRATCH_BYTES); use a longer chunk, fewer tracks, or eksb_kf_rts_scan_f64 with "
              "a caller-allocated workspace", B, T, L, need, (size_t)EKSB_SCAN_STATIC_SCRATCH_BYTES);
    return EKSB_EARG;
  }
  double* ws = nullptr;
  const int rc = scan_static_scratch(&ws);
  if (rc != EKSB_OK) return rc;
  return scan_api<double>(obs, m0, S0, C, A, Q, shared_mask, B, T, n, p, a_mode, L, ws, (size_t)EKSB_SCAN_STATIC_SCRATCH_BYTES,
                          filt, smth, ym, status, stream);
}

// eksb_scan_impl.cuh -- host-side driver of the chunked parallel-in-time scan.
#pragma once
#include "eksb_scan.cuh"

namespace eksb {

inline size_t scan_workspace_elems(long B, long Tn, int n, long L) {
  const long nc = (Tn + L - 1) / L;
  const long per = (n * n + n + 2 * (n * (n + 1) / 2) + n)      /* forward summary  */
                   + (n * n + n * (n + 1) / 2 + n)             /* backward summary */
                   + 2 * (n + n * (n + 1) / 2);                /* boundary states, both passes */
  return (size_t)num_tiles(B) * nc * per * kLanes;
}

inline long scan_auto_chunk(long B, long Tn) {
  // aim at ~32 resident warps per SM: warps = tiles * chunks
  const long want_chunks = (148L * 32 + num_tiles(B) - 1) / num_tiles(B);
  long L = (Tn + want_chunks - 1) / want_chunks;
  if (L < 64) L = 64;
  return L;
}

template <typename T, int N, int P, int AM>
static int scan_run(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int mask, long B, long Tn,
                    long L, T* ws, T* filt, T* smth, T* ym, int* status, cudaStream_t st) {
  constexpr int SG = Dim<N>::SIGMA;
  const long tiles = num_tiles(B), nc = (Tn + L - 1) / L;
  T* fsum = ws;
  T* bsum = fsum + (size_t)tiles * nc * ScanRec<N>::FWD * kLanes;
  T* finit = bsum + (size_t)tiles * nc * ScanRec<N>::BWD * kLanes;
  T* binit = finit + (size_t)tiles * nc * SG * kLanes;
  auto blocks = [&](long items) { return (unsigned)((items + kScanWarps - 1) / kScanWarps); };
  const int thr = kScanWarps * kLanes;
  // forward pass
  if (nc == 1) {
    scan_fwd_rerun_kernel<T, N, P, AM><<<blocks(tiles), thr, 0, st>>>(obs, m0, S0, C, A, Q, mask, B, Tn, L, nc, 0, 1, finit, filt, status);
  } else {
    // chunk 0 (sequential from the prior) rides in the same launch as the summaries of chunks >= 1
    scan_fwd_summary_kernel<T, N, P, AM><<<blocks(tiles * nc), thr, 0, st>>>(obs, m0, S0, C, A, Q, mask, B, Tn, L, nc, fsum, filt, status);
    scan_fwd_fold_kernel<T, N><<<(unsigned)tiles, kLanes, 0, st>>>(filt, fsum, B, Tn, L, nc, finit, status);
    scan_fwd_rerun_kernel<T, N, P, AM><<<blocks(tiles * (nc - 1)), thr, 0, st>>>(obs, m0, S0, C, A, Q, mask, B, Tn, L, nc, 1, nc, finit, filt, status);
  }
  int rc = check_launch("scan forward kernels");
  if (rc != EKSB_OK || !smth) return rc;
  // backward pass (smth may alias filt: a chunk only reads and writes its own frames)
  if (nc == 1) {
    scan_bwd_rerun_kernel<T, N, P, AM><<<blocks(tiles), thr, 0, st>>>(filt, C, A, Q, mask, B, Tn, L, nc, nc - 1, nc, binit, smth, ym, status);
  } else {
    // the terminal chunk (sequential from the end of the track) rides in the summary launch
    scan_bwd_summary_kernel<T, N, P, AM><<<blocks(tiles * nc), thr, 0, st>>>(filt, C, A, Q, mask, B, Tn, L, nc, bsum, smth, ym, status);
    scan_bwd_fold_kernel<T, N><<<(unsigned)tiles, kLanes, 0, st>>>(smth, bsum, B, Tn, L, nc, binit);
    scan_bwd_rerun_kernel<T, N, P, AM><<<blocks(tiles * (nc - 1)), thr, 0, st>>>(filt, C, A, Q, mask, B, Tn, L, nc, 0, nc - 1, binit, smth, ym, status);
  }
  return check_launch("scan backward kernels");
}

template <typename T>
static int scan_api(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int mask, long B, long Tn,
                    int n, int p, int a_mode, long chunk, T* ws, size_t ws_bytes, T* filt, T* smth, T* ym, int* status,
                    void* stream) {
  if (!obs || !m0 || !S0 || !C || !A || !Q || !filt || !ws || B <= 0 || Tn <= 0) { set_error("kf_rts_scan: null pointer or empty batch"); return EKSB_EARG; }
  if (a_mode != EKSB_A_DIAG && a_mode != EKSB_A_DENSE) { set_error("kf_rts_scan: bad a_mode %d", a_mode); return EKSB_EARG; }
  const long L = chunk > 0 ? chunk : scan_auto_chunk(B, Tn);
  if (scan_workspace_elems(B, Tn, n, L) * sizeof(T) > ws_bytes) { set_error("kf_rts_scan: workspace too small (%zu bytes needed)", scan_workspace_elems(B, Tn, n, L) * sizeof(T)); return EKSB_EARG; }
#define X(N_, P_)                                                                                                   \
  if (n == N_ && p == P_)                                                                                           \
    return a_mode == EKSB_A_DIAG                                                                                    \
               ? scan_run<T, N_, P_, 0>(obs, m0, S0, C, A, Q, mask, B, Tn, L, ws, filt, smth, ym, status, (cudaStream_t)stream) \
               : scan_run<T, N_, P_, 1>(obs, m0, S0, C, A, Q, mask, B, Tn, L, ws, filt, smth, ym, status, (cudaStream_t)stream);
  X(2, 4) X(3, 4) X(3, 6) X(3, 8) X(4, 6)
#undef X
  set_error("kf_rts_scan: (n=%d,p=%d) not instantiated (state dim <= 4)", n, p);
  return EKSB_EDIM;
}

}  // namespace eksb

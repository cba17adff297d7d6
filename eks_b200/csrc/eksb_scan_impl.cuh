// eksb_scan_impl.cuh -- host-side driver of the chunked parallel-in-time scan.
#pragma once
#include "eksb_scan.cuh"

namespace eksb {

inline long scan_env_long(const char* name, long dflt) {
  const char* v = getenv(name);
  return (v && *v) ? atol(v) : dflt;
}

inline size_t scan_workspace_elems(long B, long Tn, int n, long L) {
  const long nc = (Tn + L - 1) / L;
  const long per = (n * n + n + 2 * (n * (n + 1) / 2) + n)      /* forward summary  */
                   + (n * n + n * (n + 1) / 2 + n)             /* backward summary */
                   + 2 * (n + n * (n + 1) / 2)                 /* boundary states, both passes */
                   + 1;                                        /* the two cut positions (2 x int64 per warp) */
  return (size_t)num_tiles(B) * nc * per * kLanes;
}

inline long scan_auto_chunk(long B, long Tn, int n, int elem_size) {
  // warps = tiles * chunks.  The summary kernels keep 12 (f64) / 16 (f32) warps resident per SM (register-bound,
  // n <= 3) and stage their frames through per-lane cp.async rings, so bytes in flight do not hinge on occupancy:
  // fewer, longer chunks win because every chunk pays K warm-up frames before its summary has forgotten the
  // boundary state (eksb_math.cuh scan_fwd_chunk).  Measured, 64 tracks x 10 M frames, 8 / 12 / 16 / 24 warps per
  // SM: f64 34.8 / 35.1 / 35.5 / 36.4 ms, f32 17.8 / - / 18.2 / 20.1 ms.
  (void)n;
  const long per_sm = scan_env_long("EKSB_SCAN_WARPS_PER_SM", 8);
  const long want_chunks = (148L * per_sm + num_tiles(B) - 1) / num_tiles(B);
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
  long* fcut = reinterpret_cast<long*>(binit + (size_t)tiles * nc * SG * kLanes);   // 32 T's (>= 16 B) per (tile, chunk)
  long* bcut = fcut + tiles * nc;
  int* alive = reinterpret_cast<int*>(bcut + tiles * nc);   // [0] forward, [1] backward: "the fold must run"
  if (nc > 1 && cudaMemsetAsync(alive, 0, 2 * sizeof(int), st) != cudaSuccess) return check_launch("scan flags");
  const int early = scan_env_long("EKSB_SCAN_NO_EARLY", 0) ? 0 : 1;                 // A/B switch: old 2-read behaviour
  auto blocks = [&](long items) { return (unsigned)((items + kScanWarps - 1) / kScanWarps); };
  const int thr = kScanWarps * kLanes;
  // per-lane frame rings of the summary kernels (eksb_scan.cuh), one per warp
  const size_t smem_f = kScanWarps * ring_bytes_per_warp<T, 2 * P>(), smem_b = kScanWarps * ring_bytes_per_warp<T, SG>();
  cudaFuncSetAttribute(scan_fwd_summary_kernel<T, N, P, AM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_f);
  cudaFuncSetAttribute(scan_bwd_summary_kernel<T, N, P, AM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_b);
  // forward pass
  if (nc == 1) {
    scan_fwd_rerun_kernel<T, N, P, AM><<<blocks(tiles), thr, 0, st>>>(obs, m0, S0, C, A, Q, mask, B, Tn, L, nc, 0, 1, finit, nullptr, filt, status);
  } else {
    // chunk 0 (sequential from the prior) rides in the same launch as the summaries of chunks >= 1
    scan_fwd_summary_kernel<T, N, P, AM><<<blocks(tiles * nc), thr, smem_f, st>>>(obs, m0, S0, C, A, Q, mask, B, Tn, L, nc, early, fsum, fcut, alive, filt, status);
    scan_fwd_fold_kernel<T, N><<<(unsigned)tiles, kLanes, 0, st>>>(filt, fsum, B, Tn, L, nc, alive, finit, status);
    scan_fwd_rerun_kernel<T, N, P, AM><<<blocks(tiles * (nc - 1)), thr, 0, st>>>(obs, m0, S0, C, A, Q, mask, B, Tn, L, nc, 1, nc, finit, fcut, filt, status);
  }
  int rc = check_launch("scan forward kernels");
  if (rc != EKSB_OK || !smth) return rc;
  // backward pass (smth may alias filt: a chunk only reads and writes its own frames)
  if (nc == 1) {
    scan_bwd_rerun_kernel<T, N, P, AM><<<blocks(tiles), thr, 0, st>>>(filt, C, A, Q, mask, B, Tn, L, nc, nc - 1, nc, binit, nullptr, smth, ym, status);
  } else {
    // the terminal chunk (sequential from the end of the track) rides in the summary launch
    scan_bwd_summary_kernel<T, N, P, AM><<<blocks(tiles * nc), thr, smem_b, st>>>(filt, C, A, Q, mask, B, Tn, L, nc, early, bsum, bcut, alive + 1, smth, ym, status);
    scan_bwd_fold_kernel<T, N><<<(unsigned)tiles, kLanes, 0, st>>>(smth, bsum, B, Tn, L, nc, alive + 1, binit);
    scan_bwd_rerun_kernel<T, N, P, AM><<<blocks(tiles * (nc - 1)), thr, 0, st>>>(filt, C, A, Q, mask, B, Tn, L, nc, 0, nc - 1, binit, bcut, smth, ym, status);
  }
  return check_launch("scan backward kernels");
}

template <typename T>
static int scan_api(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int mask, long B, long Tn,
                    int n, int p, int a_mode, long chunk, T* ws, size_t ws_bytes, T* filt, T* smth, T* ym, int* status,
                    void* stream) {
  if (!obs || !m0 || !S0 || !C || !A || !Q || !filt || !ws || B <= 0 || Tn <= 0) { set_error("kf_rts_scan: null pointer or empty batch"); return EKSB_EARG; }
  if (a_mode != EKSB_A_DIAG && a_mode != EKSB_A_DENSE) { set_error("kf_rts_scan: bad a_mode %d", a_mode); return EKSB_EARG; }
  const long L = chunk > 0 ? chunk : scan_auto_chunk(B, Tn, n, (int)sizeof(T));
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

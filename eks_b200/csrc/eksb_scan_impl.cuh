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

inline bool scan_v2_enabled() {
  const char* e = getenv("EKSB_SCAN_V2");                  // A/B switch: 0 = the thread-per-track summary / fold / re-run scan only
  return !(e && e[0] == '0');
}

// Chunk length.  Primary route (time-split pipelined scan): as many chunks per tile as give the pipelined kernels a full
// GPU of virtual tiles -- 148 SMs x the tiles one SM's shared memory holds (4 for p = 4, 3 for p = 8) -- in ONE wave, but
// chunks of at least 2048 frames so that the K warm-up frames per chunk (150-800) stay a small fraction.
// EKSB_SCAN_V2=0 (summary / fold / re-run kernels only): warps = tiles * chunks, 12 resident warps per SM when the
// chunks stay >= 3 500 frames, else 8 (measured, tools/route_sweep.py, f64, 100 000 frames, 8 vs 12 warps per SM:
// 32 tiles 6.68 / 7.61 ms, 64 tiles 13.03 / 12.83, 128 tiles 26.8 / 24.0, 256 tiles 52.2 / 45.8).
inline long scan_auto_chunk(long B, long Tn, int n, int elem_size, int p = 4) {
  const long tiles = num_tiles(B);
  auto len = [&](long want_chunks) {
    if (want_chunks < 1) want_chunks = 1;
    long L = (Tn + want_chunks - 1) / want_chunks;
    return L < 64 ? 64L : L;
  };
  const long forced = scan_env_long("EKSB_SCAN_WARPS_PER_SM", 0);
  if (forced > 0) return len(((long)sm_count() * forced + tiles - 1) / tiles);
  if (scan_v2_enabled()) {
    long target = (long)pipe_vt_target_tiles(n, p, elem_size);
    if (target <= 0) target = (long)sm_count() * 4;
    long nc = target / tiles;
    const long nc_max = Tn / 2048;
    if (nc > nc_max) nc = nc_max;
    return len(nc);
  }
  const long L12 = len(((long)sm_count() * 12 + tiles - 1) / tiles);
  return L12 >= 3500 ? L12 : len(((long)sm_count() * 8 + tiles - 1) / tiles);
}

// Workspace layout shared by the forward and the backward half (they may be launched by separate calls).
template <typename T, int N> struct ScanWs {
  T *fsum, *bsum, *finit, *binit;
  long *fcut, *bcut;
  int* alive;   // [0] forward, [1] backward: "the fold must run"; [2] forward, [3] backward: "some chunk never forgot its
                // start: the fallback kernels run instead of the virtual-tile kernels" (time-split scan)
  ScanWs(T* ws, long tiles, long nc) {
    constexpr int SG = Dim<N>::SIGMA;
    fsum = ws;
    bsum = fsum + (size_t)tiles * nc * ScanRec<N>::FWD * kLanes;
    finit = bsum + (size_t)tiles * nc * ScanRec<N>::BWD * kLanes;
    binit = finit + (size_t)tiles * nc * SG * kLanes;
    fcut = reinterpret_cast<long*>(binit + (size_t)tiles * nc * SG * kLanes);   // 32 T's (>= 16 B) per (tile, chunk)
    bcut = fcut + tiles * nc;
    alive = reinterpret_cast<int*>(bcut + tiles * nc);
  }
};

inline unsigned scan_blocks(long items) { return (unsigned)((items + kScanWarps - 1) / kScanWarps); }

// forward half: obs -> filt
template <typename T, int N, int P, int AM, bool OFF>
static int scan_run_fwd_off(const T* obs, const T* yoff, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int mask,
                            long B, long Tn, long L, int first, T* ws, T* filt, int* status, cudaStream_t st) {
  const long tiles = num_tiles(B), nc = (Tn + L - 1) / L;
  ScanWs<T, N> w(ws, tiles, nc);
  const int early = scan_env_long("EKSB_SCAN_NO_EARLY", 0) ? 0 : 1;                 // A/B switch: old 2-read behaviour
  const int thr = kScanWarps * kLanes;
  const size_t smem_f = kScanWarps * ring_bytes_per_warp<T, 2 * P>();               // per-lane frame rings, one per warp
  cudaFuncSetAttribute(scan_fwd_summary_kernel<T, N, P, AM, OFF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_f);
  if (nc == 1) {
    scan_fwd_rerun_kernel<T, N, P, AM, OFF><<<scan_blocks(tiles), thr, 0, st>>>(obs, m0, S0, C, A, Q, mask, B, Tn, L, nc, 0, 1, first, yoff, w.finit, nullptr, filt, status, nullptr);
    return check_launch("scan forward kernels");
  }
  if (cudaMemsetAsync(w.alive, 0, sizeof(int), st) != cudaSuccess || cudaMemsetAsync(w.alive + 2, 0, sizeof(int), st) != cudaSuccess)
    return check_launch("scan flags");
  const int* gate = nullptr;
  if (scan_v2_enabled()) {
    // time-split pipelined scan: K warm-up frames per chunk, then every (tile, chunk) is a tile of the TMA-fed kernel
    gate = w.alive + 2;
    // three-role warm-up: double only (its scaled chain forms det(adj S + det S Lam) ~ (var)^6: below the float range
    // for process noise ~1e-7 -- the main kernels' chain has the same limit, but a warm-up has no fallback of its own)
    if constexpr (AM == 0 && N <= 3 && sizeof(T) == 8) {
      // ... and only while its three warps per chunk find schedulers of their own: measured (ncu, f64) 256 chunks 0.216 ->
      // 0.140 ms, 368 (NLL sweep) 5.6 -> 4.6 ms per sweep, 512 chunks 0.22 -> 0.26 ms
      if (scan_env_long("EKSB_SCAN_WARM1", 0) || tiles * (nc - 1) > 3L * sm_count())   // EKSB_SCAN_WARM1=1: A/B switch
        scan_fwd_warm_kernel<T, N, P, AM, OFF><<<scan_blocks(tiles * (nc - 1)), thr, 0, st>>>(obs, C, A, Q, mask, B, Tn, L, nc, yoff, w.fcut, w.alive + 2, filt, (const T*)nullptr, (T*)nullptr, 0L);
      else {
        cudaFuncSetAttribute(scan_fwd_warm3_kernel<T, N, P, OFF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Warm3Smem<T, N>::kTotal);
        scan_fwd_warm3_kernel<T, N, P, OFF><<<(unsigned)(tiles * (nc - 1)), 3 * kLanes, Warm3Smem<T, N>::kTotal, st>>>(obs, C, A, Q, mask, B, Tn, L, nc, yoff, w.fcut, w.alive + 2, filt, (const T*)nullptr, (T*)nullptr, 0L);
      }
    } else {
      scan_fwd_warm_kernel<T, N, P, AM, OFF><<<scan_blocks(tiles * (nc - 1)), thr, 0, st>>>(obs, C, A, Q, mask, B, Tn, L, nc, yoff, w.fcut, w.alive + 2, filt, (const T*)nullptr, (T*)nullptr, 0L);
    }
    int rc = check_launch("scan forward warm-up");
    if (rc != EKSB_OK) return rc;
    VTiles vt;
    vt.L = L; vt.nc = nc; vt.cut = w.fcut; vt.gate = gate;
    rc = kf_forward_pipe_vt<T>(obs, m0, S0, C, A, Q, mask, B, Tn, N, P, AM ? EKSB_A_DENSE : EKSB_A_DIAG, first, filt, status, yoff, vt, st);
    if (rc != EKSB_OK) return rc;
  }
  // summary / fold / re-run: the whole pass when gate == nullptr, else only if some chunk never forgot its start
  // (chunk 0, sequential from the prior, rides in the same launch as the summaries of chunks >= 1)
  scan_fwd_summary_kernel<T, N, P, AM, OFF><<<scan_blocks(tiles * nc), thr, smem_f, st>>>(obs, m0, S0, C, A, Q, mask, B, Tn, L, nc, early, first, yoff, w.fsum, w.fcut, w.alive, filt, status, gate);
  scan_fwd_fold_kernel<T, N><<<(unsigned)tiles, kLanes, 0, st>>>(filt, w.fsum, B, Tn, L, nc, w.alive, w.finit, status, gate);
  scan_fwd_rerun_kernel<T, N, P, AM, OFF><<<scan_blocks(tiles * (nc - 1)), thr, 0, st>>>(obs, m0, S0, C, A, Q, mask, B, Tn, L, nc, 1, nc, first, yoff, w.finit, w.fcut, filt, status, gate);
  return check_launch("scan forward kernels");
}
template <typename T, int N, int P, int AM>
static int scan_run_fwd(const T* obs, const T* yoff, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int mask,
                        long B, long Tn, long L, int first, T* ws, T* filt, int* status, cudaStream_t st) {
  return yoff ? scan_run_fwd_off<T, N, P, AM, true>(obs, yoff, m0, S0, C, A, Q, mask, B, Tn, L, first, ws, filt, status, st)
              : scan_run_fwd_off<T, N, P, AM, false>(obs, yoff, m0, S0, C, A, Q, mask, B, Tn, L, first, ws, filt, status, st);
}

// backward half: filt -> smth (+ ym); smth may alias filt (a chunk only reads and writes its own frames)
template <typename T, int N, int P, int AM>
static int scan_run_bwd(const T* filt, const T* C, const T* A, const T* Q, int mask, long B, long Tn, long L, T* ws, T* smth,
                        T* ym, int* status, cudaStream_t st) {
  constexpr int SG = Dim<N>::SIGMA;
  const long tiles = num_tiles(B), nc = (Tn + L - 1) / L;
  ScanWs<T, N> w(ws, tiles, nc);
  const int early = scan_env_long("EKSB_SCAN_NO_EARLY", 0) ? 0 : 1;
  const int thr = kScanWarps * kLanes;
  const size_t smem_b = kScanWarps * ring_bytes_per_warp<T, SG>();
  cudaFuncSetAttribute(scan_bwd_summary_kernel<T, N, P, AM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_b);
  if (nc == 1) {
    scan_bwd_rerun_kernel<T, N, P, AM><<<scan_blocks(tiles), thr, 0, st>>>(filt, C, A, Q, mask, B, Tn, L, nc, nc - 1, nc, w.binit, nullptr, smth, ym, status, nullptr);
    return check_launch("scan backward kernels");
  }
  if (cudaMemsetAsync(w.alive + 1, 0, sizeof(int), st) != cudaSuccess || cudaMemsetAsync(w.alive + 3, 0, sizeof(int), st) != cudaSuccess)
    return check_launch("scan flags");
  const int* gate = nullptr;
  if (scan_v2_enabled()) {
    gate = w.alive + 3;
    if (scan_env_long("EKSB_SCAN_WARM1", 0)) {               // A/B switch: the one-warp warm-up kernel
      scan_bwd_warm_kernel<T, N, P, AM><<<scan_blocks(tiles * (nc - 1)), thr, 0, st>>>(filt, C, A, Q, mask, B, Tn, L, nc, w.bcut, w.alive + 3, w.binit);
    } else {
      cudaFuncSetAttribute(scan_bwd_warm2_kernel<T, N, P, AM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Warm2Smem<T, N>::kTotal);
      scan_bwd_warm2_kernel<T, N, P, AM><<<(unsigned)(tiles * (nc - 1)), 2 * kLanes, Warm2Smem<T, N>::kTotal, st>>>(filt, C, A, Q, mask, B, Tn, L, nc, w.bcut, w.alive + 3, w.binit);
    }
    int rc = check_launch("scan backward warm-up");
    if (rc != EKSB_OK) return rc;
    VTiles vt;
    vt.L = L; vt.nc = nc; vt.cut = w.bcut; vt.init = w.binit; vt.gate = gate;
    rc = rts_backward_pipe_vt<T>(filt, C, A, Q, mask, B, Tn, N, P, AM ? EKSB_A_DENSE : EKSB_A_DIAG, smth, ym, status, vt, st);
    if (rc != EKSB_OK) return rc;
  }
  // (the terminal chunk, sequential from the end of the track, rides in the summary launch)
  scan_bwd_summary_kernel<T, N, P, AM><<<scan_blocks(tiles * nc), thr, smem_b, st>>>(filt, C, A, Q, mask, B, Tn, L, nc, early, w.bsum, w.bcut, w.alive + 1, smth, ym, status, gate);
  scan_bwd_fold_kernel<T, N><<<(unsigned)tiles, kLanes, 0, st>>>(smth, w.bsum, B, Tn, L, nc, w.alive + 1, w.binit, gate);
  scan_bwd_rerun_kernel<T, N, P, AM><<<scan_blocks(tiles * (nc - 1)), thr, 0, st>>>(filt, C, A, Q, mask, B, Tn, L, nc, 0, nc - 1, w.binit, w.bcut, smth, ym, status, gate);
  return check_launch("scan backward kernels");
}

#define EKSB_FOR_EACH_SCAN_DIM(X) X(2, 4) X(3, 4) X(3, 6) X(3, 8) X(4, 6)
inline bool scan_dims_ok(int n, int p) {
#define X(N_, P_) if (n == N_ && p == P_) return true;
  EKSB_FOR_EACH_SCAN_DIM(X)
#undef X
  return false;
}

// which == 1: forward half, 2: backward half, 3: both
template <typename T>
static int scan_dispatch(int which, const T* obs, const T* yoff, const T* m0, const T* S0, const T* C, const T* A, const T* Q,
                         int mask, long B, long Tn, int n, int p, int a_mode, long L, int first, T* ws, T* filt, T* smth,
                         T* ym, int* status, cudaStream_t st) {
#define X(N_, P_)                                                                                                        \
  if (n == N_ && p == P_) {                                                                                              \
    int rc = EKSB_OK;                                                                                                    \
    if (which & 1)                                                                                                       \
      rc = a_mode == EKSB_A_DIAG                                                                                         \
               ? scan_run_fwd<T, N_, P_, 0>(obs, yoff, m0, S0, C, A, Q, mask, B, Tn, L, first, ws, filt, status, st)      \
               : scan_run_fwd<T, N_, P_, 1>(obs, yoff, m0, S0, C, A, Q, mask, B, Tn, L, first, ws, filt, status, st);     \
    if (rc == EKSB_OK && (which & 2))                                                                                    \
      rc = a_mode == EKSB_A_DIAG ? scan_run_bwd<T, N_, P_, 0>(filt, C, A, Q, mask, B, Tn, L, ws, smth, ym, status, st)    \
                                 : scan_run_bwd<T, N_, P_, 1>(filt, C, A, Q, mask, B, Tn, L, ws, smth, ym, status, st);   \
    return rc;                                                                                                           \
  }
  EKSB_FOR_EACH_SCAN_DIM(X)
#undef X
  set_error("kf_rts_scan: (n=%d,p=%d) not instantiated (state dim <= 4)", n, p);
  return EKSB_EDIM;
}

template <typename T>
static int scan_api(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int mask, long B, long Tn,
                    int n, int p, int a_mode, long chunk, T* ws, size_t ws_bytes, T* filt, T* smth, T* ym, int* status,
                    void* stream) {
  if (!obs || !m0 || !S0 || !C || !A || !Q || !filt || !ws || B <= 0 || Tn <= 0) { set_error("kf_rts_scan: null pointer or empty batch"); return EKSB_EARG; }
  if (a_mode != EKSB_A_DIAG && a_mode != EKSB_A_DENSE) { set_error("kf_rts_scan: bad a_mode %d", a_mode); return EKSB_EARG; }
  const long L = chunk > 0 ? chunk : scan_auto_chunk(B, Tn, n, (int)sizeof(T));
  if (scan_workspace_elems(B, Tn, n, L) * sizeof(T) > ws_bytes) { set_error("kf_rts_scan: workspace too small (%zu bytes needed)", scan_workspace_elems(B, Tn, n, L) * sizeof(T)); return EKSB_EARG; }
  return scan_dispatch<T>(smth ? 3 : 1, obs, (const T*)nullptr, m0, S0, C, A, Q, mask, B, Tn, n, p, a_mode, L, kPriorUpdate, ws,
                          filt, smth, ym, status, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------------------
// Routed entry points (eksb_kf_forward_ws / eksb_rts_backward_ws): the same contracts as eksb_kf_forward_offset /
// eksb_rts_backward plus a caller-owned workspace.  The sequential warp-specialised kernels keep one chain warp per
// track tile, so a batch with fewer tiles than the GPU has SMs (or an awkward count between one and two tiles per
// SM) runs at one warp's pace no matter how few bytes it moves; such batches go through the two halves of the
// parallel-in-time scan instead (policy and the measurements behind it: route_scan below).
// EKSB_ROUTE=seq|scan forces a side (read per call).
// ------------------------------------------------------------------------------------------------------------
inline bool route_scan(long B, long Tn, int n, int p, int elem_size) {
  if (B <= 0 || Tn <= 0 || !scan_dims_ok(n, p)) return false;
  const char* e = getenv("EKSB_ROUTE");
  if (e && e[0] == 's' && e[1] == 'e') return false;
  if (e && e[0] == 's' && e[1] == 'c') return true;
  if (Tn < 4096) return false;
  const long tiles = num_tiles(B);
  if (scan_v2_enabled()) {
    // Worth it while the real tiles leave the pipelined kernels short of ~2 tiles per SM.  Measured (tools/route_sweep.py,
    // f64, 100 000 frames, sequential vs time-split): 16 tiles 25.3 / 4.2 ms, 64 tiles 25.2 / 12.2, 128 tiles 26.7 / 22.7,
    // 192 tiles 39.2 / 33.1, 256 tiles 44.0 / 44.2, 296 tiles 48.7 / 49.9 (f32: 64 tiles 17.5 / 6.1, 128 tiles 17.5 / 11.1,
    // 256 tiles 24.8 / 24.1); in the bench's sustained, power-capped region 2 x 256 tiles take 89.7 / 88.0 ms.
    // Hence: at least two chunks per tile and at most 0.45 x the tiles that fill the GPU (256 of 592).
    long target = (long)pipe_vt_target_tiles(n, p, elem_size);
    if (target <= 0) target = (long)sm_count() * 4;
    if (tiles * 100 > target * 45) return false;
    const long L = scan_auto_chunk(B, Tn, n, elem_size, p);
    return (Tn + L - 1) / L >= 2;
  }
  if (tiles <= 96) return true;
  return tiles <= 208 && scan_auto_chunk(B, Tn, n, elem_size, p) >= 2048;
}

// Sequential entry points of the other translation units (same shared library).
template <typename T> struct SeqApi;
template <> struct SeqApi<double> {
  static int fwd(const double* obs, const double* yoff, const double* m0, const double* S0, const double* C, const double* A,
                 const double* Q, int mask, long B, long Tn, int n, int p, int am, int first, double* filt, double* nll,
                 int* status, void* st) {
    return eksb_kf_forward_offset_f64(obs, yoff, m0, S0, C, A, Q, mask, B, Tn, n, p, am, first, filt, nll, status, st);
  }
  static int bwd(const double* filt, const double* C, const double* A, const double* Q, int mask, long B, long Tn, int n, int p,
                 int am, double* smth, double* ym, double* yv, double* cv, int* status, void* st) {
    return eksb_rts_backward_f64(filt, C, A, Q, mask, B, Tn, n, p, am, smth, ym, yv, cv, status, st);
  }
};
template <> struct SeqApi<float> {
  static int fwd(const float* obs, const float* yoff, const float* m0, const float* S0, const float* C, const float* A,
                 const float* Q, int mask, long B, long Tn, int n, int p, int am, int first, float* filt, double* nll,
                 int* status, void* st) {
    return eksb_kf_forward_offset_f32(obs, yoff, m0, S0, C, A, Q, mask, B, Tn, n, p, am, first, filt, nll, status, st);
  }
  static int bwd(const float* filt, const float* C, const float* A, const float* Q, int mask, long B, long Tn, int n, int p,
                 int am, float* smth, float* ym, float* yv, float* cv, int* status, void* st) {
    return eksb_rts_backward_f32(filt, C, A, Q, mask, B, Tn, n, p, am, smth, ym, yv, cv, status, st);
  }
};

inline size_t route_workspace_bytes(long B, long Tn, int n, int p, int elem_size) {
  if (!route_scan(B, Tn, n, p, elem_size)) return 0;
  return scan_workspace_elems(B, Tn, n, scan_auto_chunk(B, Tn, n, elem_size, p)) * (size_t)elem_size;
}

template <typename T>
static int forward_ws_api(const T* obs, const T* yoff, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int mask,
                          long B, long Tn, int n, int p, int a_mode, int first, T* filt, double* nll, int* status, T* ws,
                          size_t ws_bytes, void* stream) {
  const size_t need = (ws && filt && !nll) ? route_workspace_bytes(B, Tn, n, p, (int)sizeof(T)) : 0;
  if (need == 0 || ws_bytes < need || !obs || !m0 || !S0 || !C || !A || !Q ||
      (first != kPriorUpdate && first != kPriorOnly) || (a_mode != EKSB_A_DIAG && a_mode != EKSB_A_DENSE))
    return SeqApi<T>::fwd(obs, yoff, m0, S0, C, A, Q, mask, B, Tn, n, p, a_mode, first, filt, nll, status, stream);
  return scan_dispatch<T>(1, obs, yoff, m0, S0, C, A, Q, mask, B, Tn, n, p, a_mode, scan_auto_chunk(B, Tn, n, (int)sizeof(T), p),
                          first, ws, filt, (T*)nullptr, (T*)nullptr, status, (cudaStream_t)stream);
}

template <typename T>
static int backward_ws_api(const T* filt, const T* C, const T* A, const T* Q, int mask, long B, long Tn, int n, int p, int a_mode,
                           T* smth, T* ym, T* yv, T* cv, int* status, T* ws, size_t ws_bytes, void* stream) {
  const size_t need = (ws && !yv && !cv) ? route_workspace_bytes(B, Tn, n, p, (int)sizeof(T)) : 0;
  if (need == 0 || ws_bytes < need || !filt || !smth || !C || !A || !Q || (a_mode != EKSB_A_DIAG && a_mode != EKSB_A_DENSE))
    return SeqApi<T>::bwd(filt, C, A, Q, mask, B, Tn, n, p, a_mode, smth, ym, yv, cv, status, stream);
  return scan_dispatch<T>(2, (const T*)nullptr, (const T*)nullptr, (const T*)nullptr, (const T*)nullptr, C, A, Q, mask, B, Tn, n, p,
                          a_mode, scan_auto_chunk(B, Tn, n, (int)sizeof(T), p), kPriorUpdate, ws, const_cast<T*>(filt), smth, ym,
                          status, (cudaStream_t)stream);
}

}  // namespace eksb

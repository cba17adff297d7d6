// eksb_scan.cuh -- K4: exact parallel-in-time filter + RTS smoother for long tracks
// (config 5: 64 tracks x 10 M frames, where tracks alone give only 2 warps of parallelism).
//
// Time is cut into chunks of `L` frames; one warp handles one (track tile, chunk).  Per pass:
//   1. summary kernel : every chunk's scan element, accumulated sequentially inside the chunk
//                       (eksb_math.cuh fwd_summary_step / bwd_summary_step) -- reads the inputs once;
//   2. fold kernel    : one thread per track walks the chunk summaries and emits the exact state at
//                       every chunk boundary (fwd_apply_summary / bwd_chain) -- O(T/L) steps;
//   3. re-run kernel  : every chunk runs the ordinary sequential recursion from its exact boundary
//                       state and writes the per-frame records (kf_forward_range / rts_backward_range).
// The result is the sequential result up to rounding (tests: chunked == sequential == oracle).
// Extra HBM traffic vs the sequential kernels: one more read of the pass's input.
#pragma once
#include "eksb_common.cuh"

namespace eksb {

constexpr int kScanWarps = 4;   // warps (= chunks) per CTA

// ---- 1a. forward chunk summaries, chunks [1, nc), plus the sequential run of chunk 0 -----------------
// Work items [0, tiles): chunk 0 of each tile, run as the ordinary filter from the prior (it needs no
// summary, only its end state, and would otherwise be a serial 1-warp-per-tile launch of its own).
// Work items [tiles, tiles*nc): summary of chunk c >= 1.  Once the sensitivity Phi of a chunk's running
// mean to its unknown start state has decayed below kPhiCut for all 32 tracks of the warp (the filter
// forgets its start exponentially), J and eta have stopped changing and Phi is numerically zero, so the
// rest of the chunk is the plain filter step (2.4x cheaper); the element is exact to ~kPhiCut * |x_start|.
template <typename T> struct PhiCut;
template <> struct PhiCut<double> { static constexpr double value = 1e-26; };
template <> struct PhiCut<float> { static constexpr float value = 1e-13f; };

template <typename T, int N, int P, int AM>
__global__ void __launch_bounds__(kScanWarps* kLanes)
scan_fwd_summary_kernel(const T* __restrict__ obs, const T* __restrict__ m0, const T* __restrict__ S0,
                        const T* __restrict__ C, const T* __restrict__ A, const T* __restrict__ Q, int shared_mask,
                        long B, long Tn, long L, long nc, T* __restrict__ fsum, T* __restrict__ filt,
                        int* __restrict__ status) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, R2 = 2 * P, SR = ScanRec<N>::FWD;
  const int lane = threadIdx.x & 31;
  const long tiles = num_tiles_dev(B);
  const long w = (long)blockIdx.x * kScanWarps + (threadIdx.x >> 5);
  if (w >= tiles * nc) return;
  Params<T, N, P, AM> prm;
  if (w < tiles) {                                         // chunk 0: ordinary filter from (m0, S0)
    const long g = w;
    const long b = g * kLanes + lane, bp = b < B ? b : B - 1;
    load_track_params<T, N, P, AM>(prm, C, A, Q, shared_mask, bp);
    T m[N], V[NS];
    const T* pm = m0 + ((shared_mask & EKSB_SHARED_M0) ? 0 : bp * N);
    const T* ps = S0 + ((shared_mask & EKSB_SHARED_S0) ? 0 : bp * (N * N));
#pragma unroll
    for (int i = 0; i < N; ++i) m[i] = pm[i];
    load_sym<T, N>(V, ps);
    const long t1 = L < Tn ? L : Tn;
    double acc = 0.0;
    const bool ok0 = kf_forward_range<T, N, P, AM, kLanes, 2, false>(prm, obs + (g * Tn * R2) * kLanes + lane,
                                                                    filt + (g * Tn * SG) * kLanes + lane, t1,
                                                                    kPriorUpdate, m, V, acc);
    if (b < B && status && !ok0) status[b] = 1;
    return;
  }
  const long wi = w - tiles;                               // work item = g*(nc-1) + (c-1)
  const long g = wi / (nc - 1);
  const long c = wi - g * (nc - 1) + 1;
  const long b = g * kLanes + lane, bp = b < B ? b : B - 1;
  load_track_params<T, N, P, AM>(prm, C, A, Q, shared_mask, bp);
  T Phi[N][N], bb[N], Cc[NS], eta[N], J[NS];
#pragma unroll
  for (int i = 0; i < N; ++i) {
    bb[i] = eta[i] = T(0);
#pragma unroll
    for (int j = 0; j < N; ++j) Phi[i][j] = (i == j) ? T(1) : T(0);
  }
#pragma unroll
  for (int i = 0; i < NS; ++i) Cc[i] = J[i] = T(0);
  bool ok = true;
  const long t1 = (c + 1) * L < Tn ? (c + 1) * L : Tn;
  const T* o = obs + (g * Tn * R2) * kLanes + lane;
  long t = c * L;
  for (; t < t1; ++t) {
    T cur[R2], rec[Rec<N>::FWD];
#pragma unroll
    for (int r = 0; r < R2; ++r) cur[r] = o[(t * R2 + r) * kLanes];
    fwd_prep<T, N, P, AM>(prm, cur, cur + P, rec);
    fwd_summary_step<T, N, P, AM>(prm, rec, Phi, bb, Cc, eta, J, ok);
    if (((t - c * L) & 15) == 15) {                        // every 16 frames: has Phi died out for the whole warp?
      T mx = T(0);
#pragma unroll
      for (int i = 0; i < N; ++i)
#pragma unroll
        for (int j = 0; j < N; ++j) mx = fmax(mx, fabs(Phi[i][j]));
      if (__all_sync(0xffffffffu, mx < PhiCut<T>::value)) { ++t; break; }
    }
  }
  if (t < t1) {                                            // tail: plain filter on (bb, Cc), Phi == 0
#pragma unroll
    for (int i = 0; i < N; ++i)
#pragma unroll
      for (int j = 0; j < N; ++j) Phi[i][j] = T(0);
    for (; t < t1; ++t) {
      T cur[R2], rec[Rec<N>::FWD];
#pragma unroll
      for (int r = 0; r < R2; ++r) cur[r] = o[(t * R2 + r) * kLanes];
      fwd_prep<T, N, P, AM>(prm, cur, cur + P, rec);
      fwd_chain<T, N, P, AM>(prm, rec, true, bb, Cc, ok);
    }
  }
  T* out = fsum + ((g * nc + c) * SR) * kLanes + lane;
#pragma unroll
  for (int i = 0; i < N; ++i)
#pragma unroll
    for (int j = 0; j < N; ++j) out[(i * N + j) * kLanes] = Phi[i][j];
#pragma unroll
  for (int i = 0; i < N; ++i) { out[(N * N + i) * kLanes] = bb[i]; out[(N * N + N + NS + i) * kLanes] = eta[i]; }
#pragma unroll
  for (int i = 0; i < NS; ++i) { out[(N * N + N + i) * kLanes] = Cc[i]; out[(N * N + N + NS + N + i) * kLanes] = J[i]; }
  if (b < B && status && !ok) status[b] = 1;
}

// ---- 2a. forward fold: exact filtered state at the frame before every chunk ---------------------
template <typename T, int N>
__global__ void __launch_bounds__(kLanes)
scan_fwd_fold_kernel(const T* __restrict__ filt, const T* __restrict__ fsum, long B, long Tn, long L, long nc,
                     T* __restrict__ finit, int* __restrict__ status) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, SR = ScanRec<N>::FWD;
  const int lane = threadIdx.x;
  const long g = blockIdx.x;
  const long b = g * kLanes + lane;
  T m[N], V[NS];
  const T* f = filt + ((g * Tn + (L - 1)) * SG) * kLanes + lane;          // end of chunk 0
#pragma unroll
  for (int i = 0; i < N; ++i) m[i] = f[i * kLanes];
#pragma unroll
  for (int i = 0; i < NS; ++i) V[i] = f[(N + i) * kLanes];
  bool ok = true;
  T nxt[SR];
  {
    const T* s = fsum + ((g * nc + 1) * SR) * kLanes + lane;
#pragma unroll
    for (int r = 0; r < SR; ++r) nxt[r] = s[r * kLanes];
  }
  for (long c = 1; c < nc; ++c) {
    T* o = finit + ((g * nc + c) * SG) * kLanes + lane;
#pragma unroll
    for (int i = 0; i < N; ++i) o[i * kLanes] = m[i];
#pragma unroll
    for (int i = 0; i < NS; ++i) o[(N + i) * kLanes] = V[i];
    T sum[SR];
#pragma unroll
    for (int r = 0; r < SR; ++r) sum[r] = nxt[r];
    if (c + 1 < nc) {                                                     // prefetch the next summary
      const T* s = fsum + ((g * nc + c + 1) * SR) * kLanes + lane;
#pragma unroll
      for (int r = 0; r < SR; ++r) nxt[r] = s[r * kLanes];
    }
    fwd_apply_summary<T, N>(sum, m, V, ok);
  }
  if (b < B && status && !ok) status[b] = 1;
}

// ---- 3a. forward re-run of chunks [c_lo, c_hi) ---------------------------------------------------
template <typename T, int N, int P, int AM>
__global__ void __launch_bounds__(kScanWarps* kLanes)
scan_fwd_rerun_kernel(const T* __restrict__ obs, const T* __restrict__ m0, const T* __restrict__ S0,
                      const T* __restrict__ C, const T* __restrict__ A, const T* __restrict__ Q, int shared_mask, long B,
                      long Tn, long L, long nc, long c_lo, long c_hi, const T* __restrict__ finit, T* __restrict__ filt,
                      int* __restrict__ status) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, R2 = 2 * P;
  const int lane = threadIdx.x & 31;
  const long span = c_hi - c_lo;
  const long w = (long)blockIdx.x * kScanWarps + (threadIdx.x >> 5);
  const long g = w / span;
  const long c = w - g * span + c_lo;
  if (g >= num_tiles_dev(B)) return;
  const long b = g * kLanes + lane, bp = b < B ? b : B - 1;
  Params<T, N, P, AM> prm;
  load_track_params<T, N, P, AM>(prm, C, A, Q, shared_mask, bp);
  T m[N], V[NS];
  int first = kContinue;
  if (c == 0) {
    const T* pm = m0 + ((shared_mask & EKSB_SHARED_M0) ? 0 : bp * N);
    const T* ps = S0 + ((shared_mask & EKSB_SHARED_S0) ? 0 : bp * (N * N));
#pragma unroll
    for (int i = 0; i < N; ++i) m[i] = pm[i];
    load_sym<T, N>(V, ps);
    first = kPriorUpdate;
  } else {
    const T* s = finit + ((g * nc + c) * SG) * kLanes + lane;
#pragma unroll
    for (int i = 0; i < N; ++i) m[i] = s[i * kLanes];
#pragma unroll
    for (int i = 0; i < NS; ++i) V[i] = s[(N + i) * kLanes];
  }
  const long t0 = c * L, t1 = (c + 1) * L < Tn ? (c + 1) * L : Tn;
  double acc = 0.0;
  bool ok = kf_forward_range<T, N, P, AM, kLanes, 2, false>(prm, obs + ((g * Tn + t0) * R2) * kLanes + lane,
                                                            filt + ((g * Tn + t0) * SG) * kLanes + lane, t1 - t0, first,
                                                            m, V, acc);
  if (b < B && status && !ok) status[b] = 1;
}

// ---- 1b. backward chunk summaries, chunks [0, nc-1), plus the sequential run of the last chunk --------
// Work items [0, tiles): the terminal chunk of each tile, smoothed directly (ms = mf at T-1); it only
// touches its own frames, so it can ride in the launch that summarises the other chunks (smth may alias
// filt).  Work items [tiles, tiles*nc): (E, g, L) summary of chunk c < nc-1.
template <typename T, int N, int P, int AM>
__global__ void __launch_bounds__(kScanWarps* kLanes)
scan_bwd_summary_kernel(const T* filt, const T* __restrict__ C, const T* __restrict__ A,
                        const T* __restrict__ Q, int shared_mask, long B, long Tn, long L, long nc,
                        T* __restrict__ bsum, T* smth, T* __restrict__ ym, int* __restrict__ status) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, SR = ScanRec<N>::BWD;
  const int lane = threadIdx.x & 31;
  const long tiles = num_tiles_dev(B);
  const long w = (long)blockIdx.x * kScanWarps + (threadIdx.x >> 5);
  if (w >= tiles * nc) return;
  Params<T, N, P, AM> prm;
  if (w < tiles) {
    const long g = w;
    const long b = g * kLanes + lane, bp = b < B ? b : B - 1;
    load_track_params<T, N, P, AM>(prm, C, A, Q, shared_mask, bp);
    T ms[N], Vs[NS];
    const long t0 = (nc - 1) * L;
    const long base = (g * Tn + t0) * kLanes;
    const bool ok0 = rts_backward_range<T, N, P, AM, kLanes, 2>(prm, filt + base * SG + lane, smth + base * SG + lane,
                                                               ym ? ym + base * P + lane : nullptr, nullptr, nullptr,
                                                               Tn - t0, true, ms, Vs);
    if (b < B && status && !ok0) status[b] = 1;
    return;
  }
  const long wi = w - tiles;
  const long g = wi / (nc - 1);
  const long c = wi - g * (nc - 1);
  const long b = g * kLanes + lane, bp = b < B ? b : B - 1;
  load_track_params<T, N, P, AM>(prm, C, A, Q, shared_mask, bp);
  T acc[SR];
#pragma unroll
  for (int i = 0; i < SR; ++i) acc[i] = T(0);
#pragma unroll
  for (int i = 0; i < N; ++i) acc[i * N + i] = T(1);
  bool ok = true;
  const T* f = filt + (g * Tn * SG) * kLanes + lane;
  for (long t = (c + 1) * L - 1; t >= c * L; --t) {
    T cur[SG], rec[Rec<N>::BWD];
#pragma unroll
    for (int r = 0; r < SG; ++r) cur[r] = f[(t * SG + r) * kLanes];
    bwd_prep<T, N, P, AM>(prm, cur, cur + N, rec, ok);
    bwd_summary_step<T, N>(rec, acc);
  }
  T* out = bsum + ((g * nc + c) * SR) * kLanes + lane;
#pragma unroll
  for (int r = 0; r < SR; ++r) out[r * kLanes] = acc[r];
  if (b < B && status && !ok) status[b] = 1;
}

// ---- 2b. backward fold: exact smoothed state of the frame after every chunk ----------------------
template <typename T, int N>
__global__ void __launch_bounds__(kLanes)
scan_bwd_fold_kernel(const T* __restrict__ smth, const T* __restrict__ bsum, long B, long Tn, long L, long nc,
                     T* __restrict__ binit) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, SR = ScanRec<N>::BWD;
  const int lane = threadIdx.x;
  const long g = blockIdx.x;
  T ms[N], Vs[NS];
  const T* s = smth + ((g * Tn + (nc - 1) * L) * SG) * kLanes + lane;     // first frame of the last chunk
#pragma unroll
  for (int i = 0; i < N; ++i) ms[i] = s[i * kLanes];
#pragma unroll
  for (int i = 0; i < NS; ++i) Vs[i] = s[(N + i) * kLanes];
  for (long c = nc - 2; c >= 0; --c) {
    T* o = binit + ((g * nc + c) * SG) * kLanes + lane;
#pragma unroll
    for (int i = 0; i < N; ++i) o[i * kLanes] = ms[i];
#pragma unroll
    for (int i = 0; i < NS; ++i) o[(N + i) * kLanes] = Vs[i];
    T sum[SR];
    const T* q = bsum + ((g * nc + c) * SR) * kLanes + lane;
#pragma unroll
    for (int r = 0; r < SR; ++r) sum[r] = q[r * kLanes];
    bwd_chain<T, N>(sum, ms, Vs);
  }
}

// ---- 3b. backward re-run of chunks [c_lo, c_hi) --------------------------------------------------
template <typename T, int N, int P, int AM>
__global__ void __launch_bounds__(kScanWarps* kLanes)
scan_bwd_rerun_kernel(const T* filt, const T* __restrict__ C, const T* __restrict__ A, const T* __restrict__ Q,
                      int shared_mask, long B, long Tn, long L, long nc, long c_lo, long c_hi,
                      const T* __restrict__ binit, T* smth, T* __restrict__ ym, int* __restrict__ status) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA;
  const int lane = threadIdx.x & 31;
  const long span = c_hi - c_lo;
  const long w = (long)blockIdx.x * kScanWarps + (threadIdx.x >> 5);
  const long g = w / span;
  const long c = w - g * span + c_lo;
  if (g >= num_tiles_dev(B)) return;
  const long b = g * kLanes + lane, bp = b < B ? b : B - 1;
  Params<T, N, P, AM> prm;
  load_track_params<T, N, P, AM>(prm, C, A, Q, shared_mask, bp);
  T ms[N], Vs[NS];
  const bool terminal = (c == nc - 1);
  if (!terminal) {
    const T* s = binit + ((g * nc + c) * SG) * kLanes + lane;
#pragma unroll
    for (int i = 0; i < N; ++i) ms[i] = s[i * kLanes];
#pragma unroll
    for (int i = 0; i < NS; ++i) Vs[i] = s[(N + i) * kLanes];
  }
  const long t0 = c * L, t1 = (c + 1) * L < Tn ? (c + 1) * L : Tn;
  const long base = (g * Tn + t0) * kLanes;
  bool ok = rts_backward_range<T, N, P, AM, kLanes, 2>(prm, filt + base * SG + lane, smth + base * SG + lane,
                                                      ym ? ym + base * P + lane : nullptr, nullptr, nullptr, t1 - t0,
                                                      terminal, ms, Vs);
  if (b < B && status && !ok) status[b] = 1;
}

}  // namespace eksb

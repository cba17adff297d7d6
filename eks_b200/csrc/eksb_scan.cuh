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
// Early write-out (eksb_math.cuh scan_fwd_chunk / scan_bwd_chunk): once a chunk's summary has forgotten
// the chunk's unknown boundary state (|Phi| resp. |E| numerically zero for the whole warp), the summary
// kernel's running state is the exact state, so it writes the per-frame records itself for the rest of
// the chunk and records the cut; the re-run kernel then only covers the K frames before the cut.
// HBM traffic: (1 + K/L) reads of the pass's input + 1 write, against 2 reads + 1 write without it.
#pragma once
#include "eksb_common.cuh"
#include "eksb_ring.cuh"

namespace eksb {

constexpr int kScanWarps = 4;   // warps (= chunks) per CTA

template <typename T, int N, int P, int AM> struct FwdTailRing {
  T* ring;
  __device__ bool operator()(const Params<T, N, P, AM>& prm, const T* obs, T* filt, long n, T* m, T* V, const T* mu,
                             bool* nonpos) const {
    return kf_forward_range_ring<T, N, P, AM>(prm, obs, filt, n, kContinue, m, V, ring, nullptr, mu, nonpos);
  }
};
template <typename T, int N, int P, int AM> struct BwdTailRing {
  T* ring;
  __device__ bool operator()(const Params<T, N, P, AM>& prm, const T* filt, T* smth, T* ym, long n, T* ms, T* Vs) const {
    return rts_backward_range_ring<T, N, P, AM>(prm, filt, smth, ym, n, false, ms, Vs, ring);
  }
};

// ---- 1a. forward chunk summaries, chunks [1, nc), plus the sequential run of chunk 0 -----------------
// Work items [0, tiles): chunk 0 of each tile, run as the ordinary filter from the prior (it needs no
// summary, only its end state, and would otherwise be a serial 1-warp-per-tile launch of its own).
// Work items [tiles, tiles*nc): summary of chunk c >= 1.  Once the sensitivity Phi of a chunk's running
// mean to its unknown start state has decayed below ScanCut for all 32 tracks of the warp (the filter
// forgets its start exponentially), J and eta have stopped changing and Phi is numerically zero, so the
// rest of the chunk is the plain filter (2.4x cheaper) whose records are final and written at once;
// fcut[g*nc+c] = number of leading frames of the chunk that still need the re-run.
// three CTAs (12 warps) per SM for state dim <= 3, as the chunk policy assumes (eksb_scan_impl.cuh scan_auto_chunk)
template <typename T, int N, int P, int AM, bool OFF>
__global__ void __launch_bounds__(kScanWarps* kLanes, N <= 3 ? 3 : 1)
scan_fwd_summary_kernel(const T* __restrict__ obs, const T* __restrict__ m0, const T* __restrict__ S0,
                        const T* __restrict__ C, const T* __restrict__ A, const T* __restrict__ Q, int shared_mask,
                        long B, long Tn, long L, long nc, int early_out, int first, const T* __restrict__ yoff,
                        T* __restrict__ fsum, long* __restrict__ fcut, int* __restrict__ alive, T* __restrict__ filt,
                        int* __restrict__ status, const int* __restrict__ gate) {
  if (gate && *gate == 0) return;   // fallback path of the time-split scan: only when some chunk never forgot its start
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, R2 = 2 * P, SR = ScanRec<N>::FWD;
  extern __shared__ __align__(16) unsigned char scan_smem[];
  const int lane = threadIdx.x & 31;
  T* ring = reinterpret_cast<T*>(scan_smem + (threadIdx.x >> 5) * ring_bytes_per_warp<T, R2>()) + lane;
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
    bool ok0 = true, nonpos0 = false;
    if (first == kPriorOnly) {   // eks/newton_eks.py:132 : P starts at inv(S0)
      T Vi[NS];
      sym_inv<T, N>(V, Vi, ok0);
#pragma unroll
      for (int i = 0; i < NS; ++i) V[i] = Vi[i];
    }
    const T* mu = OFF ? yoff + bp * P : nullptr;           // read per frame (L1-resident); OFF is compile time so that
    const long t1 = L < Tn ? L : Tn;                       // the common no-offset kernel carries no extra registers
    ok0 = kf_forward_range_ring<T, N, P, AM>(prm, obs + (g * Tn * R2) * kLanes + lane,
                                             filt + (g * Tn * SG) * kLanes + lane, t1, first, m, V, ring, nullptr,
                                             mu, &nonpos0) && ok0;
    if (b < B && status) set_status(status + b, ok0, nonpos0);
    return;
  }
  const long wi = w - tiles;                               // work item = g*(nc-1) + (c-1)
  const long g = wi / (nc - 1);
  const long c = wi - g * (nc - 1) + 1;
  const long b = g * kLanes + lane, bp = b < B ? b : B - 1;
  load_track_params<T, N, P, AM>(prm, C, A, Q, shared_mask, bp);
  bool ok = true, nonpos = false;
  const long t0 = c * L, t1 = (c + 1) * L < Tn ? (c + 1) * L : Tn;
  T sum[SR];
  const T* mu = OFF ? yoff + bp * P : nullptr;
  const long cut = scan_fwd_chunk<T, N, P, AM, kLanes>(prm, obs + ((g * Tn + t0) * R2) * kLanes + lane,
                                                       filt + ((g * Tn + t0) * SG) * kLanes + lane, t1 - t0,
                                                       early_out != 0, sum, ok, FwdTailRing<T, N, P, AM>{ring}, mu, &nonpos);
  if (lane == 0) {
    fcut[g * nc + c] = cut;
    if (cut >= t1 - t0) atomicOr(alive, 1);                // its last frame is not final: the fold has to run
  }
  T* out = fsum + ((g * nc + c) * SR) * kLanes + lane;
#pragma unroll
  for (int r = 0; r < SR; ++r) out[r * kLanes] = sum[r];
  if (b < B && status) set_status(status + b, ok, nonpos);
}

// ---- 2a. forward fold: exact filtered state at the frame before every chunk ---------------------
// Only needed when some chunk did not forget its start (alive != 0); a chunk that did has already written
// its exact last frame, which is the next chunk's start state.
template <typename T, int N>
__global__ void __launch_bounds__(kLanes)
scan_fwd_fold_kernel(const T* __restrict__ filt, const T* __restrict__ fsum, long B, long Tn, long L, long nc,
                     const int* __restrict__ alive, T* __restrict__ finit, int* __restrict__ status, const int* __restrict__ gate) {
  if (gate && *gate == 0) return;   // fallback path of the time-split scan: only when some chunk never forgot its start
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, SR = ScanRec<N>::FWD;
  if (*alive == 0) return;       // every chunk wrote its own last frame: the re-run reads its start states from filt
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
  if (b < B && status) set_status(status + b, ok, false);
}

// ---- 3a. forward re-run of chunks [c_lo, c_hi) ---------------------------------------------------
template <typename T, int N, int P, int AM, bool OFF>
__global__ void __launch_bounds__(kScanWarps* kLanes)
scan_fwd_rerun_kernel(const T* __restrict__ obs, const T* __restrict__ m0, const T* __restrict__ S0,
                      const T* __restrict__ C, const T* __restrict__ A, const T* __restrict__ Q, int shared_mask, long B,
                      long Tn, long L, long nc, long c_lo, long c_hi, int first0, const T* __restrict__ yoff,
                      const T* __restrict__ finit, const long* __restrict__ fcut, T* filt, int* __restrict__ status, const int* __restrict__ gate) {
  if (gate && *gate == 0) return;   // fallback path of the time-split scan: only when some chunk never forgot its start
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
  bool ok = true;
  if (c == 0) {
    const T* pm = m0 + ((shared_mask & EKSB_SHARED_M0) ? 0 : bp * N);
    const T* ps = S0 + ((shared_mask & EKSB_SHARED_S0) ? 0 : bp * (N * N));
#pragma unroll
    for (int i = 0; i < N; ++i) m[i] = pm[i];
    load_sym<T, N>(V, ps);
    first = first0;
    if (first == kPriorOnly) {
      T Vi[NS];
      sym_inv<T, N>(V, Vi, ok);
#pragma unroll
      for (int i = 0; i < NS; ++i) V[i] = Vi[i];
    }
  } else {
    // start state: the last frame of chunk c-1 if that chunk wrote it itself (chunk 0 always did), else the fold's
    const bool prev_final = (c == 1) || (fcut && fcut[g * nc + c - 1] < L);
    const T* s = prev_final ? filt + ((g * Tn + c * L - 1) * SG) * kLanes + lane
                            : finit + ((g * nc + c) * SG) * kLanes + lane;
#pragma unroll
    for (int i = 0; i < N; ++i) m[i] = s[i * kLanes];
#pragma unroll
    for (int i = 0; i < NS; ++i) V[i] = s[(N + i) * kLanes];
  }
  const long t0 = c * L;
  long t1 = (c + 1) * L < Tn ? (c + 1) * L : Tn;
  if (fcut && c > 0) t1 = t0 + fcut[g * nc + c];           // the summary kernel wrote the frames from the cut on
  if (t1 <= t0) return;
  double acc = 0.0;
  bool nonpos = false;
  ok = kf_forward_range<T, N, P, AM, kLanes, 2, false>(prm, obs + ((g * Tn + t0) * R2) * kLanes + lane,
                                                       filt + ((g * Tn + t0) * SG) * kLanes + lane, t1 - t0, first,
                                                       m, V, acc, OFF ? yoff + bp * P : nullptr, &nonpos) && ok;
  if (b < B && status) set_status(status + b, ok, nonpos);
}

// ---- 1b. backward chunk summaries, chunks [0, nc-1), plus the sequential run of the last chunk --------
// Work items [0, tiles): the terminal chunk of each tile, smoothed directly (ms = mf at T-1); it only
// touches its own frames, so it can ride in the launch that summarises the other chunks (smth may alias
// filt).  Work items [tiles, tiles*nc): (E, g, L) summary of chunk c < nc-1; once E has died out the
// chunk's remaining (earlier) frames are smoothed and written here, bcut[g*nc+c] = first frame offset
// of the chunk that still needs the re-run.
template <typename T, int N, int P, int AM>
__global__ void __launch_bounds__(kScanWarps* kLanes, N <= 3 ? 3 : 1)
scan_bwd_summary_kernel(const T* filt, const T* __restrict__ C, const T* __restrict__ A,
                        const T* __restrict__ Q, int shared_mask, long B, long Tn, long L, long nc, int early_out,
                        T* __restrict__ bsum, long* __restrict__ bcut, int* __restrict__ alive, T* smth, T* ym,
                        int* __restrict__ status, const int* __restrict__ gate) {
  if (gate && *gate == 0) return;   // fallback path of the time-split scan: only when some chunk never forgot its start
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, SR = ScanRec<N>::BWD;
  extern __shared__ __align__(16) unsigned char scan_smem[];
  const int lane = threadIdx.x & 31;
  T* ring = reinterpret_cast<T*>(scan_smem + (threadIdx.x >> 5) * ring_bytes_per_warp<T, SG>()) + lane;
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
    const bool ok0 = rts_backward_range_ring<T, N, P, AM>(prm, filt + base * SG + lane, smth + base * SG + lane,
                                                          ym ? ym + base * P + lane : nullptr, Tn - t0, true, ms, Vs, ring);
    if (b < B && status) set_status(status + b, ok0, false);
    return;
  }
  const long wi = w - tiles;
  const long g = wi / (nc - 1);
  const long c = wi - g * (nc - 1);
  const long b = g * kLanes + lane, bp = b < B ? b : B - 1;
  load_track_params<T, N, P, AM>(prm, C, A, Q, shared_mask, bp);
  T acc[SR];
  bool ok = true;
  const long base = (g * Tn + c * L) * kLanes;
  const long cut = scan_bwd_chunk<T, N, P, AM, kLanes>(prm, filt + base * SG + lane, smth + base * SG + lane,
                                                       ym ? ym + base * P + lane : nullptr, L, early_out != 0, acc, ok,
                                                       BwdTailRing<T, N, P, AM>{ring});
  if (lane == 0) {
    bcut[g * nc + c] = cut;
    if (cut <= 0) atomicOr(alive, 1);                      // its first frame is not final: the fold has to run
  }
  T* out = bsum + ((g * nc + c) * SR) * kLanes + lane;
#pragma unroll
  for (int r = 0; r < SR; ++r) out[r * kLanes] = acc[r];
  if (b < B && status) set_status(status + b, ok, false);
}

// ---- 2b. backward fold: exact smoothed state of the frame after every chunk ----------------------
template <typename T, int N>
__global__ void __launch_bounds__(kLanes)
scan_bwd_fold_kernel(const T* __restrict__ smth, const T* __restrict__ bsum, long B, long Tn, long L, long nc,
                     const int* __restrict__ alive, T* __restrict__ binit, const int* __restrict__ gate) {
  if (gate && *gate == 0) return;   // fallback path of the time-split scan: only when some chunk never forgot its start
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, SR = ScanRec<N>::BWD;
  if (*alive == 0) return;       // every chunk smoothed its own first frame: the re-run reads it from smth
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
                      const T* __restrict__ binit, const long* __restrict__ bcut, T* smth, T* ym,
                      int* __restrict__ status, const int* __restrict__ gate) {
  if (gate && *gate == 0) return;   // fallback path of the time-split scan: only when some chunk never forgot its start
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
    // smoothed state of the frame after this chunk: chunk c+1's first frame if its summary pass (or the
    // terminal run) already wrote it -- the re-run of chunk c+1 only touches frames from its cut on
    const bool next_final = (c + 1 == nc - 1) || (bcut && bcut[g * nc + c + 1] > 0);
    const T* s = next_final ? smth + ((g * Tn + (c + 1) * L) * SG) * kLanes + lane
                            : binit + ((g * nc + c) * SG) * kLanes + lane;
#pragma unroll
    for (int i = 0; i < N; ++i) ms[i] = s[i * kLanes];
#pragma unroll
    for (int i = 0; i < NS; ++i) Vs[i] = s[(N + i) * kLanes];
  }
  long t0 = c * L;
  const long t1 = (c + 1) * L < Tn ? (c + 1) * L : Tn;
  if (bcut && !terminal) t0 += bcut[g * nc + c];           // the summary kernel smoothed the frames below the cut
  const long base = (g * Tn + t0) * kLanes;
  bool ok = rts_backward_range<T, N, P, AM, kLanes, 2>(prm, filt + base * SG + lane, smth + base * SG + lane,
                                                      ym ? ym + base * P + lane : nullptr, nullptr, nullptr, t1 - t0,
                                                      terminal, ms, Vs);
  if (b < B && status) set_status(status + b, ok, false);
}

// ---------------------------------------------------------------------------------------------------------
// Time-split pipelined scan (the primary parallel-in-time route; the summary / fold / re-run kernels above are its
// fallback).  A track tile's timeline is cut into nc chunks of L frames and every (tile, chunk) pair becomes a
// "virtual tile" of the TMA-fed warp-specialised kernels (eksb_pipe.cuh), so a batch with few real tiles still puts
// ~4 tiles on every SM -- the regime in which those kernels run on the HBM roofline.  What a virtual tile needs is an
// exact start state, and the filter provides it almost for free: it forgets its start exponentially, so
//   forward : the ordinary filter started at frame c L from (0, 0), carrying the sensitivity Phi of its state to the
//             unknown state before the chunk, IS the exact filter once |Phi| < ScanCut -- after K_c frames
//             (150-800 on the shapes of BASELINE.json).  The warm-up kernel below runs just those K_c frames per
//             chunk, one warp per (tile, chunk), writes the exact filtered record of frame c L + K_c - 1 and K_c;
//             virtual tile c then covers frames [c L + K_c, (c+1) L + K_{c+1} - 1): it carries on through the next
//             chunk's warm-up frames, so nothing is run twice except the K frames of the warm-up itself;
//   backward: the same with the smoothing element (E, g, L): after K_c frames down from the chunk's end, E is dead
//             and (g, L) is the exact smoothed state of frame w_c = (c+1) L - K_c; it goes to the workspace (the
//             records may be smoothed in place, and the fallback needs them untouched) and virtual tile c covers
//             frames (w_{c-1}, w_c], emitting frame w_c from the workspace.
// A chunk that has not forgotten by its end (tiny process noise, a blind stretch) sets the gate: the virtual-tile
// kernels then exit at once and the summary / fold / re-run kernels above do the whole pass, exactly, as before.
// ---------------------------------------------------------------------------------------------------------
template <typename T, int N, int P, int AM, bool OFF>
__global__ void __launch_bounds__(kScanWarps* kLanes)
scan_fwd_warm_kernel(const T* __restrict__ obs, const T* __restrict__ C, const T* __restrict__ A, const T* __restrict__ Q,
                     int shared_mask, long B, long Tn, long L, long nc, const T* __restrict__ yoff,
                     long* __restrict__ fcut, int* __restrict__ gate, T* __restrict__ filt,
                     const T* __restrict__ q_scale, T* __restrict__ finit, long kcap) {
  // The time-split NLL sweep (eksb_nll_pipe.cuh) launches this with grid.y = candidate scales: Q is scaled by
  // q_scale[blockIdx.y], cuts and gate are per candidate, and the start states go to finit[scale][tile][chunk]
  // (there are no filtered records to put them in).  There kcap > 0 lets the warm-up run on past the chunk's end,
  // up to kcap frames (a slowly forgetting candidate: K > L is fine, the chunk's own sum just starts later); a chunk
  // that reaches the end of the track first gets K = "to the end" (it has nothing to sum) and only kcap frames
  // without forgetting raise the gate.  The scan passes q_scale = finit = nullptr, kcap = 0, grid.y = 1.
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, R2 = 2 * P;
  const int lane = threadIdx.x & 31;
  const long tiles = num_tiles_dev(B);
  const long w = (long)blockIdx.x * kScanWarps + (threadIdx.x >> 5);
  if (w >= tiles * (nc - 1)) return;
  const long g = w / (nc - 1), c = w - g * (nc - 1) + 1;
  const long b = g * kLanes + lane, bp = b < B ? b : B - 1;
  const long sidx = blockIdx.y;
  fcut += sidx * tiles * nc;
  gate += sidx;
  Params<T, N, P, AM> prm;
  load_track_params<T, N, P, AM>(prm, C, A, Q, shared_mask, bp);
  if (q_scale) {
    const T qs = q_scale[sidx];
#pragma unroll
    for (int i = 0; i < NS; ++i) prm.Q[i] *= qs;
  }
  const T* mu = OFF ? yoff + bp * P : nullptr;
  const long t0 = c * L;
  long n = ((c + 1) * L < Tn ? (c + 1) * L : Tn) - t0;
  if (kcap > 0) n = (Tn - t0 < kcap ? Tn - t0 : kcap) + 1;   // (+1: the loop below stops one frame short of n)
  const T* o = obs + ((g * Tn + t0) * R2) * kLanes + lane;
  T Phi[N][N], bb[N], Cc[NS];
#pragma unroll
  for (int i = 0; i < N; ++i) {
    bb[i] = T(0);
#pragma unroll
    for (int j = 0; j < N; ++j) Phi[i][j] = (i == j) ? T(1) : T(0);
  }
#pragma unroll
  for (int i = 0; i < NS; ++i) Cc[i] = T(0);
  // One warp per chunk and a dependent recursion: the frames come through a register ring PF deep (a single frame of
  // look-ahead left most of a global-load latency exposed in every step: ~1 us per frame).
  constexpr int PF = (R2 * (int)sizeof(T) <= 64) ? 4 : 2;
  T ring[PF][R2];
#pragma unroll
  for (int u = 0; u < PF; ++u)
    if (t0 + u < Tn) {
#pragma unroll
      for (int r = 0; r < R2; ++r) ring[u][r] = o[(u * R2 + r) * kLanes];
    }
  bool ok = true, dead = false;
  long t = 0;
  while (t < n - 1 && !dead) {                              // the cut must leave the chunk at least one frame
#pragma unroll
    for (int u = 0; u < PF; ++u) {
      if (t >= n - 1 || dead) break;
      T cur[R2], rec[Rec<N>::FWD];
#pragma unroll
      for (int r = 0; r < R2; ++r) cur[r] = ring[u][r];
      if (t0 + t + PF < Tn) {
#pragma unroll
        for (int r = 0; r < R2; ++r) ring[u][r] = o[((t + PF) * R2 + r) * kLanes];
      }
      if (OFF) {
#pragma unroll
        for (int j = 0; j < P; ++j) cur[j] -= mu[j];
      }
      fwd_prep<T, N, P, AM>(prm, cur, cur + P, rec);
      fwd_phi_step<T, N, P, AM>(prm, rec, Phi, bb, Cc, ok);
      ++t;
      if ((t & 15) == 0) {
        T mx = T(0);
#pragma unroll
        for (int i = 0; i < N; ++i) {
#pragma unroll
          for (int j = 0; j < N; ++j) mx = fmax(mx, fabs(Phi[i][j]));
        }
        if (__all_sync(0xffffffffu, mx < ScanCut<T>::value)) dead = true;
      }
    }
  }
  if (dead) {                                               // (bb, Cc) is the exact filtered state of frame t0 + t - 1
    T* f = finit ? finit + (((sidx * tiles + g) * nc + c) * SG) * kLanes + lane
                 : filt + ((g * Tn + t0 + t - 1) * SG) * kLanes + lane;
#pragma unroll
    for (int i = 0; i < N; ++i) f[i * kLanes] = bb[i];
#pragma unroll
    for (int i = 0; i < NS; ++i) f[(N + i) * kLanes] = Cc[i];
  }
  if (lane == 0) {
    if (kcap > 0) {
      const bool ran_out = !dead && (t0 + t >= Tn);        // the track ended first: nothing left for this chunk to sum
      fcut[g * nc + c] = dead ? t : Tn - t0;
      if (!dead && !ran_out) atomicOr(gate, 1);
    } else {
      fcut[g * nc + c] = dead ? t : n;
      if (!dead) atomicOr(gate, 1);
    }
  }
}

// The forward warm-up in three roles (diagonal A, N <= 3: the scaled chain of eksb_math.cuh).  A warp's FP64 instruction
// takes its scheduler's pipe for ~4 cycles, so the one-warp kernel above runs at its instruction count: ~200 per frame,
// 0.8 us -- 1.6 ms for the K ~ 2 000 frames of a bench chunk, with most of the GPU idle.  Here one CTA owns a chunk and
// three warps on three schedulers split the frame: the helper turns obs frames into the state-independent record
// (Lam, iv) (~52), the chain warp runs the filter recursion in its scaled form (~90) and publishes (Nm, adj S,
// 1 / det Lt), from which the third warp forms the gain G = Nm adj(S) / det(Lt) and the sensitivity product
// Phi <- G A Phi (~72) and tests it.  Batches of kWarmF frames, double-buffered in shared memory, one CTA barrier per
// batch: helper on batch i, chain on i - 1, phi on i - 2.  Same contract (fcut, gate, record of frame c L + K - 1,
// q_scale / finit / kcap of the NLL sweep); K is a multiple of kWarmF instead of 16.
constexpr int kWarmF = 4;
template <typename T, int N> struct Warm3Smem {
  static constexpr size_t kTotal = sizeof(T) * kLanes * (2 * kWarmF * (Rec<N>::FWD + 2 * Dim<N>::NS + 1) + 2 * Dim<N>::SIGMA);
};
template <typename T, int N, int P, bool OFF>
__global__ void __launch_bounds__(3 * kLanes)
scan_fwd_warm3_kernel(const T* __restrict__ obs, const T* __restrict__ C, const T* __restrict__ A, const T* __restrict__ Q,
                      int shared_mask, long B, long Tn, long L, long nc, const T* __restrict__ yoff,
                      long* __restrict__ fcut, int* __restrict__ gate, T* __restrict__ filt,
                      const T* __restrict__ q_scale, T* __restrict__ finit, long kcap) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, R2 = 2 * P, REC = Rec<N>::FWD, F = kWarmF, GR = 2 * NS + 1;
  extern __shared__ __align__(16) unsigned char warm_smem[];       // Warm3Smem<T, N>::kTotal bytes (50 KB for f64, n = 3)
  typedef T (*RecB)[F][REC][kLanes];
  typedef T (*GBuf)[F][GR][kLanes];
  typedef T (*SBuf)[SG][kLanes];
  RecB recb = reinterpret_cast<RecB>(warm_smem);
  GBuf gbuf = reinterpret_cast<GBuf>(warm_smem + sizeof(T) * 2 * F * REC * kLanes);
  SBuf sbuf = reinterpret_cast<SBuf>(warm_smem + sizeof(T) * 2 * F * (REC + GR) * kLanes);
  // iteration at whose start everybody leaves; written once, by the phi warp, DURING the iteration before it (a plain
  // flag could be seen by a thread still entering the iteration it was raised in, and not by its neighbours)
  __shared__ long stop_at;
  const int lane = threadIdx.x & 31, role = threadIdx.x >> 5;      // 0 helper, 1 chain, 2 phi
  const long tiles = num_tiles_dev(B);
  const long w = blockIdx.x;                                        // (tile, chunk >= 1)
  const long g = w / (nc - 1), c = w - g * (nc - 1) + 1;
  const long b = g * kLanes + lane, bp = b < B ? b : B - 1;
  const long sidx = blockIdx.y;
  fcut += sidx * tiles * nc;
  gate += sidx;
  Params<T, N, P, 0> prm;
  load_track_params<T, N, P, 0>(prm, C, A, Q, shared_mask, bp);
  if (q_scale) {
    const T qs = q_scale[sidx];
#pragma unroll
    for (int i = 0; i < NS; ++i) prm.Q[i] *= qs;
  }
  const long t0 = c * L;
  const long clen = ((c + 1) * L < Tn ? (c + 1) * L : Tn) - t0;
  const long steps = kcap > 0 ? (Tn - t0 < kcap ? Tn - t0 : kcap) : clen - 1;   // frames the recursion may take
  const long nb = (steps + F - 1) / F;
  if (threadIdx.x == 0) stop_at = nb + 8;
  // helper state
  const T* o = obs + ((g * Tn + t0) * R2) * kLanes + lane;
  constexpr bool kOuter = (P * NS <= 36);
  T CC[kOuter ? P : 1][NS];
  T mu[P];
  T raw[F][R2];
  if (role == 0) {
    if constexpr (kOuter) make_outer<T, N, P>(prm.C, CC);
#pragma unroll
    for (int j = 0; j < P; ++j) mu[j] = OFF ? yoff[bp * P + j] : T(0);
#pragma unroll
    for (int j = 0; j < F; ++j)
      if (j < steps) {
#pragma unroll
        for (int r = 0; r < R2; ++r) raw[j][r] = o[(j * R2 + r) * kLanes];
      }
  }
  // chain state: (m, V = Nm r) started at (0, 0)
  T m[N], Nm[NS], rr = T(1);
#pragma unroll
  for (int i = 0; i < N; ++i) m[i] = T(0);
#pragma unroll
  for (int i = 0; i < NS; ++i) Nm[i] = T(0);
  // phi state
  T Phi[N][N];
#pragma unroll
  for (int i = 0; i < N; ++i) {
#pragma unroll
    for (int j = 0; j < N; ++j) Phi[i][j] = (i == j) ? T(1) : T(0);
  }
  bool dead = false;
  for (long i = 0; i < nb + 2; ++i) {
    __syncthreads();
    if (i >= stop_at) break;
    if (role == 0) {
      if (i < nb) {
#pragma unroll
        for (int j = 0; j < F; ++j) {
          if (i * F + j < steps) {
            T cur[R2], rec[REC];
#pragma unroll
            for (int r = 0; r < R2; ++r) cur[r] = raw[j][r];
            if (OFF) {
#pragma unroll
              for (int q = 0; q < P; ++q) cur[q] -= mu[q];
            }
            if constexpr (kOuter) fwd_prep_outer<T, N, P>(prm.C, CC, cur, cur + P, rec);
            else fwd_prep<T, N, P, 0>(prm, cur, cur + P, rec);
#pragma unroll
            for (int r = 0; r < REC; ++r) recb[i & 1][j][r][lane] = rec[r];
          }
        }
#pragma unroll
        for (int j = 0; j < F; ++j) {                       // the next batch's frames: in flight across the barrier
          const long t = (i + 1) * F + j;
          if (t < steps) {
#pragma unroll
            for (int r = 0; r < R2; ++r) raw[j][r] = o[(t * R2 + r) * kLanes];
          }
        }
      }
    } else if (role == 1) {
      if (i >= 1 && i <= nb) {
        const long k = i - 1;
#pragma unroll
        for (int j = 0; j < F; ++j) {
          if (k * F + j < steps) {
            T rec[REC], S[NS], adjS[NS], Lt[NS], detS, detL, Vout[NS];
#pragma unroll
            for (int r = 0; r < REC; ++r) rec[r] = recb[k & 1][j][r][lane];
#pragma unroll
            for (int q = 0; q < NS; ++q) S[q] = (prm.AA[q] * Nm[q]) * rr + prm.Q[q];
            sym_adjugate<T, N>(S, adjS, detS);
#pragma unroll
            for (int q = 0; q < NS; ++q) Lt[q] = detS * rec[q] + adjS[q];
            sym_adjugate<T, N>(Lt, Nm, detL);
            const T rcpL = rcp_chain(detL);
            rr = detS * rcpL;
            fwd_mean_scaled<T, N, P>(prm, rec, true, Nm, rr, m, Vout);
#pragma unroll
            for (int q = 0; q < NS; ++q) { gbuf[k & 1][j][q][lane] = Nm[q]; gbuf[k & 1][j][NS + q][lane] = adjS[q]; }
            gbuf[k & 1][j][2 * NS][lane] = rcpL;
            if (j == F - 1 || k * F + j == steps - 1) {      // the batch's last frame: the state a cut here would record
#pragma unroll
              for (int q = 0; q < N; ++q) sbuf[k & 1][q][lane] = m[q];
#pragma unroll
              for (int q = 0; q < NS; ++q) sbuf[k & 1][N + q][lane] = Vout[q];
            }
          }
        }
      }
    } else {
      if (i >= 2) {
        const long k = i - 2;
#pragma unroll
        for (int j = 0; j < F; ++j) {
          if (k * F + j < steps) {
            T nm[NS], aS[NS], T1[N][N];
#pragma unroll
            for (int q = 0; q < NS; ++q) { nm[q] = gbuf[k & 1][j][q][lane]; aS[q] = gbuf[k & 1][j][NS + q][lane]; }
            const T rcpL = gbuf[k & 1][j][2 * NS][lane];
#pragma unroll
            for (int a = 0; a < N; ++a) {
#pragma unroll
              for (int d = 0; d < N; ++d) {
                T acc = T(0);
#pragma unroll
                for (int q = 0; q < N; ++q) acc += aS[sym_idx<N>(a, q)] * (prm.A[q] * Phi[q][d]);
                T1[a][d] = acc;
              }
            }
#pragma unroll
            for (int a = 0; a < N; ++a) {
#pragma unroll
              for (int d = 0; d < N; ++d) {
                T acc = T(0);
#pragma unroll
                for (int q = 0; q < N; ++q) acc += nm[sym_idx<N>(a, q)] * T1[q][d];
                Phi[a][d] = acc * rcpL;
              }
            }
          }
        }
        T mx = T(0);
#pragma unroll
        for (int a = 0; a < N; ++a) {
#pragma unroll
          for (int d = 0; d < N; ++d) mx = fmax(mx, fabs(Phi[a][d]));
        }
        const long done = (k + 1) * F < steps ? (k + 1) * F : steps;      // frames the recursion has taken
        // (a cut needs a frame of the chunk left for the tile that starts there: done <= clen - 1 when kcap == 0 by
        // the choice of `steps`; with kcap the chunk's own sum may be empty)
        if (__all_sync(0xffffffffu, mx < ScanCut<T>::value)) {
          dead = true;
          T* f = finit ? finit + (((sidx * tiles + g) * nc + c) * SG) * kLanes + lane
                       : filt + ((g * Tn + t0 + done - 1) * SG) * kLanes + lane;
#pragma unroll
          for (int q = 0; q < SG; ++q) f[q * kLanes] = sbuf[k & 1][q][lane];
          if (lane == 0) { fcut[g * nc + c] = done; stop_at = i + 1; }
        }
      }
    }
  }
  if (role == 2 && !dead && lane == 0) {
    if (kcap > 0) {
      const bool ran_out = (t0 + steps >= Tn);               // the track ended first: nothing left for this chunk to sum
      fcut[g * nc + c] = Tn - t0;
      if (!ran_out) atomicOr(gate, 1);
    } else {
      fcut[g * nc + c] = clen;
      atomicOr(gate, 1);
    }
  }
}

template <typename T, int N, int P, int AM>
__global__ void __launch_bounds__(kScanWarps* kLanes)
scan_bwd_warm_kernel(const T* __restrict__ filt, const T* __restrict__ C, const T* __restrict__ A, const T* __restrict__ Q,
                     int shared_mask, long B, long Tn, long L, long nc, long* __restrict__ bcut, int* __restrict__ gate,
                     T* __restrict__ binit) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, SR = ScanRec<N>::BWD;
  const int lane = threadIdx.x & 31;
  const long tiles = num_tiles_dev(B);
  const long w = (long)blockIdx.x * kScanWarps + (threadIdx.x >> 5);
  if (w >= tiles * (nc - 1)) return;
  const long g = w / (nc - 1), c = w - g * (nc - 1);         // chunks [0, nc-1): the last one starts at the track's end
  const long b = g * kLanes + lane, bp = b < B ? b : B - 1;
  Params<T, N, P, AM> prm;
  load_track_params<T, N, P, AM>(prm, C, A, Q, shared_mask, bp);
  const T* f = filt + ((g * Tn + c * L) * SG) * kLanes + lane;   // chunk c is full length (c < nc - 1)
  T acc[SR];
#pragma unroll
  for (int i = 0; i < SR; ++i) acc[i] = T(0);
#pragma unroll
  for (int i = 0; i < N; ++i) acc[i * N + i] = T(1);
  constexpr int PF = (SG * (int)sizeof(T) <= 72) ? 4 : 2;    // register ring of records, as in the forward warm-up
  T ring[PF][SG];
#pragma unroll
  for (int u = 0; u < PF; ++u)
    if (L - 1 - u >= 0) {
#pragma unroll
      for (int r = 0; r < SG; ++r) ring[u][r] = f[((L - 1 - u) * SG + r) * kLanes];
    }
  bool ok = true, dead = false;
  long done = 0;
  while (done < L - 2 && !dead) {                            // leave the virtual tile below at least one frame of its own
#pragma unroll
    for (int u = 0; u < PF; ++u) {
      if (done >= L - 2 || dead) break;
      T cur[SG], rec[Rec<N>::BWD];
#pragma unroll
      for (int r = 0; r < SG; ++r) cur[r] = ring[u][r];     // frame L - 1 - done
      const long tn = L - 1 - done - PF;                     // the frame that takes its slot
      if (tn >= 0) {
#pragma unroll
        for (int r = 0; r < SG; ++r) ring[u][r] = f[(tn * SG + r) * kLanes];
      }
      bwd_prep<T, N, P, AM>(prm, cur, cur + N, rec, ok);
      bwd_summary_step<T, N>(rec, acc);
      ++done;
      if ((done & 15) == 0) {
        T mx = T(0);
#pragma unroll
        for (int i = 0; i < N * N; ++i) mx = fmax(mx, fabs(acc[i]));
        if (__all_sync(0xffffffffu, mx < ScanCut<T>::value)) dead = true;
      }
    }
  }
  if (dead) {                                               // (g, L) of the element = smoothed (ms, Vs) of frame (c+1) L - done
    T* o = binit + ((g * nc + c) * SG) * kLanes + lane;
#pragma unroll
    for (int i = 0; i < N; ++i) o[i * kLanes] = acc[N * N + NS + i];
#pragma unroll
    for (int i = 0; i < NS; ++i) o[(N + i) * kLanes] = acc[N * N + i];
  }
  if (lane == 0) {
    bcut[g * nc + c] = dead ? done : L;
    if (!dead) atomicOr(gate, 1);
  }
}

// The backward warm-up in two roles (any shape and type: nothing here is the scaled chain).  bwd_prep -- S = A Vf A^T + Q,
// its inverse, the RTS gain J and the frame's (L, g) -- depends on the filtered record alone, so a helper warp forms the
// records of batch i while the chain warp composes batch i - 1 onto the chunk's (E, L, g) element and tests |E|; the two
// sit on different schedulers and each carries about half of the one-warp kernel's FP64 instructions.  Same contract as
// scan_bwd_warm_kernel; K is a multiple of kWarmF instead of 16.
template <typename T, int N> struct Warm2Smem {
  static constexpr size_t kTotal = sizeof(T) * kLanes * 2 * kWarmF * Rec<N>::BWD;
};
template <typename T, int N, int P, int AM>
__global__ void __launch_bounds__(2 * kLanes)
scan_bwd_warm2_kernel(const T* __restrict__ filt, const T* __restrict__ C, const T* __restrict__ A, const T* __restrict__ Q,
                      int shared_mask, long B, long Tn, long L, long nc, long* __restrict__ bcut, int* __restrict__ gate,
                      T* __restrict__ binit) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, SR = ScanRec<N>::BWD, REC = Rec<N>::BWD, F = kWarmF;
  extern __shared__ __align__(16) unsigned char warm_smem[];
  typedef T (*RecB)[F][REC][kLanes];
  RecB recb = reinterpret_cast<RecB>(warm_smem);
  __shared__ long stop_at;                                   // as in scan_fwd_warm3_kernel
  const int lane = threadIdx.x & 31, role = threadIdx.x >> 5;   // 0 helper, 1 chain
  const long w = blockIdx.x;
  const long g = w / (nc - 1), c = w - g * (nc - 1);         // chunks [0, nc-1): the last one starts at the track's end
  const long b = g * kLanes + lane, bp = b < B ? b : B - 1;
  Params<T, N, P, AM> prm;
  load_track_params<T, N, P, AM>(prm, C, A, Q, shared_mask, bp);
  const T* f = filt + ((g * Tn + c * L) * SG) * kLanes + lane;   // chunk c is full length (c < nc - 1)
  const long steps = L - 2 > 0 ? L - 2 : 0;                  // leave the virtual tile below at least one frame of its own
  const long nb = (steps + F - 1) / F;
  if (threadIdx.x == 0) stop_at = nb + 8;
  T raw[F][SG];
  if (role == 0) {
#pragma unroll
    for (int j = 0; j < F; ++j)
      if (j < steps) {
#pragma unroll
        for (int r = 0; r < SG; ++r) raw[j][r] = f[((L - 1 - j) * SG + r) * kLanes];
      }
  }
  T acc[SR];
#pragma unroll
  for (int i = 0; i < SR; ++i) acc[i] = T(0);
#pragma unroll
  for (int i = 0; i < N; ++i) acc[i * N + i] = T(1);
  bool ok = true, dead = false;
  for (long i = 0; i < nb + 1; ++i) {
    __syncthreads();
    if (i >= stop_at) break;
    if (role == 0) {
      if (i < nb) {
#pragma unroll
        for (int j = 0; j < F; ++j) {
          if (i * F + j < steps) {
            T cur[SG], rec[REC];
#pragma unroll
            for (int r = 0; r < SG; ++r) cur[r] = raw[j][r];
            bwd_prep<T, N, P, AM>(prm, cur, cur + N, rec, ok);
#pragma unroll
            for (int r = 0; r < REC; ++r) recb[i & 1][j][r][lane] = rec[r];
          }
        }
#pragma unroll
        for (int j = 0; j < F; ++j) {                       // the next batch's records: in flight across the barrier
          const long t = (i + 1) * F + j;
          if (t < steps) {
#pragma unroll
            for (int r = 0; r < SG; ++r) raw[j][r] = f[((L - 1 - t) * SG + r) * kLanes];
          }
        }
      }
    } else if (i >= 1) {
      const long k = i - 1;
#pragma unroll
      for (int j = 0; j < F; ++j) {
        if (k * F + j < steps) {
          T rec[REC];
#pragma unroll
          for (int r = 0; r < REC; ++r) rec[r] = recb[k & 1][j][r][lane];
          bwd_summary_step<T, N>(rec, acc);
        }
      }
      T mx = T(0);
#pragma unroll
      for (int q = 0; q < N * N; ++q) mx = fmax(mx, fabs(acc[q]));
      const long done = (k + 1) * F < steps ? (k + 1) * F : steps;
      if (__all_sync(0xffffffffu, mx < ScanCut<T>::value)) {   // (g, L) of the element = smoothed (ms, Vs) of frame (c+1) L - done
        dead = true;
        T* o = binit + ((g * nc + c) * SG) * kLanes + lane;
#pragma unroll
        for (int q = 0; q < N; ++q) o[q * kLanes] = acc[N * N + NS + q];
#pragma unroll
        for (int q = 0; q < NS; ++q) o[(N + q) * kLanes] = acc[N * N + q];
        if (lane == 0) { bcut[g * nc + c] = done; stop_at = i + 1; }
      }
    }
  }
  if (role == 1 && !dead && lane == 0) {
    bcut[g * nc + c] = L;
    atomicOr(gate, 1);
  }
}

}  // namespace eksb

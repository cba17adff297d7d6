// eksb_setup.cu -- K8: per-track Kalman parameter setup on the device.
//
// Replaces, for B tracks at once and without leaving HBM, the reference's per-keypoint host block
// multiview_pca_smoother.py:684-731:
//   max_vars = max_j ev[t,j]; good = max_vars <= np.percentile(max_vars, q)      (:686-688)
//   means_j  = mean over good frames of preds[:,j]                               (:698-700)
//   PCA(3) of the centred good rows -> W = components_ (3,p)                     (:707-708, :29-31)
//   S0 = diag(var(good pcs)) [ddof 0]; Q = s * cov(diff of consecutive good pcs) [ddof 1]  (:722-728)
//   C = W^T                                                                      (:730)
// Everything downstream of the good-frame mask is linear in y, so ONE pass over the obs records
// accumulates the masked first/second moments of y and of dy (difference of consecutive good frames)
// in the observation space; W, S0 = diag(W Cov0 W^T) and Q = s W Cov_d W^T follow per track from a
// p x p Jacobi eigen-decomposition.  The percentile is an exact order statistic: an 8-bit MSB radix
// select over the IEEE bit patterns of max_vars (non-negative, so bit order == value order), then
// numpy's linear interpolation formula, evaluated without FMA contraction so the threshold (and
// with it the mask) is bit-identical to numpy's.
// Accumulation and the eigen-solve are double precision for both storage types.
#include <stdlib.h>
#include "eksb_common.cuh"

namespace eksb {

constexpr int kMomWarpsPerCta = 4;
// setup_moments_kernel, tried and measured slower per 8192 x 100000 f64 call (23.3 ms, of which this kernel 6.2 ms):
// __launch_bounds__(128, 3) (168 registers, a few spills) 28.0 ms; frames fed by cp.async.bulk + mbarrier through a
// per-warp ring of 4 stages x 4 frames in shared memory (134 registers) 23.9 ms -- one 1 KB copy per frame (the y rows
// of consecutive frames are not contiguous) costs lane 0 more issue time than the register staging costs in waits.

// time slices per tile: enough (tile, slice) work items to fill the GPU even for a few long tracks
inline int select_slices(long tiles, long Tn) {      // one 8-warp CTA per (tile, slice)
  long s = ((long)sm_count() * 4 + tiles - 1) / tiles;
  const long cap = (Tn + 2047) / 2048;
  if (s > cap) s = cap;
  return (int)(s < 1 ? 1 : s);
}
inline int moment_slices(long tiles, long Tn) {      // one warp per (tile, slice)
  long s = ((long)sm_count() * 32 + tiles - 1) / tiles;
  if (s < 16) s = 16;
  const long cap = (Tn + 255) / 256;
  if (s > cap) s = cap;
  if (s > 4096) s = 4096;
  return (int)(s < 1 ? 1 : s);
}
template <int P> constexpr int mom_rows() { return 2 + 4 * P + P * (P + 1); }
constexpr int kCandCap = 8192;    // candidate slots per track after the full-length select pass(es)
// Wide first digit of the select (round 2): sign + exponent + kWideMant mantissa bits, i.e. 8 bins per binade, in a
// window of kWideBins bins centred on 1.0 (2^-32 .. 2^32); + one bin below and one above the window.
constexpr int kMomGroups = 32;    // few tiles, long tracks: the moment slices are folded to this many before the finalize
constexpr int kWideMant = 3, kWideBins = 512, kWideRows = kWideBins + 2, kWideHalfBinades = 32;
struct SetupWs {                  // workspace layout in 8-byte words
  size_t thr, state, ghist, ghist2, gwide, part2, part, cand, flag, mx, total;
};
inline SetupWs setup_layout(long B, long Tn, int p, int elem_size) {
  const size_t tiles = (size_t)num_tiles(B);
  const int npass = elem_size;                       // 8-bit digits: 8 passes for f64, 4 for f32
  SetupWs w;
  w.thr = 0;
  w.state = w.thr + tiles * kLanes;
  w.ghist = w.state + tiles * 6 * kLanes;
  w.ghist2 = w.ghist + (size_t)npass * tiles * 256 * kLanes / 2;      // short passes of the wide-digit path
  w.gwide = w.ghist2 + (size_t)npass * tiles * 256 * kLanes / 2;
  w.part2 = w.gwide + tiles * (size_t)kWideRows * kLanes / 2;         // moment partials folded to kMomGroups per tile
  w.part = w.part2 + tiles * (size_t)kMomGroups * (2 + 4 * p + p * (p + 1)) * kLanes;
  w.cand = w.part + tiles * (size_t)moment_slices((long)tiles, Tn) * (2 + 4 * p + p * (p + 1)) * kLanes;
  w.flag = w.cand + (tiles * (size_t)kCandCap * kLanes * elem_size + 7) / 8;
  w.mx = w.flag + 2;                                 // ints: [0] candidate overflow, [1] fallback to the 8-bit passes, [2] deepen, [3] 1|2
  w.total = w.mx + (tiles * (size_t)Tn * kLanes * elem_size + 7) / 8;
  return w;
}

template <typename T> struct KeyOf;
template <> struct KeyOf<double> {
  using type = unsigned long long;
  static constexpr int kBits = 64;
  static __device__ __forceinline__ type of(double v) { return (type)__double_as_longlong(v); }
  static __device__ __forceinline__ double back(type k) { return __longlong_as_double((long long)k); }
};
template <> struct KeyOf<float> {
  using type = unsigned int;
  static constexpr int kBits = 32;
  static __device__ __forceinline__ type of(float v) { return __float_as_uint(v); }
  static __device__ __forceinline__ float back(type k) { return __uint_as_float(k); }
};

// max over the P ensemble-variance rows of one frame (rows [P,2P) of the obs record); -0.0 -> +0.0
template <typename T, int P> __device__ __forceinline__ T frame_max_var(const T* rec_lane) {
  T m = rec_lane[P * kLanes];
#pragma unroll
  for (int j = 1; j < P; ++j) {
    const T v = rec_lane[(P + j) * kLanes];
    m = v > m ? v : m;
  }
  return m + T(0);
}

// numpy.percentile(method='linear') position and interpolation (numpy/lib/_function_base_impl.py
// _quantile / _lerp), in plain IEEE double operations.
__host__ __device__ inline void percentile_position(long T, double q, long& lo, double& gamma) {
  const double quant = q / 100.0;
  double pos = (double)(T - 1) * quant;
  if (pos < 0) pos = 0;
  if (pos > (double)(T - 1)) pos = (double)(T - 1);
  lo = (long)pos;                 // floor (pos >= 0)
  gamma = pos - (double)lo;
}

// ---- 1. per-track percentile threshold ------------------------------------------------------------
// Parallel over (tile, time slice) so that a handful of long tracks still fills the GPU:
//   mx kernel    : max_vars[t] per track, once, into a compact [tile][t][32] buffer (8 B / track-frame)
//   8 (4) passes : 8-bit digit histogram of the keys that match the prefix selected so far; per-CTA
//                  shared-memory histogram [256][32 lanes], flushed with one global atomic per non-empty bin
//   pick kernel  : one warp per tile walks its 256 bins to the digit holding rank k (between passes)
//   collect      : after 3 (f64) / 2 (f32) passes only a handful of keys per track still match the prefix; they
//                  are copied to a compact [tile][kCandCap][32] buffer (plus the smallest key above the prefix) and
//                  the remaining passes and the tail run on that buffer instead of the T-long one -- 5 reads of
//                  max_vars per track-frame instead of 9.  A track with more than kCandCap matching keys (massive
//                  ties) raises the overflow flag and every later kernel goes back to the full buffer.
//   tail kernel  : count(key <= v_lo) and min(key > v_lo)  ->  sorted[lo+1]
//   thr kernel   : numpy's interpolation
struct SelState {            // per track, [tile][field][32 lanes] as 64-bit words
  static constexpr int PREFIX = 0, RANK = 1, CNT_LE = 2, NEXT = 3, NCAND = 4, RANK_C = 5, WORDS = 6;
  // NEXT: min(key > v_lo); the collect kernel seeds it with the smallest key above the candidates' prefix, the
  // tail kernel lowers it to the smallest candidate above v_lo if there is one.  NCAND: candidates appended.
  // RANK_C: rank of the wanted key among the candidates at the time they were collected.
};

template <typename T, int P>
__global__ void __launch_bounds__(256)
setup_mx_kernel(const T* __restrict__ obs, long tiles, long Tn, T* __restrict__ mx) {
  const long line = (long)blockIdx.x * 8 + (threadIdx.x >> 5);      // (g, t)
  const int lane = threadIdx.x & 31;
  if (line >= tiles * Tn) return;
  mx[line * kLanes + lane] = frame_max_var<T, P>(obs + line * (2 * P) * kLanes + lane);
}

template <typename T>
__global__ void __launch_bounds__(256)
setup_hist_kernel(const T* __restrict__ mx, long Tn, const T* __restrict__ cand, const int* __restrict__ overflow,
                  int slices, int shift, const unsigned long long* __restrict__ state,
                  unsigned int* __restrict__ ghist /* [tile][256][32], this pass */, int match_shift,
                  const int* __restrict__ gate, int gate_want, const int* __restrict__ deep = nullptr, int deep_match = 0) {
  using K = KeyOf<T>;
  using Key = typename K::type;
  __shared__ unsigned int hist[256][kLanes];
  if (gate && (*gate != 0) != (gate_want != 0)) return;    // wide-digit path vs its 8-bit fallback
  // after the extra full-length digit of the wide path the rank counts inside the longer prefix: match all of it
  // (the same keys on the candidates; it matters when the remaining passes run at full length)
  if (deep && *deep && deep_match < match_shift) match_shift = deep_match;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long g = blockIdx.x / slices;
  const int sl = (int)(blockIdx.x - g * slices);
  const bool use_c = cand && *overflow == 0;               // candidates collected: the short buffer holds every match
  long n = Tn;
  if (use_c)                                               // ... in its first max-over-the-tile NCAND slots
    n = (long)__reduce_max_sync(0xffffffffu, (unsigned)state[(g * SelState::WORDS + SelState::NCAND) * kLanes + lane]);
  const long len = (n + slices - 1) / slices;
  const long t0 = sl * len, t1 = (t0 + len < n) ? t0 + len : n;
  if (t0 >= t1) return;
  for (int i = threadIdx.x; i < 256 * kLanes; i += blockDim.x) (&hist[0][0])[i] = 0;
  __syncthreads();
  const Key prefix = (Key)state[(g * SelState::WORDS + SelState::PREFIX) * kLanes + lane];
  const bool top = (match_shift >= K::kBits);              // no prefix yet: every key takes part
  const T* base = (use_c ? cand + g * kCandCap * kLanes : mx + g * Tn * kLanes) + lane;
  for (long t = t0 + warp; t < t1; t += 8) {
    const Key key = K::of(base[t * kLanes]);
    if (use_c && key == (Key) ~(Key)0) continue;          // empty candidate slot
    if (top || (key >> (match_shift & (K::kBits - 1))) == (prefix >> (match_shift & (K::kBits - 1))))
      atomicAdd(&hist[(unsigned)(key >> shift) & 255u][lane], 1u);
  }
  __syncthreads();
  unsigned int* gh = ghist + g * 256 * kLanes;
  for (int i = threadIdx.x; i < 256 * kLanes; i += blockDim.x) {
    const unsigned int c = (&hist[0][0])[i];
    if (c) atomicAdd(&gh[i], c);
  }
}

// One CTA per tile: the tile's [256][32] histogram comes into shared memory with coalesced loads (a lone warp
// walking it in global memory paid one dependent L2 round trip per bin: 0.1 ms per pass, 0.7 ms per setup call),
// then warp 0 (lane = track) walks its 256 bins there.
template <typename T>
__global__ void __launch_bounds__(256)
setup_pick_kernel(const unsigned int* __restrict__ ghist, int shift, unsigned long long* __restrict__ state,
                  const int* __restrict__ gate, int gate_want) {
  using Key = typename KeyOf<T>::type;
  __shared__ unsigned int hist[256 * kLanes];
  if (gate && (*gate != 0) != (gate_want != 0)) return;
  const long g = blockIdx.x;
  const uint4* src = reinterpret_cast<const uint4*>(ghist + g * 256 * kLanes);
  uint4* dst = reinterpret_cast<uint4*>(hist);
  for (int i = threadIdx.x; i < 256 * kLanes / 4; i += blockDim.x) dst[i] = src[i];
  __syncthreads();
  if (threadIdx.x >= kLanes) return;
  const int lane = threadIdx.x;
  unsigned long long* st = state + g * SelState::WORDS * kLanes + lane;
  long k = (long)st[SelState::RANK * kLanes];
  int d = 0;
  for (; d < 255; ++d) {
    const unsigned c = hist[d * kLanes + lane];
    if (k < (long)c) break;
    k -= c;
  }
  st[SelState::RANK * kLanes] = (unsigned long long)k;
  st[SelState::PREFIX * kLanes] |= (unsigned long long)((Key)d << shift);
}

template <typename T>
__global__ void __launch_bounds__(256)
setup_tail_kernel(const T* __restrict__ mx, long Tn, const T* __restrict__ cand, const int* __restrict__ overflow,
                  int slices, unsigned long long* __restrict__ state) {
  using K = KeyOf<T>;
  using Key = typename K::type;
  __shared__ unsigned int s_cnt[kLanes];
  __shared__ unsigned long long s_next[kLanes];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long g = blockIdx.x / slices;
  const int sl = (int)(blockIdx.x - g * slices);
  const bool use_c = cand && *overflow == 0;
  long n = Tn;
  if (use_c) n = (long)__reduce_max_sync(0xffffffffu, (unsigned)state[(g * SelState::WORDS + SelState::NCAND) * kLanes + lane]);
  const long len = (n + slices - 1) / slices;
  const long t0 = sl * len, t1 = (t0 + len < n) ? t0 + len : n;
  if (t0 >= t1) return;
  if (warp == 0) { s_cnt[lane] = 0; s_next[lane] = ~0ull; }
  __syncthreads();
  unsigned long long* st = state + g * SelState::WORDS * kLanes + lane;
  const Key klo = (Key)st[SelState::PREFIX * kLanes];
  const T* base = (use_c ? cand + g * kCandCap * kLanes : mx + g * Tn * kLanes) + lane;
  unsigned cnt = 0;
  unsigned long long nxt = ~0ull;
  for (long t = t0 + warp; t < t1; t += 8) {
    const Key key = K::of(base[t * kLanes]);
    if (use_c && key == (Key) ~(Key)0) continue;          // empty candidate slot
    if (key <= klo) ++cnt;
    else if ((unsigned long long)key < nxt) nxt = key;
  }
  atomicAdd(&s_cnt[lane], cnt);
  atomicMin(&s_next[lane], nxt);
  __syncthreads();
  if (warp == 0) {
    if (s_cnt[lane]) atomicAdd(&st[SelState::CNT_LE * kLanes], (unsigned long long)s_cnt[lane]);
    if (s_next[lane] != ~0ull) atomicMin(&st[SelState::NEXT * kLanes], s_next[lane]);
  }
}

// Copy the keys that still match the selected prefix (bits >= low_bit) to the compact candidate buffer and
// find the smallest key above that prefix (it is sorted[lo+1] when the rank-lo key is the largest candidate).
template <typename T>
__global__ void __launch_bounds__(256)
setup_collect_kernel(const T* __restrict__ mx, long Tn, int slices, int low_bit, unsigned long long* __restrict__ state,
                     T* __restrict__ cand, int* __restrict__ overflow, const int* __restrict__ gate, int gate_want) {
  using K = KeyOf<T>;
  using Key = typename K::type;
  // The wide-digit path collects ~T/25 keys per track: one global atomic per key on the track's counter was half of
  // this kernel's time.  Each thread (one track, every 8th frame of the slice) stages its matches in a private strip
  // of shared memory and reserves their slots with one atomic per kStage matches; no CTA-wide synchronisation.
  constexpr int kStage = 8;
  __shared__ unsigned long long s_above[kLanes];
  __shared__ T s_buf[8][kStage][kLanes];
  if (gate && (*gate != 0) != (gate_want != 0)) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long g = blockIdx.x / slices;
  const int sl = (int)(blockIdx.x - g * slices);
  if (warp == 0) s_above[lane] = ~0ull;
  __syncthreads();
  unsigned long long* st = state + g * SelState::WORDS * kLanes + lane;
  const Key pp = (Key)st[SelState::PREFIX * kLanes] >> low_bit;
  if (sl == 0 && warp == 0) st[SelState::RANK_C * kLanes] = st[SelState::RANK * kLanes];
  const long len = (Tn + slices - 1) / slices;
  const long t0 = sl * len, t1 = (t0 + len < Tn) ? t0 + len : Tn;
  const T* base = mx + g * Tn * kLanes + lane;
  T* out = cand + g * kCandCap * kLanes + lane;
  unsigned long long above = ~0ull;
  int cnt = 0;
  auto flush = [&]() {
    if (cnt == 0) return;
    unsigned long long idx = atomicAdd(&st[SelState::NCAND * kLanes], (unsigned long long)cnt);
    for (int i = 0; i < cnt; ++i, ++idx) {
      if (idx < (unsigned long long)kCandCap) out[idx * kLanes] = s_buf[warp][i][lane];
      else atomicExch(overflow, 1);
    }
    cnt = 0;
  };
  auto take = [&](T v) {
    const Key key = K::of(v);
    const Key kp = key >> low_bit;
    if (kp == pp) {
      s_buf[warp][cnt++][lane] = v;
    } else if (kp > pp && (unsigned long long)key < above) {
      above = key;
    }
  };
  long t = t0 + warp;
  for (; t + 24 < t1; t += 32) {                           // four lines in flight per warp
    const T v0 = base[t * kLanes], v1 = base[(t + 8) * kLanes], v2 = base[(t + 16) * kLanes], v3 = base[(t + 24) * kLanes];
    if (cnt > kStage - 4) flush();
    take(v0); take(v1); take(v2); take(v3);
  }
  for (; t < t1; t += 8) {
    if (cnt >= kStage) flush();
    take(base[t * kLanes]);
  }
  flush();
  if (above != ~0ull) atomicMin(&s_above[lane], above);
  __syncthreads();
  if (warp == 0 && s_above[lane] != ~0ull) atomicMin(&st[SelState::NEXT * kLanes], s_above[lane]);
}

// ---- wide first digit (round 2): ONE full-length pass instead of three ------------------------------------------
// The top byte of a positive float is almost all exponent: ensemble variances of a session span a few binades, so the
// first 8-bit pass fills 3 bins and the select needs three full-length passes before a handful of keys are left.  Here
// the first digit is sign + exponent + 3 mantissa bits inside a window of 64 binades around 1.0: one pass leaves
// ~T/50 keys per track (one eighth of a binade), which are collected at once; the remaining bits are resolved on the
// candidates.  Off the common case, decided on the device behind flags (the other side's kernels exit at once):
//   * the bin holds more keys than candidate slots (the pick kernel has the count): one more full-length 8-bit digit
//     below the wide one, then collect -- four full-length reads, what round 1 always paid;
//   * still more than the slots (tied keys): the remaining passes run at full length, as in round 1;
//   * the rank falls below / above the window (zero or absurd variances): the round-1 8-bit passes from the top.
template <typename T> struct WideOf;
template <> struct WideOf<double> { static constexpr int shift = 52 - kWideMant; static constexpr long base = (long)(1023 - kWideHalfBinades) << kWideMant; };
template <> struct WideOf<float> { static constexpr int shift = 23 - kWideMant; static constexpr long base = (long)(127 - kWideHalfBinades) << kWideMant; };

template <typename T>
__global__ void __launch_bounds__(256)
setup_wide_kernel(const T* __restrict__ mx, long Tn, int slices, unsigned int* __restrict__ gwide /* [tile][kWideRows][32] */) {
  using K = KeyOf<T>;
  using Key = typename K::type;
  extern __shared__ unsigned int whist[];                  // [kWideRows][32]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long g = blockIdx.x / slices;
  const int sl = (int)(blockIdx.x - g * slices);
  const long len = (Tn + slices - 1) / slices;
  const long t0 = sl * len, t1 = (t0 + len < Tn) ? t0 + len : Tn;
  if (t0 >= t1) return;
  for (int i = threadIdx.x; i < kWideRows * kLanes; i += blockDim.x) whist[i] = 0;
  __syncthreads();
  const T* base = mx + g * Tn * kLanes + lane;
  auto bin_of = [](Key key) {
    const long d = (long)(key >> WideOf<T>::shift) - WideOf<T>::base;
    return d < 0 ? kWideBins : (d >= kWideBins ? kWideBins + 1 : (int)d);
  };
  long t = t0 + warp;
  for (; t + 24 < t1; t += 32) {                           // four lines in flight per warp
    const Key k0 = K::of(base[t * kLanes]), k1 = K::of(base[(t + 8) * kLanes]), k2 = K::of(base[(t + 16) * kLanes]),
              k3 = K::of(base[(t + 24) * kLanes]);
    atomicAdd(&whist[bin_of(k0) * kLanes + lane], 1u);
    atomicAdd(&whist[bin_of(k1) * kLanes + lane], 1u);
    atomicAdd(&whist[bin_of(k2) * kLanes + lane], 1u);
    atomicAdd(&whist[bin_of(k3) * kLanes + lane], 1u);
  }
  for (; t < t1; t += 8) atomicAdd(&whist[bin_of(K::of(base[t * kLanes])) * kLanes + lane], 1u);
  __syncthreads();
  unsigned int* gh = gwide + g * kWideRows * kLanes;
  for (int i = threadIdx.x; i < kWideRows * kLanes; i += blockDim.x) {
    const unsigned int c = whist[i];
    if (c) atomicAdd(&gh[i], c);
  }
}

// one CTA per tile: walk the wide histogram to the bin holding rank k; below / above the window -> fallback
template <typename T>
__global__ void __launch_bounds__(256)
setup_wide_pick_kernel(const unsigned int* __restrict__ gwide, unsigned long long* __restrict__ state, long B,
                       int* __restrict__ fallback, int* __restrict__ deep, int* __restrict__ skip_collect) {
  using Key = typename KeyOf<T>::type;
  extern __shared__ unsigned int whist[];
  const long g = blockIdx.x;
  const uint4* src = reinterpret_cast<const uint4*>(gwide + g * kWideRows * kLanes);
  uint4* dst = reinterpret_cast<uint4*>(whist);
  for (int i = threadIdx.x; i < kWideRows * kLanes / 4; i += blockDim.x) dst[i] = src[i];
  __syncthreads();
  if (threadIdx.x >= kLanes) return;
  const int lane = threadIdx.x;
  if (g * kLanes + lane >= B) return;                      // padding lanes of the last tile (all-zero records) do not vote
  unsigned long long* st = state + g * SelState::WORDS * kLanes + lane;
  long k = (long)st[SelState::RANK * kLanes];
  const long under = whist[kWideBins * kLanes + lane];
  if (k < under) { atomicExch(fallback, 1); atomicExch(skip_collect, 1); return; }
  k -= under;
  int d = 0;
  for (; d < kWideBins; ++d) {
    const unsigned c = whist[d * kLanes + lane];
    if (k < (long)c) break;
    k -= c;
  }
  if (d == kWideBins) { atomicExch(fallback, 1); atomicExch(skip_collect, 1); return; }
  // more keys in the bin than candidate slots (long or narrowly distributed tracks): one more full-length digit first
  if (whist[d * kLanes + lane] > (unsigned)kCandCap) { atomicExch(deep, 1); atomicExch(skip_collect, 1); }
  st[SelState::RANK * kLanes] = (unsigned long long)k;
  st[SelState::PREFIX * kLanes] = (unsigned long long)((Key)(WideOf<T>::base + d) << WideOf<T>::shift);
}

// fallback only: bring the select state and the candidate buffer back to their initial values
template <typename T>
__global__ void __launch_bounds__(256)
setup_fallback_reset_kernel(unsigned long long* __restrict__ state, long tiles, long lo, T* __restrict__ cand,
                            const int* __restrict__ fallback, int* __restrict__ overflow) {
  if (*fallback == 0) return;
  if (blockIdx.x == 0 && threadIdx.x == 0) *overflow = 0;
  const long nthr = (long)gridDim.x * blockDim.x, i0 = (long)blockIdx.x * blockDim.x + threadIdx.x;
  for (long i = i0; i < tiles * kLanes; i += nthr) {
    unsigned long long* st = state + (i / kLanes) * SelState::WORDS * kLanes + (i % kLanes);
    st[SelState::PREFIX * kLanes] = 0; st[SelState::RANK * kLanes] = (unsigned long long)lo; st[SelState::CNT_LE * kLanes] = 0;
    st[SelState::NEXT * kLanes] = ~0ull; st[SelState::NCAND * kLanes] = 0; st[SelState::RANK_C * kLanes] = 0;
  }
  using Key = typename KeyOf<T>::type;
  Key* c = reinterpret_cast<Key*>(cand);
  for (long i = i0; i < tiles * (long)kCandCap * kLanes; i += nthr) c[i] = (Key) ~(Key)0;
}

__global__ void setup_init_state_kernel(unsigned long long* __restrict__ state, long lo) {
  unsigned long long* st = state + (long)blockIdx.x * SelState::WORDS * kLanes + threadIdx.x;
  st[SelState::RANK * kLanes] = (unsigned long long)lo;
  st[SelState::NEXT * kLanes] = ~0ull;
}

template <typename T>
__global__ void setup_thr_kernel(const unsigned long long* __restrict__ state, const int* __restrict__ overflow,
                                 int collected, long B, long Tn, double q, double* __restrict__ thr) {
  using K = KeyOf<T>;
  using Key = typename K::type;
  const long b = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const long g = b / kLanes;
  const int lane = (int)(b - g * kLanes);
  const unsigned long long* st = state + g * SelState::WORDS * kLanes + lane;
  long lo;
  double gamma;
  percentile_position(Tn, q, lo, gamma);
  const double a = (double)K::back((Key)st[SelState::PREFIX * kLanes]);
  // with candidates, CNT_LE counts inside the candidate buffer and the rank to compare with is the rank among them
  const long need = (collected && *overflow == 0) ? (long)st[SelState::RANK_C * kLanes] + 2 : lo + 2;
  const bool same = (long)st[SelState::CNT_LE * kLanes] >= need || lo + 1 >= Tn || st[SelState::NEXT * kLanes] == ~0ull;
  const double b2 = same ? a : (double)K::back((Key)st[SelState::NEXT * kLanes]);
  const double diff = __dsub_rn(b2, a);
  double v = __dadd_rn(a, __dmul_rn(diff, gamma));
  if (gamma >= 0.5) v = __dsub_rn(b2, __dmul_rn(diff, __dsub_rn(1.0, gamma)));
  thr[b] = v;
}

// ---- 2. masked moments of y and of dy over the good frames ----------------------------------------
// One warp per (tile, time slice).  Partial record rows (doubles, [tile][slice][row][32 lanes]):
//   0: n_good | 1..P: sum(y-s) | sum (y-s)(y-s)^T packed | n_d | sum dy | sum dy dy^T packed |
//   first good y | last good y          with s = y[frame 0] (shift against cancellation)
template <int P> struct MomRows {
  static constexpr int PS = P * (P + 1) / 2;
  static constexpr int N0 = 0, SY = 1, SYY = SY + P, ND = SYY + PS, SD = ND + 1, SDD = SD + P, FIRST = SDD + PS,
                       LAST = FIRST + P, ROWS = LAST + P;
};

template <typename T, int P>
__global__ void __launch_bounds__(kMomWarpsPerCta* kLanes)
setup_moments_kernel(const T* __restrict__ obs, const T* __restrict__ mx, long B, long Tn, int slices,
                     const double* __restrict__ thr, double* __restrict__ part) {
  using MR = MomRows<P>;
  constexpr int PS = MR::PS;
  const int lane = threadIdx.x & 31;
  const long w = (long)blockIdx.x * kMomWarpsPerCta + (threadIdx.x >> 5);
  const long g = w / slices;
  const int sl = (int)(w - g * slices);
  if (g >= num_tiles_dev(B)) return;
  const long b = g * kLanes + lane;
  const double th = thr[b < B ? b : B - 1];
  const T* base = obs + (g * Tn * (2 * P)) * kLanes + lane;
  const T* mxb = mx + g * Tn * kLanes + lane;          // max_vars from the select passes: 1 row instead of the P variance rows
  const long len = (Tn + slices - 1) / slices;
  const long t0 = sl * len, t1 = (t0 + len < Tn) ? t0 + len : Tn;
  double sh[P], sy[P], syy[PS], sd[P], sdd[PS], first[P], last[P];
  double n0 = 0, nd = 0;
#pragma unroll
  for (int j = 0; j < P; ++j) { sh[j] = (double)base[j * kLanes]; sy[j] = sd[j] = first[j] = last[j] = 0; }
#pragma unroll
  for (int j = 0; j < PS; ++j) syy[j] = sdd[j] = 0;
  bool have = false;
  constexpr int kUn = P <= 4 ? 8 : 4;          // frames whose records are loaded before any is processed (registers)
  for (long tb = t0; tb < t1; tb += kUn) {
    T raw[kUn][P], mxv[kUn];
#pragma unroll
    for (int f = 0; f < kUn; ++f)
      if (tb + f < t1) {
        const T* r = base + (tb + f) * (2 * P) * kLanes;
        mxv[f] = mxb[(tb + f) * kLanes];
#pragma unroll
        for (int j = 0; j < P; ++j) raw[f][j] = r[j * kLanes];
      }
#pragma unroll
    for (int f = 0; f < kUn; ++f) {
      if (tb + f >= t1) break;
      if ((double)mxv[f] <= th) {
        double y[P];
#pragma unroll
        for (int j = 0; j < P; ++j) y[j] = (double)raw[f][j];
        if (have) {
          double d[P];
#pragma unroll
          for (int j = 0; j < P; ++j) { d[j] = y[j] - last[j]; sd[j] += d[j]; }
          int k = 0;
#pragma unroll
          for (int i = 0; i < P; ++i)
#pragma unroll
            for (int j = i; j < P; ++j) sdd[k++] += d[i] * d[j];
          nd += 1;
        } else {
#pragma unroll
          for (int j = 0; j < P; ++j) first[j] = y[j];
          have = true;
        }
        double c[P];
#pragma unroll
        for (int j = 0; j < P; ++j) { last[j] = y[j]; c[j] = y[j] - sh[j]; sy[j] += c[j]; }
        int k = 0;
#pragma unroll
        for (int i = 0; i < P; ++i)
#pragma unroll
          for (int j = i; j < P; ++j) syy[k++] += c[i] * c[j];
        n0 += 1;
      }
    }
  }
  double* o = part + ((g * slices + sl) * MR::ROWS) * kLanes + lane;
  o[MR::N0 * kLanes] = n0;
  o[MR::ND * kLanes] = nd;
#pragma unroll
  for (int j = 0; j < P; ++j) {
    o[(MR::SY + j) * kLanes] = sy[j];
    o[(MR::SD + j) * kLanes] = sd[j];
    o[(MR::FIRST + j) * kLanes] = first[j];
    o[(MR::LAST + j) * kLanes] = last[j];
  }
#pragma unroll
  for (int j = 0; j < PS; ++j) { o[(MR::SYY + j) * kLanes] = syy[j]; o[(MR::SDD + j) * kLanes] = sdd[j]; }
}

// ---- 3. per-track finalize: merge slices, eigen-decomposition, (means, C, S0, Q) ---------------------
// cyclic Jacobi (Rutishauser's update form) on a symmetric P x P matrix held in a flat array; on return
// d[] holds the eigenvalues and column c of v (v[i*P+c]) the c-th eigenvector.  Only the upper triangle of a
// is referenced.
template <int P> __device__ __noinline__ void jacobi_eig(double* a, double* d, double* v) {
  double bb[P], z[P];
  for (int i = 0; i < P; ++i) {
    for (int j = 0; j < P; ++j) v[i * P + j] = (i == j) ? 1.0 : 0.0;
    bb[i] = d[i] = a[i * P + i];
    z[i] = 0.0;
  }
  for (int sweep = 0; sweep < 50; ++sweep) {
    double off = 0.0;
    for (int i = 0; i < P - 1; ++i)
      for (int j = i + 1; j < P; ++j) off += fabs(a[i * P + j]);
    if (off == 0.0) break;
    for (int ip = 0; ip < P - 1; ++ip)
      for (int iq = ip + 1; iq < P; ++iq) {
        const double apq = a[ip * P + iq];
        const double g = 100.0 * fabs(apq);
        if (sweep > 3 && fabs(d[ip]) + g == fabs(d[ip]) && fabs(d[iq]) + g == fabs(d[iq])) {
          a[ip * P + iq] = 0.0;
        } else if (apq != 0.0) {
          const double h = d[iq] - d[ip];
          double t;
          if (fabs(h) + g == fabs(h)) {
            t = apq / h;
          } else {
            const double theta = 0.5 * h / apq;
            t = 1.0 / (fabs(theta) + sqrt(1.0 + theta * theta));
            if (theta < 0.0) t = -t;
          }
          const double c = 1.0 / sqrt(1.0 + t * t), sn = t * c, tau = sn / (1.0 + c), hh = t * apq;
          z[ip] -= hh; z[iq] += hh; d[ip] -= hh; d[iq] += hh;
          a[ip * P + iq] = 0.0;
          for (int j = 0; j < ip; ++j) {                       // rotate the upper triangle
            const double x = a[j * P + ip], y = a[j * P + iq];
            a[j * P + ip] = x - sn * (y + x * tau);
            a[j * P + iq] = y + sn * (x - y * tau);
          }
          for (int j = ip + 1; j < iq; ++j) {
            const double x = a[ip * P + j], y = a[j * P + iq];
            a[ip * P + j] = x - sn * (y + x * tau);
            a[j * P + iq] = y + sn * (x - y * tau);
          }
          for (int j = iq + 1; j < P; ++j) {
            const double x = a[ip * P + j], y = a[iq * P + j];
            a[ip * P + j] = x - sn * (y + x * tau);
            a[iq * P + j] = y + sn * (x - y * tau);
          }
          for (int j = 0; j < P; ++j) {
            const double x = v[j * P + ip], y = v[j * P + iq];
            v[j * P + ip] = x - sn * (y + x * tau);
            v[j * P + iq] = y + sn * (x - y * tau);
          }
        }
      }
    for (int i = 0; i < P; ++i) { bb[i] += z[i]; d[i] = bb[i]; z[i] = 0.0; }
  }
}

// Few tiles and long tracks leave thousands of moment slices per tile (one warp each, to fill the GPU); the finalize
// kernel walks them in frame order, one thread per track (6 ms at 64 tracks x 2M frames).  One warp per (tile, group)
// folds a run of consecutive slices into one partial of the same layout first; the finalize then walks kMomGroups.
template <int P>
__global__ void __launch_bounds__(256)
setup_reduce_kernel(const double* __restrict__ part, long tiles, int slices, int groups, double* __restrict__ part2) {
  using MR = MomRows<P>;
  constexpr int PS = MR::PS;
  const int lane = threadIdx.x & 31;
  const long w = (long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const long g = w / groups;
  const int grp = (int)(w - g * groups);
  if (g >= tiles) return;
  const int per = (slices + groups - 1) / groups, s0 = grp * per, s1 = (s0 + per < slices) ? s0 + per : slices;
  double n0 = 0, nd = 0, sy[P], syy[PS], sd[P], sdd[PS], first[P], last[P];
  for (int j = 0; j < P; ++j) sy[j] = sd[j] = first[j] = last[j] = 0;
  for (int j = 0; j < PS; ++j) syy[j] = sdd[j] = 0;
  bool have = false;
  for (int sl = s0; sl < s1; ++sl) {
    const double* r = part + ((g * slices + sl) * MR::ROWS) * kLanes + lane;
    const double n_s = r[MR::N0 * kLanes];
    if (n_s == 0) continue;
    if (have) {                                 // stitch, as in the finalize kernel
      double d[P];
      for (int j = 0; j < P; ++j) { d[j] = r[(MR::FIRST + j) * kLanes] - last[j]; sd[j] += d[j]; }
      int k = 0;
      for (int i = 0; i < P; ++i)
        for (int j = i; j < P; ++j) sdd[k++] += d[i] * d[j];
      nd += 1;
    } else {
      for (int j = 0; j < P; ++j) first[j] = r[(MR::FIRST + j) * kLanes];
    }
    have = true;
    n0 += n_s;
    nd += r[MR::ND * kLanes];
    for (int j = 0; j < P; ++j) {
      sy[j] += r[(MR::SY + j) * kLanes];
      sd[j] += r[(MR::SD + j) * kLanes];
      last[j] = r[(MR::LAST + j) * kLanes];
    }
    for (int j = 0; j < PS; ++j) { syy[j] += r[(MR::SYY + j) * kLanes]; sdd[j] += r[(MR::SDD + j) * kLanes]; }
  }
  double* o = part2 + ((g * groups + grp) * MR::ROWS) * kLanes + lane;
  o[MR::N0 * kLanes] = n0;
  o[MR::ND * kLanes] = nd;
  for (int j = 0; j < P; ++j) {
    o[(MR::SY + j) * kLanes] = sy[j]; o[(MR::SD + j) * kLanes] = sd[j];
    o[(MR::FIRST + j) * kLanes] = first[j]; o[(MR::LAST + j) * kLanes] = last[j];
  }
  for (int j = 0; j < PS; ++j) { o[(MR::SYY + j) * kLanes] = syy[j]; o[(MR::SDD + j) * kLanes] = sdd[j]; }
}

template <typename T, int P, int N>
__global__ void setup_finalize_kernel(const T* __restrict__ obs, const double* __restrict__ part, long B, long Tn,
                                      int slices, double smooth_param, T* __restrict__ means, T* __restrict__ Cout,
                                      T* __restrict__ S0, T* __restrict__ Q, int* __restrict__ status) {
  using MR = MomRows<P>;
  constexpr int PS = MR::PS;
  const long b = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const long g = b / kLanes;
  const int lane = (int)(b - g * kLanes);
  double n0 = 0, nd = 0, sy[P], syy[PS], sd[P], sdd[PS], last[P];
  for (int j = 0; j < P; ++j) sy[j] = sd[j] = last[j] = 0;
  for (int j = 0; j < PS; ++j) syy[j] = sdd[j] = 0;
  bool have = false;
  for (int sl = 0; sl < slices; ++sl) {
    const double* r = part + ((g * slices + sl) * MR::ROWS) * kLanes + lane;
    const double n_s = r[MR::N0 * kLanes];
    if (n_s == 0) continue;
    if (have) {                                 // stitch: first good frame of this slice follows `last`
      double d[P];
      for (int j = 0; j < P; ++j) { d[j] = r[(MR::FIRST + j) * kLanes] - last[j]; sd[j] += d[j]; }
      int k = 0;
      for (int i = 0; i < P; ++i)
        for (int j = i; j < P; ++j) sdd[k++] += d[i] * d[j];
      nd += 1;
    }
    have = true;
    n0 += n_s;
    nd += r[MR::ND * kLanes];
    for (int j = 0; j < P; ++j) {
      sy[j] += r[(MR::SY + j) * kLanes];
      sd[j] += r[(MR::SD + j) * kLanes];
      last[j] = r[(MR::LAST + j) * kLanes];
    }
    for (int j = 0; j < PS; ++j) { syy[j] += r[(MR::SYY + j) * kLanes]; sdd[j] += r[(MR::SDD + j) * kLanes]; }
  }
  bool ok = n0 >= 2 && nd >= 2;
  const T* base = obs + (g * Tn * (2 * P)) * kLanes + lane;
  double cov0[P][P], covd[P][P], a[P * P], v[P * P], ev[P];
  {
    int k = 0;
    for (int i = 0; i < P; ++i)
      for (int j = i; j < P; ++j, ++k) {
        const double c0 = (syy[k] - sy[i] * sy[j] / n0) / n0;              // ddof 0 (np.var, :722)
        const double cd = (sdd[k] - sd[i] * sd[j] / nd) / (nd - 1.0);      // ddof 1 (np.cov, :728)
        cov0[i][j] = cov0[j][i] = c0;
        covd[i][j] = covd[j][i] = cd;
        a[i * P + j] = a[j * P + i] = c0;
      }
  }
  for (int j = 0; j < P; ++j) means[b * P + j] = (T)((double)base[j * kLanes] + sy[j] / n0);   // (:698-700)
  jacobi_eig<P>(a, ev, v);
  int order[P];
  for (int i = 0; i < P; ++i) order[i] = i;
  for (int i = 0; i < P - 1; ++i)              // eigenvalues descending (sklearn PCA order)
    for (int j = i + 1; j < P; ++j)
      if (ev[order[j]] > ev[order[i]]) { const int tmp = order[i]; order[i] = order[j]; order[j] = tmp; }
  double W[N][P];
  for (int c = 0; c < N; ++c) {
    const int col = order[c];
    int jm = 0;                                 // sklearn svd_flip(u_based_decision=False): largest |entry| positive
    for (int j = 1; j < P; ++j)
      if (fabs(v[j * P + col]) > fabs(v[jm * P + col])) jm = j;
    const double sgn = v[jm * P + col] < 0 ? -1.0 : 1.0;
    for (int j = 0; j < P; ++j) W[c][j] = sgn * v[j * P + col];
  }
  for (int j = 0; j < P; ++j)
    for (int c = 0; c < N; ++c) Cout[(b * P + j) * N + c] = (T)W[c][j];     // C = W^T (:730)
  for (int c = 0; c < N; ++c)
    for (int e = 0; e < N; ++e) {
      double s0 = 0, q = 0;
      for (int i = 0; i < P; ++i)
        for (int j = 0; j < P; ++j) {
          s0 += W[c][i] * cov0[i][j] * W[e][j];
          q += W[c][i] * covd[i][j] * W[e][j];
        }
      S0[(b * N + c) * N + e] = (T)((c == e) ? s0 : 0.0);                  // np.diag(np.var(..)) (:722)
      Q[(b * N + c) * N + e] = (T)(smooth_param * q);
    }
  if (status && !ok) status[b] = 1;
}

// ---- 4. centre the observations in place: y[t,j] -= means[b,j]  (:702-705) ---------------------------
// One warp per (tile, block of kCenterFrames frames): all R x kCenterFrames lines are loaded before the
// first store so that enough bytes are in flight per SM to saturate HBM.
constexpr int kCenterFrames = 8;
template <typename T, int R>
__global__ void __launch_bounds__(256)
center_rows_kernel(T* __restrict__ rec, const T* __restrict__ means, long B, long Tn, int rows) {
  const long nblk = (Tn + kCenterFrames - 1) / kCenterFrames;
  const long w = (long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (w >= num_tiles_dev(B) * nblk) return;
  const long g = w / nblk;
  const long t0 = (w - g * nblk) * kCenterFrames;
  const long b = g * kLanes + lane;
  if (b >= B) return;
  T mu[R], v[kCenterFrames][R];
#pragma unroll
  for (int r = 0; r < R; ++r) mu[r] = means[b * R + r];
  T* base = rec + ((g * Tn + t0) * rows) * kLanes + lane;
  const int nf = (int)(Tn - t0 < kCenterFrames ? Tn - t0 : kCenterFrames);
#pragma unroll
  for (int f = 0; f < kCenterFrames; ++f)
    if (f < nf) {
#pragma unroll
      for (int r = 0; r < R; ++r) v[f][r] = base[((long)f * rows + r) * kLanes];
    }
#pragma unroll
  for (int f = 0; f < kCenterFrames; ++f)
    if (f < nf) {
#pragma unroll
      for (int r = 0; r < R; ++r) base[((long)f * rows + r) * kLanes] = v[f][r] - mu[r];
    }
}

template <typename T, int P>
static int setup_run(T* obs, long B, long Tn, int n, double q, double s, int center, T* means, T* C, T* S0, T* Q,
                     double* ws, int* status, cudaStream_t st, int mx_ready) {
  using Key = typename KeyOf<T>::type;
  static_assert(MomRows<P>::ROWS == mom_rows<P>(), "workspace row count");
  if (n != 3) { set_error("pca_setup: n=%d (the reference fits 3 components)", n); return EKSB_EDIM; }
  const long tiles = num_tiles(B);
  const SetupWs L = setup_layout(B, Tn, P, (int)sizeof(T));
  double* thr = ws + L.thr;
  unsigned long long* state = reinterpret_cast<unsigned long long*>(ws + L.state);
  unsigned int* ghist = reinterpret_cast<unsigned int*>(ws + L.ghist);
  double* part = ws + L.part;
  T* mx = reinterpret_cast<T*>(ws + L.mx);
  T* cand = reinterpret_cast<T*>(ws + L.cand);
  int* overflow = reinterpret_cast<int*>(ws + L.flag);
  const int npass = (int)sizeof(Key), ssl = select_slices(tiles, Tn), msl = moment_slices(tiles, Tn);
  unsigned int* ghist2 = reinterpret_cast<unsigned int*>(ws + L.ghist2);
  unsigned int* gwide = reinterpret_cast<unsigned int*>(ws + L.gwide);
  int* fallback = overflow + 1;
  // full-length passes before the candidates are collected (EKSB_SETUP_NO_COLLECT=1: never, every pass full length)
  const char* nc = getenv("EKSB_SETUP_NO_COLLECT");
  const int collect_after = (nc && nc[0] == '1') ? npass : (sizeof(T) == 8 ? 3 : 2);
  const char* nw = getenv("EKSB_SETUP_NO_WIDE");          // A/B switch: 1 = 8-bit passes only (round 1)
  // state: prefix 0, rank = lo, cnt_le 0, next = ~0 ; histograms 0
  long lo;
  double gamma;
  percentile_position(Tn, q, lo, gamma);
  const bool collect = collect_after < npass;
  const bool wide = collect && !(nw && nw[0] == '1');
  cudaMemsetAsync(ghist, 0, (size_t)npass * tiles * 256 * kLanes * sizeof(unsigned int), st);
  if (wide) {
    cudaMemsetAsync(ghist2, 0, (size_t)npass * tiles * 256 * kLanes * sizeof(unsigned int), st);
    cudaMemsetAsync(gwide, 0, (size_t)tiles * kWideRows * kLanes * sizeof(unsigned int), st);
  }
  cudaMemsetAsync(state, 0, (size_t)tiles * SelState::WORDS * kLanes * 8, st);
  cudaMemsetAsync(overflow, 0, 16, st);                                                          // overflow, fallback, deepen
  if (collect) cudaMemsetAsync(cand, 0xFF, (size_t)tiles * kCandCap * kLanes * sizeof(T), st);   // all-ones key = empty slot
  setup_init_state_kernel<<<(unsigned)tiles, kLanes, 0, st>>>(state, lo);
  if (!mx_ready)     // else the ensemble kernel wrote max_vars into the workspace already (eksb_ensemble_tiled_mx)
    setup_mx_kernel<T, P><<<(unsigned)((tiles * Tn + 7) / 8), 256, 0, st>>>(obs, tiles, Tn, mx);
  const int* gate = nullptr;
  if (wide) {
    // one full-length pass with a wide first digit, the candidates of its bin, the remaining bits on the candidates
    constexpr int kWideSmem = kWideRows * kLanes * (int)sizeof(unsigned int);
    cudaFuncSetAttribute(setup_wide_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, kWideSmem);
    cudaFuncSetAttribute(setup_wide_pick_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, kWideSmem);
    setup_wide_kernel<T><<<(unsigned)(tiles * ssl), 256, kWideSmem, st>>>(mx, Tn, ssl, gwide);
    int* deep = overflow + 2;
    int* skip_collect = overflow + 3;                       // fallback or deepen
    setup_wide_pick_kernel<T><<<(unsigned)tiles, 256, kWideSmem, st>>>(gwide, state, B, fallback, deep, skip_collect);
    setup_collect_kernel<T><<<(unsigned)(tiles * ssl), 256, 0, st>>>(mx, Tn, ssl, WideOf<T>::shift, state, cand, overflow, skip_collect, 0);
    // a bin with more keys than candidate slots (the pick kernel knows the count): one more full-length 8-bit digit
    // below the wide one, then collect -- four full-length reads, what round 1 always paid
    constexpr int dshift = WideOf<T>::shift - 8;
    {
      unsigned int* gh = ghist2 + (size_t)(npass - 1) * tiles * 256 * kLanes;
      setup_hist_kernel<T><<<(unsigned)(tiles * ssl), 256, 0, st>>>(mx, Tn, nullptr, overflow, ssl, dshift, state, gh, WideOf<T>::shift, deep, 1);
      setup_pick_kernel<T><<<(unsigned)tiles, 256, 0, st>>>(gh, dshift, state, deep, 1);
      setup_collect_kernel<T><<<(unsigned)(tiles * ssl), 256, 0, st>>>(mx, Tn, ssl, dshift, state, cand, overflow, deep, 1);
    }
    int pass = 0;
    for (int shift = ((WideOf<T>::shift - 1) / 8) * 8; shift >= 0; shift -= 8, ++pass) {
      unsigned int* gh = ghist2 + (size_t)pass * tiles * 256 * kLanes;
      const int match = shift + 8 < WideOf<T>::shift ? shift + 8 : WideOf<T>::shift;
      setup_hist_kernel<T><<<(unsigned)(tiles * ssl), 256, 0, st>>>(mx, Tn, cand, overflow, ssl, shift, state, gh, match, fallback, 0, deep, dshift);
      setup_pick_kernel<T><<<(unsigned)tiles, 256, 0, st>>>(gh, shift, state, fallback, 0);
    }
    setup_fallback_reset_kernel<T><<<(unsigned)(4 * sm_count()), 256, 0, st>>>(state, tiles, lo, cand, fallback, overflow);
    gate = fallback;                                        // the 8-bit passes below run only if the flag is up
  }
  for (int pass = 0; pass < npass; ++pass) {
    const int shift = 8 * (npass - 1 - pass);
    unsigned int* gh = ghist + (size_t)pass * tiles * 256 * kLanes;
    if (pass == collect_after)
      setup_collect_kernel<T><<<(unsigned)(tiles * ssl), 256, 0, st>>>(mx, Tn, ssl, shift + 8, state, cand, overflow, gate, 1);
    setup_hist_kernel<T><<<(unsigned)(tiles * ssl), 256, 0, st>>>(mx, Tn, pass >= collect_after ? cand : nullptr, overflow,
                                                                  ssl, shift, state, gh, shift + 8, gate, 1);
    setup_pick_kernel<T><<<(unsigned)tiles, 256, 0, st>>>(gh, shift, state, gate, 1);
  }
  setup_tail_kernel<T><<<(unsigned)(tiles * ssl), 256, 0, st>>>(mx, Tn, collect ? cand : nullptr, overflow, ssl, state);
  setup_thr_kernel<T><<<(unsigned)((B + 127) / 128), 128, 0, st>>>(state, overflow, collect ? 1 : 0, B, Tn, q, thr);
  const long warps = tiles * msl;
  setup_moments_kernel<T, P><<<(unsigned)((warps + kMomWarpsPerCta - 1) / kMomWarpsPerCta), kMomWarpsPerCta * kLanes, 0, st>>>(
      obs, mx, B, Tn, msl, thr, part);
  const double* fpart = part;
  int fsl = msl;
  if (msl > 2 * kMomGroups) {                   // few tiles, long tracks: fold the slices first (see setup_reduce_kernel)
    double* part2 = ws + L.part2;
    setup_reduce_kernel<P><<<(unsigned)((tiles * kMomGroups + 7) / 8), 256, 0, st>>>(part, tiles, msl, kMomGroups, part2);
    fpart = part2;
    fsl = kMomGroups;
  }
  setup_finalize_kernel<T, P, 3><<<(unsigned)((B + 63) / 64), 64, 0, st>>>(obs, fpart, B, Tn, fsl, s, means, C, S0, Q, status);
  if (center) {
    const long warps_c = tiles * ((Tn + kCenterFrames - 1) / kCenterFrames);
    center_rows_kernel<T, P><<<(unsigned)((warps_c + 7) / 8), 256, 0, st>>>(obs, means, B, Tn, 2 * P);
  }
  return check_launch("pca_setup kernels");
}

template <typename T>
static int setup_api(T* obs, long B, long Tn, int n, int p, double q, double s, int center, T* means, T* C, T* S0, T* Q,
                     double* ws, size_t ws_bytes, int* status, void* stream, int mx_ready = 0) {
  if (!obs || !means || !C || !S0 || !Q || !ws || B <= 0 || Tn <= 1) { set_error("pca_setup: null pointer or empty batch"); return EKSB_EARG; }
  if (!(q >= 0.0 && q <= 100.0)) { set_error("pca_setup: quantile %g outside [0,100]", q); return EKSB_EARG; }
  if (ws_bytes < eksb_pca_setup_workspace_bytes(B, Tn, p, (int)sizeof(T))) { set_error("pca_setup: workspace too small"); return EKSB_EARG; }
  cudaStream_t st = (cudaStream_t)stream;
  switch (p) {
    case 4: return setup_run<T, 4>(obs, B, Tn, n, q, s, center, means, C, S0, Q, ws, status, st, mx_ready);
    case 6: return setup_run<T, 6>(obs, B, Tn, n, q, s, center, means, C, S0, Q, ws, status, st, mx_ready);
    case 8: return setup_run<T, 8>(obs, B, Tn, n, q, s, center, means, C, S0, Q, ws, status, st, mx_ready);
  }
  set_error("pca_setup: p=%d not instantiated (4, 6, 8)", p);
  return EKSB_EDIM;
}

}  // namespace eksb

using namespace eksb;

extern "C" size_t eksb_pca_setup_workspace_bytes(long B, long T, int p, int elem_size) {
  if (B <= 0 || T <= 0 || p <= 0 || (elem_size != 4 && elem_size != 8)) return 0;
  return setup_layout(B, T, p, elem_size).total * 8;
}
extern "C" int eksb_pca_setup_f64(double* obs, long B, long T, int n, int p, double quantile, double smooth_param,
                                  int center, double* means, double* C, double* S0, double* Q, double* workspace,
                                  size_t workspace_bytes, int* status, void* stream) {
  return setup_api<double>(obs, B, T, n, p, quantile, smooth_param, center, means, C, S0, Q, workspace, workspace_bytes, status, stream);
}
extern "C" int eksb_pca_setup_f32(float* obs, long B, long T, int n, int p, double quantile, double smooth_param,
                                  int center, float* means, float* C, float* S0, float* Q, double* workspace,
                                  size_t workspace_bytes, int* status, void* stream) {
  return setup_api<float>(obs, B, T, n, p, quantile, smooth_param, center, means, C, S0, Q, workspace, workspace_bytes, status, stream);
}

/* max_vars produced by eksb_ensemble_tiled_mx: the buffer lives inside the setup workspace at this byte offset
 * ([tile][t][32] scalars of the obs type); pass mx_ready = 1 and the setup skips its own first pass. */
extern "C" size_t eksb_pca_setup_mx_offset_bytes(long B, long T, int p, int elem_size) {
  if (B <= 0 || T <= 0 || p <= 0 || (elem_size != 4 && elem_size != 8)) return 0;
  return setup_layout(B, T, p, elem_size).mx * 8;
}
extern "C" int eksb_pca_setup_mx_f64(double* obs, long B, long T, int n, int p, double quantile, double smooth_param,
                                     int center, int mx_ready, double* means, double* C, double* S0, double* Q,
                                     double* workspace, size_t workspace_bytes, int* status, void* stream) {
  return setup_api<double>(obs, B, T, n, p, quantile, smooth_param, center, means, C, S0, Q, workspace, workspace_bytes, status, stream, mx_ready);
}
extern "C" int eksb_pca_setup_mx_f32(float* obs, long B, long T, int n, int p, double quantile, double smooth_param,
                                     int center, int mx_ready, float* means, float* C, float* S0, float* Q,
                                     double* workspace, size_t workspace_bytes, int* status, void* stream) {
  return setup_api<float>(obs, B, T, n, p, quantile, smooth_param, center, means, C, S0, Q, workspace, workspace_bytes, status, stream, mx_ready);
}

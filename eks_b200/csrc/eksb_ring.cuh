// eksb_ring.cuh -- per-lane shared-memory frame rings for the thread-per-track recursions (device only).
#pragma once
#include "eksb_common.cuh"

namespace eksb {

// ---- per-lane shared-memory frame ring (device only) ----------------------------------------------
// A register prefetch ring cannot keep more than ~2 frames of loads in flight per warp (scoreboard
// slots, see eksb_pipe.cuh), and with 12 resident warps per SM that is too few bytes in flight to
// saturate HBM.  Here every lane copies ITS OWN column of the next kRingDepth frames global -> shared
// with cp.async (LDGSTS: no register, no scoreboard), and reads it back after cp.async.wait_group:
// a lane only ever consumes what it copied itself, so no barrier of any kind is needed.
// Ring layout per warp: [slot][row][32 lanes]; `ring` points at this lane's column.
constexpr int kRingDepth = 6;

template <typename T> __device__ __forceinline__ void cp_async_elem(T* dst_smem, const T* src) {
  static_assert(sizeof(T) == 4 || sizeof(T) == 8, "element size");
  const unsigned d = (unsigned)__cvta_generic_to_shared(dst_smem);
  if (sizeof(T) == 8) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(src) : "memory");
  else                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int NPENDING> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(NPENDING) : "memory");
}
template <typename T, int ROWS> __device__ __forceinline__ void ring_issue(T* ring, int slot, const T* frame) {
#pragma unroll
  for (int r = 0; r < ROWS; ++r) cp_async_elem(ring + (slot * ROWS + r) * kLanes, frame + r * kLanes);
}
template <typename T, int ROWS> __host__ __device__ constexpr size_t ring_bytes_per_warp() { return (size_t)kRingDepth * ROWS * kLanes * sizeof(T); }

// kf_forward_range (eksb_math.cuh) with the obs frames staged through the ring; same arithmetic.
template <typename T, int N, int P, int AM, bool NLL = false>
__device__ bool kf_forward_range_ring(const Params<T, N, P, AM>& prm, const T* __restrict__ obs, T* __restrict__ filt,
                                      long n, int first, T* m, T* V, T* ring, double* nll = nullptr,
                                      const T* mu = nullptr /* P per-track observation offsets in registers, or null */,
                                      bool* nonpos = nullptr /* set when a frame has a non-positive ensemble variance */) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, R2 = 2 * P, D = kRingDepth;
  bool ok = true;
#pragma unroll
  for (int k = 0; k < D; ++k) {
    if (k < n) ring_issue<T, R2>(ring, k, obs + (long)k * R2 * kLanes);
    cp_async_commit();
  }
  int slot = 0;
  for (long t = 0; t < n; ++t) {
    cp_async_wait<D - 1>();
    T cur[R2];
#pragma unroll
    for (int r = 0; r < R2; ++r) cur[r] = ring[(slot * R2 + r) * kLanes];
    if (t + D < n) ring_issue<T, R2>(ring, slot, obs + (t + D) * R2 * kLanes);
    cp_async_commit();
    slot = (slot + 1 == D) ? 0 : slot + 1;
    if (mu) {
#pragma unroll
      for (int j = 0; j < P; ++j) cur[j] -= mu[j];
    }
    const bool is_first = (t == 0);
    if (nonpos && !(is_first && first == kPriorOnly)) *nonpos = *nonpos || any_nonpos<T, P>(cur + P);
    if (!(is_first && first == kPriorOnly)) {
      T mp[N], S[NS];
      if (is_first && first == kPriorUpdate) {
#pragma unroll
        for (int i = 0; i < N; ++i) mp[i] = m[i];
#pragma unroll
        for (int i = 0; i < NS; ++i) S[i] = V[i];
      } else {
        predict<T, N, P, AM>(prm, m, V, mp, S);
      }
      const double e = update<T, N, P, AM, NLL>(prm, cur, cur + P, mp, S, m, V, ok);
      if (NLL) *nll += e;
    }
    if (filt) {
      T* o = filt + t * SG * kLanes;
#pragma unroll
      for (int i = 0; i < N; ++i) o[i * kLanes] = m[i];
#pragma unroll
      for (int i = 0; i < NS; ++i) o[(N + i) * kLanes] = V[i];
    }
  }
  cp_async_wait<0>();
  return ok;
}

// rts_backward_range (eksb_math.cuh; no yv / cv outputs) with the filtered records staged through the ring.
// smth may alias filt: frame t is copied to shared memory before the loop reaches (and overwrites) it.
template <typename T, int N, int P, int AM>
__device__ bool rts_backward_range_ring(const Params<T, N, P, AM>& prm, const T* filt, T* smth, T* ym, long n,
                                        bool terminal, T* ms, T* Vs, T* ring) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, D = kRingDepth;
  bool ok = true;
#pragma unroll
  for (int k = 0; k < D; ++k) {
    if (n - 1 - k >= 0) ring_issue<T, SG>(ring, k, filt + (n - 1 - k) * SG * kLanes);
    cp_async_commit();
  }
  int slot = 0;
  for (long t = n - 1; t >= 0; --t) {
    cp_async_wait<D - 1>();
    T cur[SG];
#pragma unroll
    for (int r = 0; r < SG; ++r) cur[r] = ring[(slot * SG + r) * kLanes];
    if (t - D >= 0) ring_issue<T, SG>(ring, slot, filt + (t - D) * SG * kLanes);
    cp_async_commit();
    slot = (slot + 1 == D) ? 0 : slot + 1;
    if (terminal && t == n - 1) {
#pragma unroll
      for (int i = 0; i < N; ++i) ms[i] = cur[i];
#pragma unroll
      for (int i = 0; i < NS; ++i) Vs[i] = cur[N + i];
    } else {
      T ms1[N], Vs1[NS];
#pragma unroll
      for (int i = 0; i < N; ++i) ms1[i] = ms[i];
#pragma unroll
      for (int i = 0; i < NS; ++i) Vs1[i] = Vs[i];
      rts_step<T, N, P, AM>(prm, cur, cur + N, ms1, Vs1, ms, Vs, nullptr, ok);
    }
    T* o = smth + t * SG * kLanes;
#pragma unroll
    for (int i = 0; i < N; ++i) o[i * kLanes] = ms[i];
#pragma unroll
    for (int i = 0; i < NS; ++i) o[(N + i) * kLanes] = Vs[i];
    if (ym) {
#pragma unroll
      for (int j = 0; j < P; ++j) {
        T sacc = T(0);
#pragma unroll
        for (int i = 0; i < N; ++i) sacc += prm.C[j][i] * ms[i];
        ym[(t * P + j) * kLanes] = sacc;
      }
    }
  }
  cp_async_wait<0>();
  return ok;
}


}  // namespace eksb

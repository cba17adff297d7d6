// eksb_kf.cuh -- forward Kalman filter (A2/A3, A6) and RTS smoother (A4/A5) kernels.
//
// Mapping: one track per thread, one track-tile (32 tracks) per warp, one warp per CTA.
// Every global access of a warp is one contiguous 128 B (f32) / 256 B (f64) line of the
// track-tile layout (eksb_math.cuh), so the kernels stream HBM at full sector efficiency
// The forward kernel stages its obs frames through a per-lane cp.async ring in shared memory
// (eksb_ring.cuh: no barrier, no register or scoreboard tied up by the prefetch); the backward kernel
// (CV / yv outputs, single-track API calls) keeps a PF-frame register prefetch ring.
// All matrices live in registers (see -Xptxas -v: no local memory).
#pragma once
#include <stdlib.h>
#include "eksb_common.cuh"
#include "eksb_ring.cuh"

namespace eksb {

// Frames in flight per thread: as many as fit in ~64 32-bit registers of prefetch ring.
template <typename T, int ROWS> struct Prefetch {
  static constexpr int kRegsPerFrame = ROWS * (int)(sizeof(T) / 4);
  static constexpr int value = (64 / kRegsPerFrame) < 1 ? 1 : (64 / kRegsPerFrame);
};

// grid = (n_tiles, n_scale), block = 32
template <typename T, int N, int P, int AM, bool NLL>
__global__ void __launch_bounds__(kLanes)
kf_forward_kernel(const T* __restrict__ obs, const T* __restrict__ m0, const T* __restrict__ S0,
                  const T* __restrict__ C, const T* __restrict__ A, const T* __restrict__ Q, int shared_mask,
                  long B, long Tn, int first, const T* __restrict__ q_scale, T* __restrict__ filt,
                  double* __restrict__ nll, int* __restrict__ status, const T* __restrict__ yoff) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA;
  extern __shared__ __align__(16) unsigned char kf_smem[];
  const long g = blockIdx.x;
  const int lane = threadIdx.x;
  const long b = g * kLanes + lane;
  const long bp = b < B ? b : B - 1;   // padding lanes replay the last track's parameters
  Params<T, N, P, AM> prm;
  load_track_params<T, N, P, AM>(prm, C, A, Q, shared_mask, bp);
  if (q_scale) {
    const T s = q_scale[blockIdx.y];
#pragma unroll
    for (int i = 0; i < NS; ++i) prm.Q[i] *= s;
  }
  T m[N], V[NS];
  {
    const T* pm = m0 + ((shared_mask & EKSB_SHARED_M0) ? 0 : bp * N);
    const T* ps = S0 + ((shared_mask & EKSB_SHARED_S0) ? 0 : bp * (N * N));
#pragma unroll
    for (int i = 0; i < N; ++i) m[i] = pm[i];
    load_sym<T, N>(V, ps);
  }
  bool ok = true;
  if (first == kPriorOnly) {   // eks/newton_eks.py:132 : P starts at inv(S0)
    T Vi[NS];
    sym_inv<T, N>(V, Vi, ok);
#pragma unroll
    for (int i = 0; i < NS; ++i) V[i] = Vi[i];
  }
  const T* o = obs + (g * Tn * (2 * P)) * kLanes + lane;
  T* f = filt ? filt + (g * Tn * SG) * kLanes + lane : nullptr;
  double acc = 0.0;
  T mu[P];                                 // optional per-track observation offset (camera means): y - mu, like the pipelined kernels
#pragma unroll
  for (int j = 0; j < P; ++j) mu[j] = yoff ? yoff[bp * P + j] : T(0);
  bool nonpos = false;
  ok = kf_forward_range_ring<T, N, P, AM, NLL>(prm, o, f, Tn, first, m, V, reinterpret_cast<T*>(kf_smem) + lane, &acc,
                                               yoff ? mu : nullptr, &nonpos) && ok;
  if (b < B) {
    if (NLL) nll[(long)blockIdx.y * B + b] = 0.5 * (acc + (double)Tn * P * 1.8378770664093453 /* log 2pi */);
    if (status) set_status(status + b, ok, nonpos);
  }
}

// grid = n_tiles, block = 32
template <typename T, int N, int P, int AM>
__global__ void __launch_bounds__(kLanes)
rts_backward_kernel(const T* filt, const T* __restrict__ C, const T* __restrict__ A, const T* __restrict__ Q,
                    int shared_mask, long B, long Tn, T* smth, T* __restrict__ ym, T* __restrict__ yv,
                    T* __restrict__ cv, int* __restrict__ status) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA;
  const long g = blockIdx.x;
  const int lane = threadIdx.x;
  const long b = g * kLanes + lane;
  const long bp = b < B ? b : B - 1;
  Params<T, N, P, AM> prm;
  load_track_params<T, N, P, AM>(prm, C, A, Q, shared_mask, bp);
  const long base = g * Tn * kLanes;
  T ms[N], Vs[NS];
  bool ok = rts_backward_range<T, N, P, AM, kLanes, Prefetch<T, Dim<N>::SIGMA>::value>(
      prm, filt + base * SG + lane, smth + base * SG + lane, ym ? ym + base * P + lane : nullptr,
      yv ? yv + base * P + lane : nullptr, cv ? cv + base * (N * N) + lane : nullptr, Tn, true, ms, Vs);
  if (b < B && status) set_status(status + b, ok, false);
}

inline bool use_nll_pipe() {
  const char* e = getenv("EKSB_NLL_PIPE");                 // A/B switch, read per call: 0 = plain thread-per-track NLL kernel
  return !(e && e[0] == '0');
}

template <typename T, int N, int P>
static int launch_forward(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int shared_mask,
                          long B, long Tn, int a_mode, int first, const T* q_scale, int n_scale, T* filt, double* nll,
                          int* status, const T* yoff, cudaStream_t st) {
  if (filt && !nll && !q_scale && use_pipe(B)) {           // warp-specialised kernel, in groups of tiles it handles best
    const long tiles = num_tiles(B), ng = pipe_groups(tiles, (int)sizeof(T)), per = (tiles + ng - 1) / ng;
    int rc = EKSB_OK;
    for (long g0 = 0; g0 < tiles && rc == EKSB_OK; g0 += per) {
      const long b0 = g0 * kLanes, Bs = (B - b0 < per * kLanes) ? B - b0 : per * kLanes;
      auto off = [&](const T* ptr, int flag, long stride) { return ptr + ((shared_mask & flag) ? 0 : b0 * stride); };
      rc = kf_forward_pipe<T>(obs + g0 * Tn * (2 * P) * kLanes, off(m0, EKSB_SHARED_M0, N), off(S0, EKSB_SHARED_S0, N * N),
                              off(C, EKSB_SHARED_C, P * N), off(A, EKSB_SHARED_A, N * N), off(Q, EKSB_SHARED_Q, N * N),
                              shared_mask, Bs, Tn, N, P, a_mode, first, filt + g0 * Tn * Dim<N>::SIGMA * kLanes,
                              status ? status + b0 : nullptr, yoff ? yoff + b0 * P : nullptr, st);
    }
    if (rc != EKSB_EDIM) return rc;
  }
  if (nll && !filt && a_mode == EKSB_A_DIAG && use_pipe(B) && use_nll_pipe()) {   // NLL / grid search: pipelined form
    const int rc = kf_nll_pipe<T>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, N, P, first, q_scale, n_scale, nll, status, yoff, st);
    if (rc != EKSB_EDIM) return rc;
  }
  dim3 grid((unsigned)num_tiles(B), (unsigned)(q_scale ? n_scale : 1));
#define EKSB_GO(AM_, NLL_)                                                                                      \
  kf_forward_kernel<T, N, P, AM_, NLL_><<<grid, kLanes, ring_bytes_per_warp<T, 2 * P>(), st>>>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, first, \
                                                                 q_scale, filt, nll, status, yoff)
  if (a_mode == EKSB_A_DIAG) { if (nll) EKSB_GO(0, true); else EKSB_GO(0, false); }
  else                       { if (nll) EKSB_GO(1, true); else EKSB_GO(1, false); }
#undef EKSB_GO
  return check_launch("kf_forward_kernel");
}

template <typename T, int N, int P>
static int launch_backward(const T* filt, const T* C, const T* A, const T* Q, int shared_mask, long B, long Tn,
                           int a_mode, T* smth, T* ym, T* yv, T* cv, int* status, cudaStream_t st) {
  if (!cv && use_pipe(B)) {
    const long tiles = num_tiles(B), ng = pipe_groups(tiles, (int)sizeof(T)), per = (tiles + ng - 1) / ng;
    int rc = EKSB_OK;
    for (long g0 = 0; g0 < tiles && rc == EKSB_OK; g0 += per) {
      const long b0 = g0 * kLanes, Bs = (B - b0 < per * kLanes) ? B - b0 : per * kLanes;
      auto off = [&](const T* ptr, int flag, long stride) { return ptr + ((shared_mask & flag) ? 0 : b0 * stride); };
      const long rec0 = g0 * Tn * kLanes;
      rc = rts_backward_pipe<T>(filt + rec0 * Dim<N>::SIGMA, off(C, EKSB_SHARED_C, P * N), off(A, EKSB_SHARED_A, N * N),
                                off(Q, EKSB_SHARED_Q, N * N), shared_mask, Bs, Tn, N, P, a_mode,
                                smth + rec0 * Dim<N>::SIGMA, ym ? ym + rec0 * P : nullptr, yv ? yv + rec0 * P : nullptr,
                                status ? status + b0 : nullptr, st);
    }
    if (rc != EKSB_EDIM) return rc;
  }
  unsigned grid = (unsigned)num_tiles(B);
  if (a_mode == EKSB_A_DIAG)
    rts_backward_kernel<T, N, P, 0><<<grid, kLanes, 0, st>>>(filt, C, A, Q, shared_mask, B, Tn, smth, ym, yv, cv, status);
  else
    rts_backward_kernel<T, N, P, 1><<<grid, kLanes, 0, st>>>(filt, C, A, Q, shared_mask, B, Tn, smth, ym, yv, cv, status);
  return check_launch("rts_backward_kernel");
}

template <typename T>
static int kf_forward_api(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int shared_mask,
                          long B, long Tn, int n, int p, int a_mode, int first, const T* q_scale, int n_scale, T* filt,
                          double* nll, int* status, void* stream, const T* yoff = nullptr) {
  if (!obs || !m0 || !S0 || !C || !A || !Q || B <= 0 || Tn <= 0) { set_error("kf_forward: null pointer or empty batch"); return EKSB_EARG; }
  if (first != kPriorUpdate && first != kPriorOnly) { set_error("kf_forward: bad first-frame mode %d", first); return EKSB_EARG; }
  if (a_mode != EKSB_A_DIAG && a_mode != EKSB_A_DENSE) { set_error("kf_forward: bad a_mode %d", a_mode); return EKSB_EARG; }
  if (q_scale && n_scale > 1 && filt) { set_error("kf_forward: filt must be NULL when sweeping n_scale > 1"); return EKSB_EARG; }
  if (q_scale && n_scale < 1) { set_error("kf_forward: n_scale < 1"); return EKSB_EARG; }
#define X(N_, P_)                                                                                                 \
  if (n == N_ && p == P_)                                                                                         \
    return launch_forward<T, N_, P_>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, a_mode, first, q_scale, n_scale,   \
                                     filt, nll, status, yoff, (cudaStream_t)stream);
  EKSB_FOR_EACH_DIM(X)
#undef X
  set_error("kf_forward: (n=%d,p=%d) not instantiated", n, p);
  return EKSB_EDIM;
}

template <typename T>
static int rts_backward_api(const T* filt, const T* C, const T* A, const T* Q, int shared_mask, long B, long Tn, int n,
                            int p, int a_mode, T* smth, T* ym, T* yv, T* cv, int* status, void* stream) {
  if (!filt || !smth || !C || !A || !Q || B <= 0 || Tn <= 0) { set_error("rts_backward: null pointer or empty batch"); return EKSB_EARG; }
  if (a_mode != EKSB_A_DIAG && a_mode != EKSB_A_DENSE) { set_error("rts_backward: bad a_mode %d", a_mode); return EKSB_EARG; }
#define X(N_, P_)                                                                                           \
  if (n == N_ && p == P_)                                                                                   \
    return launch_backward<T, N_, P_>(filt, C, A, Q, shared_mask, B, Tn, a_mode, smth, ym, yv, cv, status, \
                                      (cudaStream_t)stream);
  EKSB_FOR_EACH_DIM(X)
#undef X
  set_error("rts_backward: (n=%d,p=%d) not instantiated", n, p);
  return EKSB_EDIM;
}

}  // namespace eksb

// eksb_nll_pipe.cuh -- A8: the NLL of the forward filter in the warp-specialised form, for the smoothing-parameter
// grid search (eks_b200/selection.py; replaces the CLI flag of scripts/multicam_example.py:38-43).
//
// The plain thread-per-track kernel (eksb_kf.cuh, grid.y = candidates) re-does the whole measurement update -- two
// symmetric inverses, the reciprocals of every ensemble variance, a double-precision log -- per candidate and per frame.
// Nothing of the helper record (Lam = C^T R^-1 C, iv = C^T R^-1 y, yy = y^T R^-1 y, log det R) depends on the process
// noise, so here ONE CTA per track tile runs H helper warps that turn the TMA-fed obs frames into records, and up to
// kNllChains chain warps that each carry the (m, Nm, r) recursion of ONE candidate scale over the same records and
// accumulate its NLL (eksb_math.cuh fwd_nll_scaled): per candidate ~110 FP64 instructions per frame instead of ~250,
// one log per 4 frames, obs read once per 8 candidates.  Diagonal A, N <= 3 (every shipped configuration); other
// shapes keep the plain kernel.  No state records are written.
#pragma once
#include "eksb_pipe.cuh"

namespace eksb {

constexpr int kNllChains = 8, kNllHelpers = 2, kNllF = 4, kNllStages = 3;

template <typename T, int ROWS_IN, int REC> struct NllSmem {
  static constexpr size_t kRaw = (size_t)kNllStages * kNllF * ROWS_IN * kLanes * sizeof(T);
  static constexpr size_t kRec = (size_t)2 * kNllF * REC * kLanes * sizeof(T);
  static constexpr size_t kTotal = kRaw + kRec + 128;
};

// grid = (n_tiles, ceil(n_scale / sc)), block = (kNllHelpers + sc) warps; chain warp c of block y evaluates scale y*sc + c
template <typename T, int N, int P>
__global__ void __launch_bounds__((kNllHelpers + kNllChains) * kLanes)
kf_nll_pipe_kernel(const T* __restrict__ obs, const T* __restrict__ m0, const T* __restrict__ S0, const T* __restrict__ C,
                   const T* __restrict__ A, const T* __restrict__ Q, int shared_mask, long B, long Tn, int first,
                   const T* __restrict__ q_scale, int n_scale, double* __restrict__ nll, int* __restrict__ status,
                   const T* __restrict__ yoff, int sc) {
  constexpr int NS = Dim<N>::NS, R2 = 2 * P, REC = RecNll<N>::FWD;
  constexpr int H = kNllHelpers, F = kNllF, NST = kNllStages;
  using SM = NllSmem<T, R2, REC>;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  T* raw = reinterpret_cast<T*>(smem_raw);
  T* recb = reinterpret_cast<T*>(smem_raw + SM::kRaw);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + SM::kRaw + SM::kRec);
  const int bar_cnt = (H + sc) * kLanes;
  const long g = blockIdx.x;
  const long b = g * kLanes + lane, bp = b < B ? b : B - 1;
  Params<T, N, P, 0> prm;
  load_track_params<T, N, P, 0>(prm, C, A, Q, shared_mask, bp);
  const long nb = (Tn + F - 1) / F;
  constexpr uint32_t kFrameBytes = R2 * kLanes * sizeof(T);

  if (warp < H) {                                          // ---------------- helpers: obs frames -> (Lam, iv, yy, ldr)
    const int h = warp;
    const T* gsrc = obs + (g * Tn * R2) * kLanes;
    const bool issuer = (h == 0 && lane == 0);
    constexpr bool kOuter = (P * NS <= 36);
    T CC[kOuter ? P : 1][NS];
    if constexpr (kOuter) make_outer<T, N, P>(prm.C, CC);
    T mu[P];
#pragma unroll
    for (int j = 0; j < P; ++j) mu[j] = yoff ? yoff[bp * P + j] : T(0);
    bool nonpos = false;
    if (issuer) {
      for (int s = 0; s < NST; ++s) mbar_init(&full[s], 1);
      mbar_fence_init();
    }
    tile_barrier(1, bar_cnt);                              // barriers initialised
    if (issuer) {
      for (int s = 0; s < NST; ++s)
        if (s < nb) {
          const long t0 = (long)s * F;
          const uint32_t bytes = (uint32_t)((Tn - t0 < F ? Tn - t0 : F)) * kFrameBytes;
          mbar_expect_tx(&full[s], bytes);
          bulk_g2s(raw + (size_t)s * F * R2 * kLanes, gsrc + t0 * R2 * kLanes, bytes, &full[s]);
        }
    }
    for (long k = 0; k <= nb; ++k) {
      const int st = (int)(k % NST);
      if (k < nb) {
        mbar_wait(&full[st], (uint32_t)((k / NST) & 1));
        const T* src = raw + (size_t)st * F * R2 * kLanes + lane;
        T* dst = recb + (k & 1) * (F * REC * kLanes) + lane;
        const long rem = Tn - k * F;
        const int nfr = (int)(rem < F ? rem : F);
#pragma unroll
        for (int jj = 0; jj < (F + H - 1) / H; ++jj) {
          const int j = jj * H + h;
          if (j < nfr) {
            T cur[R2], rec[REC];
#pragma unroll
            for (int r = 0; r < R2; ++r) cur[r] = src[(j * R2 + r) * kLanes];
            if (yoff) {
#pragma unroll
              for (int q = 0; q < P; ++q) cur[q] -= mu[q];
            }
            nonpos = nonpos || (any_nonpos<T, P>(cur + P) && !(first == kPriorOnly && k == 0 && j == 0));
            if constexpr (kOuter) fwd_prep_outer_nll<T, N, P>(prm.C, CC, cur, cur + P, rec);
            else fwd_prep_nll<T, N, P>(prm.C, cur, cur + P, rec);
#pragma unroll
            for (int r = 0; r < REC; ++r) dst[(j * REC + r) * kLanes] = rec[r];
          }
        }
      }
      tile_barrier(1, bar_cnt);                            // batch k published; raw stage st is free
      if (issuer && k + NST < nb) {
        const long t0 = (k + NST) * F;
        const uint32_t bytes = (uint32_t)((Tn - t0 < F ? Tn - t0 : F)) * kFrameBytes;
        mbar_expect_tx(&full[st], bytes);
        bulk_g2s(raw + (size_t)st * F * R2 * kLanes, gsrc + t0 * R2 * kLanes, bytes, &full[st]);
      }
    }
    if (blockIdx.y == 0 && b < B && status) set_status(status + b, true, nonpos);
    return;
  }

  // ---------------- chain warp: one candidate scale
  const int c = warp - H;
  long s = (long)blockIdx.y * sc + c;
  const bool valid = (n_scale <= 0) ? (s == 0) : (s < n_scale);
  if (!valid) s = 0;                                       // (still takes part in the barriers)
  if (q_scale) {
    const T qs = q_scale[s];
#pragma unroll
    for (int i = 0; i < NS; ++i) prm.Q[i] *= qs;
  }
  T m[N], Nm[NS], rr = T(1);
  bool ok = true;
  {
    const T* pm = m0 + ((shared_mask & EKSB_SHARED_M0) ? 0 : bp * N);
    const T* ps = S0 + ((shared_mask & EKSB_SHARED_S0) ? 0 : bp * (N * N));
#pragma unroll
    for (int i = 0; i < N; ++i) m[i] = pm[i];
    load_sym<T, N>(Nm, ps);
    if (first == kPriorOnly) {                             // eks/newton_eks.py:132 : P starts at inv(S0), frame 0 has no update
      T Vi[NS];
      sym_inv<T, N>(Nm, Vi, ok);
#pragma unroll
      for (int i = 0; i < NS; ++i) Nm[i] = Vi[i];
    }
  }
  double acc = 0.0;
  T xprod = T(1);
  auto frame = [&](const T* src, int j, bool is_first) {
    if (is_first && first == kPriorOnly) return;
    T rec[REC], xdet;
#pragma unroll
    for (int r = 0; r < REC; ++r) rec[r] = src[(j * REC + r) * kLanes];
    const T e = fwd_nll_scaled<T, N, P>(prm, rec, !(is_first && first == kPriorUpdate), m, Nm, rr, xdet, ok);
    acc += (double)e + (double)rec[NS + N + 1];
    xprod *= xdet;
  };
  tile_barrier(1, bar_cnt);                                // (barrier init)
  tile_barrier(1, bar_cnt);                                // batch 0 ready
  for (long k = 0; k < nb; ++k) {
    const T* src = recb + (k & 1) * (F * REC * kLanes) + lane;
    if (k > 0 && (k + 1) * F < Tn) {
#pragma unroll
      for (int j = 0; j < F; ++j) frame(src, j, false);
    } else {
#pragma unroll
      for (int j = 0; j < F; ++j)
        if (k * F + j < Tn) frame(src, j, k == 0 && j == 0);
    }
    acc -= log((double)xprod);                             // one log per F = 4 frames: log det(Id + Lam S) = -log xdet
    xprod = T(1);
    tile_barrier(1, bar_cnt);                              // batch k consumed / batch k+1 ready
  }
  if (valid && b < B) {
    nll[s * B + b] = 0.5 * (acc + (double)Tn * P * 1.8378770664093453 /* log 2pi */);
    if (status) set_status(status + b, ok, false);
  }
}

template <typename T, int N, int P>
static int launch_nll_pipe(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int shared_mask, long B,
                           long Tn, int first, const T* q_scale, int n_scale, double* nll, int* status, const T* yoff,
                           cudaStream_t st) {
  using SM = NllSmem<T, 2 * P, RecNll<N>::FWD>;
  const int ns = (q_scale && n_scale > 0) ? n_scale : 1;
  const int sc = ns < kNllChains ? ns : kNllChains;
  auto kern = kf_nll_pipe_kernel<T, N, P>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SM::kTotal);
  dim3 grid((unsigned)num_tiles(B), (unsigned)((ns + sc - 1) / sc));
  kern<<<grid, (kNllHelpers + sc) * kLanes, SM::kTotal, st>>>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, first, q_scale,
                                                            q_scale ? n_scale : 0, nll, status, yoff, sc);
  return check_launch("kf_nll_pipe_kernel");
}

// diagonal A, N <= 3 only; EKSB_EDIM otherwise (the caller keeps the plain kernel)
template <typename T>
int kf_nll_pipe(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int shared_mask, long B, long Tn,
                int n, int p, int first, const T* q_scale, int n_scale, double* nll, int* status, const T* yoff, cudaStream_t st) {
#define X(N_, P_) \
  if (n == N_ && p == P_) return launch_nll_pipe<T, N_, P_>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, first, q_scale, n_scale, nll, status, yoff, st);
  X(2, 4) X(3, 4) X(3, 6) X(3, 8)
#undef X
  return EKSB_EDIM;
}

}  // namespace eksb

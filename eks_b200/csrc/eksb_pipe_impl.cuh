// eksb_pipe_impl.cuh -- launch configuration + dispatch for the warp-specialised kernels.
#pragma once
#include "eksb_pipe.cuh"

namespace eksb {

constexpr size_t kCtaSmemBudget = 113 * 1024;   // two CTAs per SM (227 KB usable)
constexpr size_t kTileSmemBudget = 56 * 1024;   // four tiles per SM

#ifndef EKSB_PIPE_TPC
#define EKSB_PIPE_TPC 4
#endif

template <typename T, int ROWS_IN, int REC, int H, int F, int NST> struct TileBytes {
  static constexpr size_t value = PipeSmem<T, ROWS_IN, REC, PipeCfg<H, F, NST, 1>>::kTile;
};

// Largest (F, NST) whose per-tile ring fits the budget: F frames per hand-off batch (fewer
// barriers when larger), NST raw batches in flight (deeper HBM prefetch when larger).
template <typename T, int H, int ROWS_IN, int REC> struct AutoCfg {
  static constexpr bool f8n3 = TileBytes<T, ROWS_IN, REC, H, 8, 3>::value <= kTileSmemBudget;
  static constexpr bool f4n3 = TileBytes<T, ROWS_IN, REC, H, 4, 3>::value <= kTileSmemBudget;
  static constexpr bool f4n2 = TileBytes<T, ROWS_IN, REC, H, 4, 2>::value <= kTileSmemBudget;
  static constexpr bool f2n3 = TileBytes<T, ROWS_IN, REC, H, 2, 3>::value <= kTileSmemBudget;
  static constexpr int F = f8n3 ? 8 : (f4n3 || f4n2) ? 4 : 2;
  static constexpr int NST = f8n3 ? 3 : f4n3 ? 3 : f4n2 ? 2 : 3;
  static constexpr size_t kTileSmem = TileBytes<T, ROWS_IN, REC, H, F, NST>::value;
  static constexpr int TPC0 = (int)(kCtaSmemBudget / kTileSmem);
  static constexpr int TPC = TPC0 > EKSB_PIPE_TPC ? EKSB_PIPE_TPC : TPC0;
  static constexpr bool ok = TPC >= 1 && (f8n3 || f4n3 || f4n2 || f2n3);
  using type = PipeCfg<H, F, NST, (TPC >= 1 ? TPC : 1)>;
};

constexpr int kFwdHelpers = 1, kBwdHelpers = 2;

template <typename T, int N, int P, int AM, int OP = 0>
static int launch_fwd_pipe(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int shared_mask,
                           long B, long Tn, int first, T* filt, int* status, cudaStream_t st) {
  using Auto = AutoCfg<T, kFwdHelpers, 2 * P, Rec<N>::FWD>;
  if (!Auto::ok) return EKSB_EDIM;
  using Cfg = typename Auto::type;
  auto kern = kf_forward_pipe_kernel<T, N, P, AM, OP, Cfg>;
  const size_t smem = Auto::kTileSmem * Cfg::TPC;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const unsigned grid = (unsigned)((num_tiles(B) + Cfg::TPC - 1) / Cfg::TPC);
  kern<<<grid, Cfg::THREADS, smem, st>>>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, first, filt, status);
  return check_launch("kf_forward_pipe_kernel");
}

template <typename T, int N, int P, int AM, int OP = 0>
static int launch_bwd_pipe(const T* filt, const T* C, const T* A, const T* Q, int shared_mask, long B, long Tn, T* smth,
                           T* ym, T* yv, int* status, cudaStream_t st) {
  using Auto = AutoCfg<T, kBwdHelpers, Dim<N>::SIGMA, Rec<N>::BWD>;
  if (!Auto::ok) return EKSB_EDIM;
  using Cfg = typename Auto::type;
  auto kern = rts_backward_pipe_kernel<T, N, P, AM, OP, Cfg>;
  const size_t smem = Auto::kTileSmem * Cfg::TPC;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const unsigned grid = (unsigned)((num_tiles(B) + Cfg::TPC - 1) / Cfg::TPC);
  kern<<<grid, Cfg::THREADS, smem, st>>>(filt, C, A, Q, shared_mask, B, Tn, smth, ym, yv, status);
  return check_launch("rts_backward_pipe_kernel");
}

// pipelined instantiations: state dim <= 4 (larger records do not fit the shared-memory ring)
#define EKSB_FOR_EACH_PIPE_DIM(X) X(2, 4) X(3, 4) X(3, 6) X(3, 8) X(4, 6)

#ifdef EKSB_PIPE_FWD
template <typename T>
int kf_forward_pipe(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int shared_mask, long B,
                    long Tn, int n, int p, int a_mode, int first, T* filt, int* status, cudaStream_t st) {
#define X(N_, P_)                                                                                                      \
  if (n == N_ && p == P_)                                                                                              \
    return a_mode == EKSB_A_DIAG                                                                                       \
               ? launch_fwd_pipe<T, N_, P_, 0>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, first, filt, status, st)      \
               : launch_fwd_pipe<T, N_, P_, 1>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, first, filt, status, st);
  EKSB_FOR_EACH_PIPE_DIM(X)
#undef X
  return EKSB_EDIM;
}
#endif

#ifdef EKSB_PIPE_BWD
template <typename T>
int rts_backward_pipe(const T* filt, const T* C, const T* A, const T* Q, int shared_mask, long B, long Tn, int n, int p,
                      int a_mode, T* smth, T* ym, T* yv, int* status, cudaStream_t st) {
#define X(N_, P_)                                                                                                 \
  if (n == N_ && p == P_)                                                                                         \
    return a_mode == EKSB_A_DIAG                                                                                  \
               ? launch_bwd_pipe<T, N_, P_, 0>(filt, C, A, Q, shared_mask, B, Tn, smth, ym, yv, status, st)       \
               : launch_bwd_pipe<T, N_, P_, 1>(filt, C, A, Q, shared_mask, B, Tn, smth, ym, yv, status, st);
  EKSB_FOR_EACH_PIPE_DIM(X)
#undef X
  return EKSB_EDIM;
}
#endif

#ifdef EKSB_PIPE_BTD
// A7: forward elimination then back-substitution; `work` receives (z_t, M_t^-1), `sol` (x_t, Sig_t)
template <typename T>
int btd_solve_pipe(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int shared_mask, long B,
                   long Tn, int n, int p, int a_mode, T* work, T* sol, T* ym, int* status, cudaStream_t st) {
#define X(N_, P_)                                                                                                       \
  if (n == N_ && p == P_) {                                                                                             \
    int rc = a_mode == EKSB_A_DIAG                                                                                      \
                 ? launch_fwd_pipe<T, N_, P_, 0, 1>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, kPriorUpdate, work, status, st) \
                 : launch_fwd_pipe<T, N_, P_, 1, 1>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, kPriorUpdate, work, status, st); \
    if (rc != EKSB_OK) return rc;                                                                                       \
    return a_mode == EKSB_A_DIAG                                                                                        \
               ? launch_bwd_pipe<T, N_, P_, 0, 1>(work, C, A, Q, shared_mask, B, Tn, sol, ym, nullptr, status, st)      \
               : launch_bwd_pipe<T, N_, P_, 1, 1>(work, C, A, Q, shared_mask, B, Tn, sol, ym, nullptr, status, st);     \
  }
  EKSB_FOR_EACH_PIPE_DIM(X)
#undef X
  set_error("btd_solve: (n=%d,p=%d) not instantiated (state dim <= 4)", n, p);
  return EKSB_EDIM;
}
#endif

}  // namespace eksb

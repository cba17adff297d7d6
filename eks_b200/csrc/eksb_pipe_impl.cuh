// eksb_pipe_impl.cuh -- launch configuration + dispatch for the warp-specialised kernels.
#pragma once
#include <stdlib.h>
#include "eksb_pipe.cuh"

namespace eksb {

constexpr size_t kCtaSmemBudget = 227 * 1024;   // one CTA may take the whole SM
constexpr size_t kTileSmemBudget = 56 * 1024;   // four tiles per SM

// Tiles per CTA for a launch over `tiles` track tiles.  A tile lives on one SM for the whole pass, so
// the pass takes as long as the most loaded SM: aim for ceil(tiles / #SMs) tiles on every busy SM, in ONE
// CTA (warp w runs on scheduler w % 4, so the chain warps 0..tpc-1 land on distinct schedulers).
inline int pick_tpc(long tiles, int tpc_max, size_t tile_smem) {
  const int kSMs = sm_count();
  long per_sm = (tiles + kSMs - 1) / kSMs;
  if (per_sm < 1) per_sm = 1;
  for (int c = 1; c <= 8; ++c) {                  // c co-resident CTAs per SM when one CTA cannot hold per_sm tiles
    const long tpc = (per_sm + c - 1) / c;
    if (tpc <= tpc_max && c * (tpc * tile_smem + 1024) <= kCtaSmemBudget) return (int)tpc;
  }
  return tpc_max;                                 // more tiles than fit at once: several waves
}

// Place the warps of a CTA (tpc chain warps + tpc*H helper warps) on the SM's four schedulers; warp index =
// 4*k + s for the k-th warp placed on scheduler s, unused slots become filler warps that exit at once.
// Returns warps per CTA.
//  * tpc <= 2: every chain warp gets a scheduler to ITSELF and the helpers share the remaining ones.  The pass
//    runs at the chain warp's pace, and a helper on the same scheduler steals its issue slots (measured, f64
//    8192 tracks sustained: backward 24.6 ms with a shared scheduler, 22.2 ms with a dedicated one).
//  * tpc >= 3: not enough schedulers to dedicate; greedy by instruction weight (w_chain per chain warp,
//    w_helper for all helpers of a tile together), at most ceil(total/4) warps per scheduler.
inline int roles_warps(int tpc, int H) {
  if (tpc <= 2) {
    const int per = (tpc * H + (4 - tpc) - 1) / (4 - tpc);
    return 4 * (per < 1 ? 1 : per);
  }
  return 4 * ((tpc * (H + 1) + 3) / 4);
}
inline int make_roles(int tpc, int H, float w_chain, float w_helper, RoleMap& rm, bool split_chain = false) {
  for (int w = 0; w < kMaxWarps; ++w) { rm.tile[w] = -1; rm.h[w] = -1; }
  rm.chain_warps = 1;
  int len[4] = {0, 0, 0, 0};
  auto put = [&](int s2, int tile, int h) {
    const int w = 4 * len[s2] + s2;
    rm.tile[w] = (signed char)tile;
    rm.h[w] = (signed char)h;
    ++len[s2];
  };
  if (tpc == 1 && split_chain) {                    // backward: covariance chain, mean chain, helpers on the rest
    rm.chain_warps = 2;
    put(0, 0, kRoleBwdCov);
    put(1, 0, kRoleBwdMean);
    for (int h = 0; h < H; ++h) put(2 + (h & 1), 0, h);
    return 4 * ((H + 1) / 2 < 1 ? 1 : (H + 1) / 2);
  }
  if (tpc <= 2) {
    for (int t = 0; t < tpc; ++t) put(t, t, -1);
    int s2 = tpc;
    for (int h = 0; h < H; ++h)
      for (int t = 0; t < tpc; ++t) {
        put(s2, t, h);
        if (++s2 == 4) s2 = tpc;
      }
    return roles_warps(tpc, H);
  }
  const int total = tpc * (H + 1), cap = (total + 3) / 4;
  float load[4] = {0, 0, 0, 0};
  auto place = [&](int tile, int h, float wgt) {
    int best = -1;
    for (int s2 = 0; s2 < 4; ++s2)
      if (len[s2] < cap && (best < 0 || load[s2] < load[best])) best = s2;
    put(best, tile, h);
    load[best] += wgt;
  };
  for (int t = 0; t < tpc; ++t) place(t, -1, w_chain);
  for (int h = 0; h < H; ++h)
    for (int t = 0; t < tpc; ++t) place(t, h, w_helper / H);
  return 4 * cap;
}

#ifndef EKSB_PIPE_TPC
#define EKSB_PIPE_TPC 4
#endif

// tiles the launch covers: real track tiles, or (tile, chunk) pairs of the time-split scan
inline long vtile_count(long B, const VTiles& vt) { return num_tiles(B) * (vt.nc > 1 ? vt.nc : 1); }

template <typename T, int ROWS_IN, int REC, int H, int F, int NST> struct TileBytes {
  static constexpr size_t value = PipeSmem<T, ROWS_IN, REC, PipeCfg<H, F, NST, 1>>::kTile;
};

// Largest (F, NST) whose per-tile ring fits the budget: F frames per hand-off batch (fewer
// barriers when larger), NST raw batches in flight (deeper HBM prefetch when larger).
template <typename T, int H, int ROWS_IN, int REC, int FORCE_F = 0, int FORCE_NST = 0> struct AutoCfg {
  static constexpr bool f8n3 = TileBytes<T, ROWS_IN, REC, H, 8, 3>::value <= kTileSmemBudget;
  static constexpr bool f4n3 = TileBytes<T, ROWS_IN, REC, H, 4, 3>::value <= kTileSmemBudget;
  static constexpr bool f4n2 = TileBytes<T, ROWS_IN, REC, H, 4, 2>::value <= kTileSmemBudget;
  static constexpr bool f2n3 = TileBytes<T, ROWS_IN, REC, H, 2, 3>::value <= kTileSmemBudget;
  static constexpr int F = FORCE_F ? FORCE_F : f8n3 ? 8 : (f4n3 || f4n2) ? 4 : 2;
  static constexpr int NST = FORCE_NST ? FORCE_NST : f8n3 ? 3 : f4n3 ? 3 : f4n2 ? 2 : 3;
  static constexpr size_t kTileSmem = TileBytes<T, ROWS_IN, REC, H, F, NST>::value;
  static constexpr int TPC0 = (int)((kCtaSmemBudget - 1024) / kTileSmem);
  static constexpr int TPC1 = TPC0 > EKSB_PIPE_TPC ? EKSB_PIPE_TPC : TPC0;
  static constexpr int TPCW = kMaxWarps / (H + 1);           // warps of a CTA: tpc * (H + 1) <= 32
  static constexpr int TPC = TPC1 > TPCW ? TPCW : TPC1;
  static constexpr bool ok = TPC >= 1 && (FORCE_F || f8n3 || f4n3 || f4n2 || f2n3);
  using type = PipeCfg<H, F, NST, (TPC >= 1 ? TPC : 1)>;
};

#ifndef EKSB_FWD_H
#define EKSB_FWD_H 2
#endif
#ifndef EKSB_BWD_H
#define EKSB_BWD_H 2
#endif
#ifndef EKSB_FWD_F      // 0 = automatic (largest batch/stage count that fits the per-tile budget)
#define EKSB_FWD_F 0
#endif
#ifndef EKSB_FWD_NST
#define EKSB_FWD_NST 0
#endif
#ifndef EKSB_BWD_F
#define EKSB_BWD_F 0
#endif
#ifndef EKSB_BWD_NST
#define EKSB_BWD_NST 0
#endif
#ifndef EKSB_BWD_WIDE_H
#define EKSB_BWD_WIDE_H 2
#endif
#ifndef EKSB_FWD_DEEP_H
#define EKSB_FWD_DEEP_H 2
#endif
#ifndef EKSB_BWD_WIDE_NST
#define EKSB_BWD_WIDE_NST 3
#endif
#ifndef EKSB_BWD_WIDE_F
#define EKSB_BWD_WIDE_F 6
#endif
#ifndef EKSB_BWD_SOLO_H
#define EKSB_BWD_SOLO_H 4
#endif
#ifndef EKSB_BWD_SOLO_F
#define EKSB_BWD_SOLO_F 8
#endif
constexpr int kFwdHelpers = EKSB_FWD_H, kBwdHelpers = EKSB_BWD_H;
// instructions per frame (n=3, p=4, f64 SASS): forward chain / helper, backward chain / helpers
constexpr float kFwdChainW = 118, kFwdHelperW = 125, kBwdChainW = 103, kBwdHelperW = 200;

template <typename T, int N, int P, int AM, int OP, typename Auto>
static int launch_fwd_cfg(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int shared_mask,
                          long B, long Tn, int first, T* filt, int* status, const T* yoff, cudaStream_t st,
                          const VTiles& vt = VTiles()) {
  using Cfg = typename Auto::type;
  auto kern = kf_forward_pipe_kernel<T, N, P, AM, OP, Cfg>;
  const long nvt = vtile_count(B, vt);
  const int tpc = pick_tpc(nvt, Cfg::TPC, Auto::kTileSmem);
  const size_t smem = Auto::kTileSmem * tpc;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(Auto::kTileSmem * Cfg::TPC));
  const unsigned grid = (unsigned)((nvt + tpc - 1) / tpc);
  RoleMap rm;
  const int warps = make_roles(tpc, Cfg::H, kFwdChainW, kFwdHelperW, rm);
  kern<<<grid, warps * kLanes, smem, st>>>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, first, filt, status, yoff, tpc, rm, vt);
  return check_launch("kf_forward_pipe_kernel");
}

// Three-role forward (kf_forward_pipe3_kernel): covariance warp of tile t alone on scheduler t, the mean
// warps and helpers round-robin on the other schedulers.  tpc <= 2 only.
inline int make_roles3(int tpc, int H, RoleMap& rm) {
  for (int w = 0; w < kMaxWarps; ++w) { rm.tile[w] = -1; rm.h[w] = 0; }
  rm.chain_warps = 2;
  int len[4] = {0, 0, 0, 0};
  auto put = [&](int s2, int tile, int h) {
    const int w = 4 * len[s2] + s2;
    rm.tile[w] = (signed char)tile;
    rm.h[w] = (signed char)h;
    ++len[s2];
  };
  for (int t = 0; t < tpc; ++t) put(t, t, kRoleCov);
  int s2 = tpc;
  auto next = [&]() { if (++s2 == 4) s2 = tpc; };
  for (int t = 0; t < tpc; ++t) { put(s2, t, kRoleMean); next(); }
  for (int h = 0; h < H; ++h)
    for (int t = 0; t < tpc; ++t) { put(s2, t, h); next(); }
  int mx = 1;
  for (int i = 0; i < 4; ++i) mx = len[i] > mx ? len[i] : mx;
  return 4 * mx;
}

template <int H_, int F_, int NST_> struct Pipe3Cfg {
  static constexpr int H = H_, F = F_, NST = NST_;
  static constexpr int THREADS = 4 * ((2 * (H + 1) + 1) / 2) * 32;   // two tiles: (H + 1) warps on each of 2 schedulers
};
template <typename T, int N, int P, int H> struct Auto3 {
  template <int F, int NST> static constexpr size_t bytes() {
    return Pipe3Smem<T, 2 * P, Rec<N>::FWD, Dim<N>::NS + 1, Pipe3Cfg<H, F, NST>>::kTile;
  }
  static constexpr size_t kTwo = (kCtaSmemBudget - 1024) / 2;
  static constexpr bool a = bytes<8, 4>() <= kTwo, b = bytes<6, 4>() <= kTwo, c = bytes<6, 3>() <= kTwo,
                        d = bytes<4, 3>() <= kTwo;
  static constexpr int F = a ? 8 : (b || c) ? 6 : 4;
  static constexpr int NST = (a || b) ? 4 : 3;
  static constexpr bool ok = a || b || c || d;
  using type = Pipe3Cfg<H, F, NST>;
  static constexpr size_t kTileSmem = bytes<F, NST>();
};

// helpers of the three-role forward kernel.  Measured at one tile per SM, 100 000 frames: f64 13.6 / 11.7 / 11.2 ms
// with 2 / 3 / 4 helpers, f32 7.9 / 9.7 / 10.2 ms.
#ifndef EKSB_FWD3_H64
#define EKSB_FWD3_H64 4
#endif
#ifndef EKSB_FWD3_H32
#define EKSB_FWD3_H32 2
#endif
template <typename T> struct Fwd3Helpers { static constexpr int value = sizeof(T) == 8 ? EKSB_FWD3_H64 : EKSB_FWD3_H32; };

template <typename T, int N, int P>
static int launch_fwd_pipe3(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int shared_mask,
                            long B, long Tn, int first, T* filt, int* status, const T* yoff, int tpc, cudaStream_t st) {
  using Auto = Auto3<T, N, P, Fwd3Helpers<T>::value>;
  using Cfg = typename Auto::type;
  auto kern = kf_forward_pipe3_kernel<T, N, P, Cfg>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(Auto::kTileSmem * 2));
  const unsigned grid = (unsigned)((num_tiles(B) + tpc - 1) / tpc);
  RoleMap rm;
  const int warps = make_roles3(tpc, Cfg::H, rm);
  kern<<<grid, warps * kLanes, Auto::kTileSmem * tpc, st>>>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, first, filt,
                                                            status, yoff, tpc, rm);
  return check_launch("kf_forward_pipe3_kernel");
}

// Two ring shapes, like the backward pass.  With few tiles per SM the bytes in flight per SM (stages x
// frames x record) bound the read bandwidth (Little's law: 256 tiles x 24 KB = 6 MB in flight sustained
// ~5.7 TB/s algorithmic in f64); the "deep" ring (8-frame batches, 4 stages: 4x the bytes in flight) lifts
// that to ~6.4 TB/s.  It needs ~100 KB per f64 tile, so it is used when the tiles of an SM fit.
template <typename T, int N, int P, int AM, int OP = 0>
static int launch_fwd_pipe(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int shared_mask,
                           long B, long Tn, int first, T* filt, int* status, const T* yoff, cudaStream_t st,
                           const VTiles& vt = VTiles()) {
  using Shallow = AutoCfg<T, kFwdHelpers, 2 * P, Rec<N>::FWD, EKSB_FWD_F, EKSB_FWD_NST>;
  using Deep = AutoCfg<T, EKSB_FWD_DEEP_H, 2 * P, Rec<N>::FWD, 8, 4>;
  const long nvt = vtile_count(B, vt);
  const int kSMs = sm_count();
  if constexpr (OP == 0 && AM == 0 && N <= 3 && Auto3<T, N, P, Fwd3Helpers<T>::value>::ok) {
    const long per_sm = (nvt + kSMs - 1) / kSMs;
    // Measured (f64, 100 000 frames, sustained): one tile per SM 16.2 -> 13.5 ms (the pass runs at the chain's
    // pace); two tiles per SM 18.3 -> 20.8 ms (the FP64 pipes of the two helper schedulers saturate once they
    // also carry the mean warps, and the two-role kernel is already HBM-bound there).  Hence: one tile per SM only.
    static const int max_per_sm = getenv("EKSB_FWD3_MAX") ? atoi(getenv("EKSB_FWD3_MAX")) : 1;
    if (vt.nc <= 1 && per_sm <= max_per_sm && per_sm <= 2)     // (the three-role kernel has no virtual-tile mode)
      return launch_fwd_pipe3<T, N, P>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, first, filt, status, yoff,
                                       (int)(per_sm < 1 ? 1 : per_sm), st);
  }
  if constexpr (N <= 3 && Deep::ok) {
    const long per_sm = (nvt + kSMs - 1) / kSMs;
    static const bool shallow_only = getenv("EKSB_FWD_SHALLOW") != nullptr;
    if (!shallow_only && per_sm <= Deep::TPC && per_sm * Deep::kTileSmem + 1024 <= kCtaSmemBudget)
      return launch_fwd_cfg<T, N, P, AM, OP, Deep>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, first, filt, status, yoff, st, vt);
  }
  if (!Shallow::ok) return EKSB_EDIM;
  return launch_fwd_cfg<T, N, P, AM, OP, Shallow>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, first, filt, status, yoff, st, vt);
}

template <typename T, int N, int P, int AM, int OP, int OUT, typename Auto>
static int launch_bwd_cfg(const T* filt, const T* C, const T* A, const T* Q, int shared_mask, long B, long Tn,
                          T* smth, T* ym, T* yv, int* status, cudaStream_t st, const VTiles& vt = VTiles()) {
  using Cfg = typename Auto::type;
  auto kern = rts_backward_pipe_kernel<T, N, P, AM, OP, OUT, Cfg>;
  const long nvt = vtile_count(B, vt);
  const int tpc = pick_tpc(nvt, Cfg::TPC, Auto::kTileSmem);
  const size_t smem = Auto::kTileSmem * tpc;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(Auto::kTileSmem * Cfg::TPC));
  const unsigned grid = (unsigned)((nvt + tpc - 1) / tpc);
  RoleMap rm;
  static const bool no_split = getenv("EKSB_BWD_NO_SPLIT") != nullptr;
  const int warps = make_roles(tpc, Cfg::H, kBwdChainW, kBwdHelperW, rm, !no_split);
  kern<<<grid, warps * kLanes, smem, st>>>(filt, C, A, Q, shared_mask, B, Tn, smth, ym, yv, status, tpc, rm, vt);
  return check_launch("rts_backward_pipe_kernel");
}

// Two ring shapes.  "Wide" (3 helpers, 6-frame batches) keeps the helpers ahead of the chain warp when
// an SM holds few tiles (measured, f64 8192 tracks: 5.82 -> 4.85 ms per 20 000 frames); it needs
// 1.5x the shared memory per tile, so when more tiles must share an SM the "narrow" ring (2 helpers,
// 4-frame batches) is used.  Wide is instantiated for state dim <= 3 only (every shipped config).
template <typename T, int N, int P, int AM, int OP, int OUT>
static int launch_bwd_pipe_out(const T* filt, const T* C, const T* A, const T* Q, int shared_mask, long B, long Tn,
                               T* smth, T* ym, T* yv, int* status, cudaStream_t st, const VTiles& vt = VTiles()) {
  const long nvt = vtile_count(B, vt);
  const int kSMs = sm_count();
  using Narrow = AutoCfg<T, kBwdHelpers, Dim<N>::SIGMA, Rec<N>::BWD, EKSB_BWD_F, EKSB_BWD_NST>;
  using Wide = AutoCfg<T, EKSB_BWD_WIDE_H, Dim<N>::SIGMA, Rec<N>::BWD, EKSB_BWD_WIDE_F, EKSB_BWD_WIDE_NST>;
  // "Solo": an SM holds one tile, so the whole shared memory and all four schedulers serve it -- 4 helpers on
  // 8-frame batches, covariance and mean chains on two warps (f64 4736 tracks x 100 000 frames: 15.8 -> 13.0 ms).
  using Solo = AutoCfg<T, EKSB_BWD_SOLO_H, Dim<N>::SIGMA, Rec<N>::BWD, EKSB_BWD_SOLO_F, 3>;
  if constexpr (N <= 3 && Solo::ok) {
    static const bool no_solo = getenv("EKSB_BWD_NO_SOLO") != nullptr;
    if (!no_solo && nvt <= kSMs)
      return launch_bwd_cfg<T, N, P, AM, OP, OUT, Solo>(filt, C, A, Q, shared_mask, B, Tn, smth, ym, yv, status, st, vt);
  }
  if constexpr (N <= 3 && Wide::ok) {
    const long per_sm = (nvt + kSMs - 1) / kSMs;
    static const bool narrow_only = getenv("EKSB_BWD_NARROW") != nullptr;
    if (!narrow_only && per_sm <= Wide::TPC && per_sm * Wide::kTileSmem + 1024 <= kCtaSmemBudget)
      return launch_bwd_cfg<T, N, P, AM, OP, OUT, Wide>(filt, C, A, Q, shared_mask, B, Tn, smth, ym, yv, status, st, vt);
  }
  if (!Narrow::ok) return EKSB_EDIM;
  return launch_bwd_cfg<T, N, P, AM, OP, OUT, Narrow>(filt, C, A, Q, shared_mask, B, Tn, smth, ym, yv, status, st, vt);
}

// The projections are compile-time variants: a runtime `if (yv)` leaves ~50 predicated-off
// FMAs per frame on the chain warp's issue path (seen in the ncu source page, round 1).
template <typename T, int N, int P, int AM, int OP = 0>
static int launch_bwd_pipe(const T* filt, const T* C, const T* A, const T* Q, int shared_mask, long B, long Tn, T* smth,
                           T* ym, T* yv, int* status, cudaStream_t st, const VTiles& vt = VTiles()) {
  if (ym && yv) return launch_bwd_pipe_out<T, N, P, AM, OP, 3>(filt, C, A, Q, shared_mask, B, Tn, smth, ym, yv, status, st, vt);
  if (ym) return launch_bwd_pipe_out<T, N, P, AM, OP, 1>(filt, C, A, Q, shared_mask, B, Tn, smth, ym, yv, status, st, vt);
  if (yv) return launch_bwd_pipe_out<T, N, P, AM, OP, 2>(filt, C, A, Q, shared_mask, B, Tn, smth, ym, yv, status, st, vt);
  return launch_bwd_pipe_out<T, N, P, AM, OP, 0>(filt, C, A, Q, shared_mask, B, Tn, smth, ym, yv, status, st, vt);
}

// pipelined instantiations: state dim <= 4 (larger records do not fit the shared-memory ring)
#define EKSB_FOR_EACH_PIPE_DIM(X) X(2, 4) X(3, 4) X(3, 6) X(3, 8) X(4, 6)

#ifdef EKSB_PIPE_FWD
template <typename T>
int kf_forward_pipe(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int shared_mask, long B,
                    long Tn, int n, int p, int a_mode, int first, T* filt, int* status, const T* yoff, cudaStream_t st) {
#define X(N_, P_)                                                                                                      \
  if (n == N_ && p == P_)                                                                                              \
    return a_mode == EKSB_A_DIAG                                                                                       \
               ? launch_fwd_pipe<T, N_, P_, 0>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, first, filt, status, yoff, st) \
               : launch_fwd_pipe<T, N_, P_, 1>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, first, filt, status, yoff, st);
  EKSB_FOR_EACH_PIPE_DIM(X)
#undef X
  return EKSB_EDIM;
}
#endif

#ifdef EKSB_PIPE_FWD
template <typename T>
int kf_forward_pipe_vt(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int shared_mask, long B,
                       long Tn, int n, int p, int a_mode, int first, T* filt, int* status, const T* yoff, VTiles vt,
                       cudaStream_t st) {
#define X(N_, P_)                                                                                                          \
  if (n == N_ && p == P_)                                                                                                  \
    return a_mode == EKSB_A_DIAG                                                                                           \
               ? launch_fwd_pipe<T, N_, P_, 0>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, first, filt, status, yoff, st, vt) \
               : launch_fwd_pipe<T, N_, P_, 1>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, first, filt, status, yoff, st, vt);
  EKSB_FOR_EACH_PIPE_DIM(X)
#undef X
  return EKSB_EDIM;
}
// Virtual tiles that fill the GPU: 148 SMs x the tiles one SM holds with the ring the forward pass would use
// (the backward rings are no larger).
inline size_t pipe_vt_target_tiles_impl(int n, int p, int elem_size) {
#define X(N_, P_)                                                                                              \
  if (n == N_ && p == P_)                                                                                      \
    return (size_t)sm_count() * (elem_size == 8 ? AutoCfg<double, kFwdHelpers, 2 * P_, Rec<N_>::FWD, EKSB_FWD_F, EKSB_FWD_NST>::TPC \
                                          : AutoCfg<float, kFwdHelpers, 2 * P_, Rec<N_>::FWD, EKSB_FWD_F, EKSB_FWD_NST>::TPC);
  EKSB_FOR_EACH_PIPE_DIM(X)
#undef X
  return 0;
}
#endif

#ifdef EKSB_PIPE_BWD
template <typename T>
int rts_backward_pipe_vt(const T* filt, const T* C, const T* A, const T* Q, int shared_mask, long B, long Tn, int n, int p,
                         int a_mode, T* smth, T* ym, int* status, VTiles vt, cudaStream_t st) {
#define X(N_, P_)                                                                                                       \
  if (n == N_ && p == P_)                                                                                               \
    return a_mode == EKSB_A_DIAG                                                                                        \
               ? launch_bwd_pipe<T, N_, P_, 0>(filt, C, A, Q, shared_mask, B, Tn, smth, ym, (T*)nullptr, status, st, vt) \
               : launch_bwd_pipe<T, N_, P_, 1>(filt, C, A, Q, shared_mask, B, Tn, smth, ym, (T*)nullptr, status, st, vt);
  EKSB_FOR_EACH_PIPE_DIM(X)
#undef X
  return EKSB_EDIM;
}

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
                 ? launch_fwd_pipe<T, N_, P_, 0, 1>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, kPriorUpdate, work, status, (const T*)nullptr, st) \
                 : launch_fwd_pipe<T, N_, P_, 1, 1>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, kPriorUpdate, work, status, (const T*)nullptr, st); \
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

// large batches: groups of tiles, like launch_forward / launch_backward (eksb_kf.cuh)
template <typename T>
int btd_solve_grouped(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int shared_mask, long B,
                      long Tn, int n, int p, int a_mode, T* work, T* sol, T* ym, int* status, cudaStream_t st) {
  const long tiles = num_tiles(B), ng = pipe_groups(tiles, (int)sizeof(T)), per = (tiles + ng - 1) / ng;
  const long sg = n + n * (n + 1) / 2;
  int rc = EKSB_OK;
  for (long g0 = 0; g0 < tiles && rc == EKSB_OK; g0 += per) {
    const long b0 = g0 * kLanes, Bs = (B - b0 < per * kLanes) ? B - b0 : per * kLanes, rec0 = g0 * Tn * kLanes;
    auto off = [&](const T* ptr, int flag, long stride) { return ptr + ((shared_mask & flag) ? 0 : b0 * stride); };
    rc = btd_solve_pipe<T>(obs + rec0 * 2 * p, off(m0, EKSB_SHARED_M0, n), off(S0, EKSB_SHARED_S0, n * n),
                           off(C, EKSB_SHARED_C, p * n), off(A, EKSB_SHARED_A, n * n), off(Q, EKSB_SHARED_Q, n * n),
                           shared_mask, Bs, Tn, n, p, a_mode, work + rec0 * sg, sol + rec0 * sg,
                           ym ? ym + rec0 * p : nullptr, status ? status + b0 : nullptr, st);
  }
  return rc;
}
#endif

}  // namespace eksb

// eksb_pipe.cuh -- warp-specialised, TMA-fed forward-filter / RTS kernels for the latency-bound
// regime (few thousand tracks x very long T, e.g. 16 384 x 100 000: ~3.5 track tiles per SM).
//
// A sequential recursion exposes one warp per 32 tracks.  Two things then dominate a plain
// thread-per-track kernel (eksb_kf.cuh): (1) a lone warp issues ~1 dependent instruction every
// 4 cycles, so the frame time is the instruction latency of the WHOLE frame program; (2) a
// register prefetch ring cannot hide HBM latency because a warp has only 6 scoreboard slots:
// the wait on frame t's loads also waits on the younger prefetches sharing the slot (measured:
// ~600 ns per frame regardless of ring depth).  This kernel removes both:
//   * raw frame records are brought HBM -> shared memory by 1-D bulk async copies
//     (cp.async.bulk + mbarrier complete_tx; the track-tile layout makes F consecutive frames of
//     a tile one contiguous run), NST batches in flight per tile, no registers, no scoreboards;
//   * each tile gets 1 "chain" warp + H "helper" warps.  Helpers turn raw frames into compact
//     state-independent records (eksb_math.cuh fwd_prep / bwd_prep) in a double-buffered
//     shared-memory ring; the chain warp runs only the true dependency chain (fwd_chain /
//     bwd_chain) and stores the results with full-line writes.
// Hand-off is one named barrier (bar.sync id, (H+1)*32) per batch of F frames, per tile, so
// tiles sharing a CTA do not couple.  The launch (eksb_pipe_impl.cuh) packs ceil(tiles/148) tiles
// into a CTA so every busy SM carries the same number of tiles, and hands the kernel a RoleMap that
// puts each chain warp on a scheduler (SM sub-partition) of its own where the CTA is small enough.
#pragma once
#include <stdint.h>
#include "eksb_common.cuh"

namespace eksb {

__device__ __forceinline__ void tile_barrier(int id, int count) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t done;
  do {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(addr), "r"(parity)
        : "memory");
  } while (!done);
}
// global -> shared bulk async copy (TMA engine, 1-D); bytes % 16 == 0, both addresses 16 B aligned
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// H helper warps per tile, F frames per batch, NST raw batches in flight, at most TPC tiles per CTA
// (the launch picks the actual tiles per CTA, `tpc` <= TPC, so that tiles spread evenly over the SMs)
template <int H_, int F_, int NST_, int TPC_> struct PipeCfg {
  static constexpr int H = H_, F = F_, NST = NST_, TPC = TPC_;
  // upper bound on warps per CTA incl. filler warps: the largest value eksb_pipe_impl.cuh roles_warps returns
  // for 1..TPC tiles per CTA (dedicated-scheduler placement for 1-2 tiles, packed placement for more)
  static constexpr int warps_for(int tpc) {
    if (tpc <= 2) {
      const int per = (tpc * H + (4 - tpc) - 1) / (4 - tpc);
      return 4 * (per < 1 ? 1 : per);
    }
    return 4 * ((tpc * (H + 1) + 3) / 4);
  }
  static constexpr int max_warps() {
    int m = 4;
    for (int t = 1; t <= TPC; ++t) m = warps_for(t) > m ? warps_for(t) : m;
    return m;
  }
  static constexpr int WARPS_C = max_warps();
  static constexpr int THREADS = (WARPS_C > 32 ? 32 : WARPS_C) * kLanes;
};

// Which job each warp of a CTA does.  Warp w of a CTA issues from scheduler (SM sub-partition)
// w % 4, and with so few warps per SM the pass runs at the pace of the busiest scheduler, so the
// launch code (eksb_pipe_impl.cuh: make_roles) places chain and helper warps to even out the four
// schedulers' instruction load instead of using a fixed warp -> role formula.
constexpr int kMaxWarps = 32;
struct RoleMap {
  signed char tile[kMaxWarps];   // tile within the CTA, or -1: idle warp (exits at once)
  signed char h[kMaxWarps];      // -1: chain warp; 0..H-1: helper index
  int chain_warps;               // backward kernel: 1, or 2 = covariance chain (h = -2) and mean chain (h = -3) apart
};
constexpr int kRoleBwdCov = -2, kRoleBwdMean = -3;

template <typename T, int ROWS_IN, int REC, typename Cfg> struct PipeSmem {
  static constexpr size_t kRaw = (size_t)Cfg::NST * Cfg::F * ROWS_IN * kLanes * sizeof(T);
  static constexpr size_t kRec = (size_t)2 * Cfg::F * REC * kLanes * sizeof(T);
  static constexpr size_t kBar = 128;                      // NST mbarriers (8 B each), padded
  static constexpr size_t kTile = kRaw + kRec + kBar;
  static_assert(Cfg::NST * 8 <= 128, "too many stages");
};

// ------------------------------------------------------------------------------------------
// forward:  obs records -> filt records
// ------------------------------------------------------------------------------------------
// OP = 0: Kalman filter (A2/A6).  OP = 1: forward elimination of the block-tridiagonal Newton
// system (A7, eksb_math.cuh btd_fwd_chain); the record written is (z_t, M_t^-1).
template <typename T, int N, int P, int AM, int OP, typename Cfg>
__global__ void __launch_bounds__(Cfg::THREADS)
kf_forward_pipe_kernel(const T* __restrict__ obs, const T* __restrict__ m0, const T* __restrict__ S0,
                       const T* __restrict__ C, const T* __restrict__ A, const T* __restrict__ Q, int shared_mask,
                       long B, long Tn_full, int first, T* __restrict__ filt, int* __restrict__ status,
                       const T* __restrict__ yoff, int tpc, const __grid_constant__ RoleMap roles,
                       const __grid_constant__ VTiles vt) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, R2 = 2 * P, REC = Rec<N>::FWD;
  constexpr int H = Cfg::H, F = Cfg::F, NST = Cfg::NST;
  const int TPC = tpc;
  using SM = PipeSmem<T, R2, REC, Cfg>;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile_in_cta = roles.tile[warp];
  if (tile_in_cta < 0) return;                             // filler warp (keeps w % 4 placement)
  const bool is_chain = roles.h[warp] < 0;
  const int h = is_chain ? 0 : roles.h[warp];
  // virtual tiles (eksb_scan.cuh): slot v = (tile g, chunk c) covers frames [t_begin, t_end) of tile g; it starts from
  // the prior (c == 0) or from the exact filtered record of frame t_begin - 1 the warm-up kernel left in filt
  const bool virt = vt.nc > 1;
  if (virt && *vt.gate != 0) return;                       // some chunk never forgot its start: the fallback kernels run
  const long v = (long)blockIdx.x * TPC + tile_in_cta;
  if (v >= num_tiles_dev(B) * (virt ? vt.nc : 1)) return;  // chain + helpers of a tile leave together
  long g = v, t_begin = 0, t_end = Tn_full;
  if (virt) {
    g = v / vt.nc;
    const long c = v - g * vt.nc;
    t_begin = c ? c * vt.L + vt.cut[g * vt.nc + c] : 0;
    t_end = (c + 1 < vt.nc) ? (c + 1) * vt.L + vt.cut[g * vt.nc + c + 1] - 1 : Tn_full;
  }
  const long Tn = t_end - t_begin;                         // frames of THIS (virtual) tile
  const bool has_prior = (t_begin == 0);
  unsigned char* tile_smem = smem_raw + (size_t)tile_in_cta * SM::kTile;
  T* raw = reinterpret_cast<T*>(tile_smem);
  T* recb = reinterpret_cast<T*>(tile_smem + SM::kRaw);
  uint64_t* full = reinterpret_cast<uint64_t*>(tile_smem + SM::kRaw + SM::kRec);
  const int bar_id = tile_in_cta + 1, bar_cnt = (H + 1) * kLanes;
  const long b = g * kLanes + lane;
  const long bp = b < B ? b : B - 1;
  Params<T, N, P, AM> prm;
  load_track_params<T, N, P, AM>(prm, C, A, Q, shared_mask, bp);
  const long nb = (Tn + F - 1) / F;
  const T* gsrc = obs + ((g * Tn_full + t_begin) * R2) * kLanes;   // (virtual) tile base (lane 0)
  constexpr uint32_t kFrameBytes = R2 * kLanes * sizeof(T);

  if (!is_chain) {
    const bool issuer = (h == 0 && lane == 0);
    constexpr bool kOuter = (P * NS <= 36);                // outer products C_j C_j^T kept in registers
    T CC[kOuter ? P : 1][NS];
    if (kOuter) make_outer<T, N, kOuter ? P : 1>(prm.C, CC);
    // optional per-track observation offset (camera means, multiview_pca_smoother.py:702-705): y - mu is formed here,
    // with the same rounding as a separate centring pass over the records, which this replaces
    T mu[P];
#pragma unroll
    for (int j = 0; j < P; ++j) mu[j] = yoff ? yoff[bp * P + j] : T(0);
    bool nonpos = false;                                   // a frame with a non-positive ensemble variance (status bit 2)
    if (issuer) {
      for (int s = 0; s < NST; ++s) mbar_init(&full[s], 1);
      mbar_fence_init();
    }
    tile_barrier(bar_id, bar_cnt);                         // barriers initialised
    if (issuer) {
      for (int s = 0; s < NST; ++s) {
        if (s < nb) {
          const long t0 = (long)s * F;
          const uint32_t bytes = (uint32_t)((Tn - t0 < F ? Tn - t0 : F)) * kFrameBytes;
          mbar_expect_tx(&full[s], bytes);
          bulk_g2s(raw + (size_t)s * F * R2 * kLanes, gsrc + t0 * R2 * kLanes, bytes, &full[s]);
        }
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
            nonpos = nonpos || (any_nonpos<T, P>(cur + P) && !(has_prior && first == kPriorOnly && k == 0 && j == 0));
            if constexpr (kOuter) fwd_prep_outer<T, N, P>(prm.C, CC, cur, cur + P, rec);
            else fwd_prep<T, N, P, AM>(prm, cur, cur + P, rec);
#pragma unroll
            for (int r = 0; r < REC; ++r) dst[(j * REC + r) * kLanes] = rec[r];
          }
        }
      }
      tile_barrier(bar_id, bar_cnt);                       // batch k published; raw stage st is free
      if (issuer && k + NST < nb) {
        const long t0 = (k + NST) * F;
        const uint32_t bytes = (uint32_t)((Tn - t0 < F ? Tn - t0 : F)) * kFrameBytes;
        mbar_expect_tx(&full[st], bytes);
        bulk_g2s(raw + (size_t)st * F * R2 * kLanes, gsrc + t0 * R2 * kLanes, bytes, &full[st]);
      }
    }
    if (b < B && status) set_status(status + b, true, nonpos);
    return;
  }

  // ---------------- chain warp: the sequential dependency only
  T m[N], V[NS];
  {
    const T* pm = m0 + ((shared_mask & EKSB_SHARED_M0) ? 0 : bp * N);
    const T* ps = S0 + ((shared_mask & EKSB_SHARED_S0) ? 0 : bp * (N * N));
#pragma unroll
    for (int i = 0; i < N; ++i) m[i] = pm[i];
    load_sym<T, N>(V, ps);
  }
  if (!has_prior) {                                        // exact filtered state of frame t_begin - 1 (warm-up kernel)
    const T* s0 = filt + ((g * Tn_full + t_begin - 1) * SG) * kLanes + lane;
#pragma unroll
    for (int i = 0; i < N; ++i) m[i] = s0[i * kLanes];
#pragma unroll
    for (int i = 0; i < NS; ++i) V[i] = s0[(N + i) * kLanes];
  }
  bool ok = true;
  BtdConst<T, N> bc;
  T iS0[NS], iS0m0[N];
  if (OP == 1) {
    btd_make_const<T, N, P, AM>(prm, bc, ok);
    sym_inv<T, N>(V, iS0, ok);
    sym_matvec<T, N>(iS0, m, iS0m0);
#pragma unroll
    for (int i = 0; i < N; ++i) m[i] = T(0);
#pragma unroll
    for (int i = 0; i < NS; ++i) V[i] = T(0);
  } else if (has_prior && first == kPriorOnly) {
    T Vi[NS];
    sym_inv<T, N>(V, Vi, ok);
#pragma unroll
    for (int i = 0; i < NS; ++i) V[i] = Vi[i];
  }
  T* f = filt + ((g * Tn_full + t_begin) * SG) * kLanes + lane;
  // covariance carried as V = Nm * rr between frames (fwd_chain_scaled) where the fast path applies
  constexpr bool kScaled = (OP == 0 && AM == 0 && N <= 3);
  T Nm[NS], rr = T(1);
#pragma unroll
  for (int i = 0; i < NS; ++i) Nm[i] = V[i];
  // one frame of the chain.  FIRST: frame 0 of the track (prior handling); LAST only matters to OP 1.
  auto frame = [&](const T* src, int j, bool is_first, bool is_last) {
    T rec[REC];
#pragma unroll
    for (int r = 0; r < REC; ++r) rec[r] = src[(j * REC + r) * kLanes];
    if constexpr (OP == 1) {
      btd_fwd_chain<T, N>(bc, rec, is_first, is_last, iS0, iS0m0, m, V, ok);
    } else if constexpr (kScaled) {
      if (!(is_first && first == kPriorOnly))
        fwd_chain_scaled<T, N, P>(prm, rec, !(is_first && first == kPriorUpdate), m, Nm, rr, V, ok);
    } else {
      if (!(is_first && first == kPriorOnly))
        fwd_chain<T, N, P, AM>(prm, rec, !(is_first && first == kPriorUpdate), m, V, ok);
    }
    T* o = f + j * (SG * kLanes);
#pragma unroll
    for (int i = 0; i < N; ++i) o[i * kLanes] = m[i];
#pragma unroll
    for (int i = 0; i < NS; ++i) o[(N + i) * kLanes] = V[i];
  };
  tile_barrier(bar_id, bar_cnt);                           // (barrier init)
  tile_barrier(bar_id, bar_cnt);                           // batch 0 ready
  // interior batches (all F frames present, neither frame 0 nor frame T-1 inside): no guards
  for (long k = 0; k < nb; ++k) {
    const T* src = recb + (k & 1) * (F * REC * kLanes) + lane;
    if (k > 0 && (k + 1) * F < Tn) {
#pragma unroll
      for (int j = 0; j < F; ++j) frame(src, j, false, false);
    } else {
#pragma unroll
      for (int j = 0; j < F; ++j) {
        const long t = k * F + j;
        if (t < Tn) frame(src, j, has_prior && t == 0, t == Tn - 1);
      }
    }
    f += F * SG * kLanes;
    tile_barrier(bar_id, bar_cnt);                         // batch k consumed / batch k+1 ready
  }
  if (b < B && status) set_status(status + b, ok, false);
}

// ------------------------------------------------------------------------------------------
// forward, three roles per tile (Kalman filter with diagonal A, N <= 3): the chain is split in two.
//   helpers   : obs frames -> info records (Lam, iv)                      batch k     in period k
//   cov warp  : (Nm, r) recursion, needs only Lam; writes (Nm, r) records  batch k - 1 in period k
//   mean warp : m recursion from (Lam, iv) and (Nm, r); stores (m, V)      batch k - 2 in period k
// One named barrier per period among the H + 2 warps of the tile.  The info ring has 3 slots (written at k,
// read at k-1 and k-2), the (Nm, r) ring 2.  The covariance recursion is the pass's critical path; with the
// mean update, the normalisation V = Nm r and all global stores on another warp (and scheduler) it carries
// ~70 instead of ~110 instructions per frame.  Used when every chain warp can have a scheduler of its own
// (<= 2 tiles per CTA); results are identical to kf_forward_pipe_kernel (same arithmetic, same order).
// ------------------------------------------------------------------------------------------
constexpr int kRoleCov = -1, kRoleMean = -2;

template <typename T, int ROWS_IN, int REC, int VR, typename Cfg> struct Pipe3Smem {
  static constexpr size_t kRaw = (size_t)Cfg::NST * Cfg::F * ROWS_IN * kLanes * sizeof(T);
  static constexpr size_t kRec = (size_t)3 * Cfg::F * REC * kLanes * sizeof(T);
  static constexpr size_t kV = (size_t)2 * Cfg::F * VR * kLanes * sizeof(T);
  static constexpr size_t kBar = 128;
  static constexpr size_t kTile = kRaw + kRec + kV + kBar;
};

template <typename T, int N, int P, typename Cfg>
__global__ void __launch_bounds__(Cfg::THREADS)
kf_forward_pipe3_kernel(const T* __restrict__ obs, const T* __restrict__ m0, const T* __restrict__ S0,
                        const T* __restrict__ C, const T* __restrict__ A, const T* __restrict__ Q, int shared_mask,
                        long B, long Tn, int first, T* __restrict__ filt, int* __restrict__ status,
                        const T* __restrict__ yoff, int tpc, const __grid_constant__ RoleMap roles) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, R2 = 2 * P, REC = Rec<N>::FWD, VR = NS + 1;
  constexpr int H = Cfg::H, F = Cfg::F, NST = Cfg::NST;
  using SM = Pipe3Smem<T, R2, REC, VR, Cfg>;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile_in_cta = roles.tile[warp];
  if (tile_in_cta < 0) return;
  const int role = roles.h[warp];
  const long g = (long)blockIdx.x * tpc + tile_in_cta;
  if (g >= num_tiles_dev(B)) return;
  unsigned char* tile_smem = smem_raw + (size_t)tile_in_cta * SM::kTile;
  T* raw = reinterpret_cast<T*>(tile_smem);
  T* recb = reinterpret_cast<T*>(tile_smem + SM::kRaw);
  T* vb = reinterpret_cast<T*>(tile_smem + SM::kRaw + SM::kRec);
  uint64_t* full = reinterpret_cast<uint64_t*>(tile_smem + SM::kRaw + SM::kRec + SM::kV);
  const int bar_id = tile_in_cta + 1, bar_cnt = (H + 2) * kLanes;
  const long b = g * kLanes + lane;
  const long bp = b < B ? b : B - 1;
  Params<T, N, P, 0> prm;
  load_track_params<T, N, P, 0>(prm, C, A, Q, shared_mask, bp);
  const long nb = (Tn + F - 1) / F;
  constexpr uint32_t kFrameBytes = R2 * kLanes * sizeof(T);

  if (role >= 0) {                                         // ---------------- helper
    const int h = role;
    const T* gsrc = obs + (g * Tn * R2) * kLanes;
    const bool issuer = (h == 0 && lane == 0);
    constexpr bool kOuter = (P * NS <= 36);
    T CC[kOuter ? P : 1][NS];
    if constexpr (kOuter) make_outer<T, N, P>(prm.C, CC);
    T mu[P];                                               // optional per-track observation offset, as in the two-role kernel
#pragma unroll
    for (int j = 0; j < P; ++j) mu[j] = yoff ? yoff[bp * P + j] : T(0);
    bool nonpos = false;                                   // a frame with a non-positive ensemble variance (status bit 2)
    if (issuer) {
      for (int s = 0; s < NST; ++s) mbar_init(&full[s], 1);
      mbar_fence_init();
    }
    tile_barrier(bar_id, bar_cnt);                         // barriers initialised
    if (issuer) {
      for (int s = 0; s < NST; ++s)
        if (s < nb) {
          const long t0 = (long)s * F;
          const uint32_t bytes = (uint32_t)((Tn - t0 < F ? Tn - t0 : F)) * kFrameBytes;
          mbar_expect_tx(&full[s], bytes);
          bulk_g2s(raw + (size_t)s * F * R2 * kLanes, gsrc + t0 * R2 * kLanes, bytes, &full[s]);
        }
    }
    for (long k = 0; k < nb + 2; ++k) {
      const int st = (int)(k % NST);
      if (k < nb) {
        mbar_wait(&full[st], (uint32_t)((k / NST) & 1));
        const T* src = raw + (size_t)st * F * R2 * kLanes + lane;
        T* dst = recb + (k % 3) * (F * REC * kLanes) + lane;
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
            if constexpr (kOuter) fwd_prep_outer<T, N, P>(prm.C, CC, cur, cur + P, rec);
            else fwd_prep<T, N, P, 0>(prm, cur, cur + P, rec);
#pragma unroll
            for (int r = 0; r < REC; ++r) dst[(j * REC + r) * kLanes] = rec[r];
          }
        }
      }
      tile_barrier(bar_id, bar_cnt);                       // end of period k
      if (issuer && k + NST < nb) {
        const long t0 = (k + NST) * F;
        const uint32_t bytes = (uint32_t)((Tn - t0 < F ? Tn - t0 : F)) * kFrameBytes;
        mbar_expect_tx(&full[st], bytes);
        bulk_g2s(raw + (size_t)st * F * R2 * kLanes, gsrc + t0 * R2 * kLanes, bytes, &full[st]);
      }
    }
    if (b < B && status) set_status(status + b, true, nonpos);
    return;
  }

  // prior: both chain warps start from (m0, S0); kPriorOnly starts from (m0, S0^-1) and frame 0 is emitted as is
  T m[N], Nm[NS], rr = T(1);
  bool ok = true;
  {
    const T* pm = m0 + ((shared_mask & EKSB_SHARED_M0) ? 0 : bp * N);
    const T* ps = S0 + ((shared_mask & EKSB_SHARED_S0) ? 0 : bp * (N * N));
#pragma unroll
    for (int i = 0; i < N; ++i) m[i] = pm[i];
    load_sym<T, N>(Nm, ps);
    if (first == kPriorOnly) {
      T Vi[NS];
      sym_inv<T, N>(Nm, Vi, ok);
#pragma unroll
      for (int i = 0; i < NS; ++i) Nm[i] = Vi[i];
    }
  }
  tile_barrier(bar_id, bar_cnt);                           // (barrier init)

  if (role == kRoleCov) {                                  // ---------------- covariance chain
    for (long k = 0; k < nb + 2; ++k) {
      if (k >= 1 && k <= nb) {
        const long kb = k - 1;
        const T* src = recb + (kb % 3) * (F * REC * kLanes) + lane;
        T* dst = vb + (kb & 1) * (F * VR * kLanes) + lane;
        const long rem = Tn - kb * F;
        const int nfr = (int)(rem < F ? rem : F);
        auto frame = [&](int j, bool is_first) {
          if (!(is_first && first == kPriorOnly)) {
            T lam[NS];
#pragma unroll
            for (int r = 0; r < NS; ++r) lam[r] = src[(j * REC + r) * kLanes];
            fwd_cov_scaled<T, N, P>(prm, lam, !(is_first && first == kPriorUpdate), Nm, rr, ok);
          }
#pragma unroll
          for (int r = 0; r < NS; ++r) dst[(j * VR + r) * kLanes] = Nm[r];
          dst[(j * VR + NS) * kLanes] = rr;
        };
        if (kb > 0 && nfr == F) {
#pragma unroll
          for (int j = 0; j < F; ++j) frame(j, false);
        } else {
#pragma unroll
          for (int j = 0; j < F; ++j)
            if (j < nfr) frame(j, kb == 0 && j == 0);
        }
      }
      tile_barrier(bar_id, bar_cnt);
    }
    if (b < B && status) set_status(status + b, ok, false);
    return;
  }

  // ---------------- mean chain + stores
  T* f = filt + (g * Tn * SG) * kLanes + lane;
  for (long k = 0; k < nb + 2; ++k) {
    if (k >= 2) {
      const long kb = k - 2;
      const T* src = recb + (kb % 3) * (F * REC * kLanes) + lane;
      const T* vsrc = vb + (kb & 1) * (F * VR * kLanes) + lane;
      const long rem = Tn - kb * F;
      const int nfr = (int)(rem < F ? rem : F);
      auto frame = [&](int j, bool is_first) {
        T nm[NS], V[NS];
#pragma unroll
        for (int r = 0; r < NS; ++r) nm[r] = vsrc[(j * VR + r) * kLanes];
        const T r1 = vsrc[(j * VR + NS) * kLanes];
        if (!(is_first && first == kPriorOnly)) {
          T rec[REC];
#pragma unroll
          for (int r = 0; r < REC; ++r) rec[r] = src[(j * REC + r) * kLanes];
          fwd_mean_scaled<T, N, P>(prm, rec, !(is_first && first == kPriorUpdate), nm, r1, m, V);
        } else {
#pragma unroll
          for (int r = 0; r < NS; ++r) V[r] = nm[r] * r1;
        }
        T* o = f + j * (SG * kLanes);
#pragma unroll
        for (int i = 0; i < N; ++i) o[i * kLanes] = m[i];
#pragma unroll
        for (int i = 0; i < NS; ++i) o[(N + i) * kLanes] = V[i];
      };
      if (kb > 0 && nfr == F) {
#pragma unroll
        for (int j = 0; j < F; ++j) frame(j, false);
      } else {
#pragma unroll
        for (int j = 0; j < F; ++j)
          if (j < nfr) frame(j, kb == 0 && j == 0);
      }
      f += F * SG * kLanes;
    }
    tile_barrier(bar_id, bar_cnt);
  }
}

// ------------------------------------------------------------------------------------------
// backward:  filt records -> smth records (+ ym, yv); walks u = T-1-t upward.
// Batch k holds frames t_hi = T-1-kF down to t_lo = max(0, T-(k+1)F): one contiguous run in
// HBM; local frame j (u = kF+j) sits at run offset nfr-1-j.
// ------------------------------------------------------------------------------------------
// OP = 0: RTS smoother (A4/A5).  OP = 1: back-substitution + selected inverse of the
// block-tridiagonal system from the (z_t, M_t^-1) records (btd_bwd_prep); same chain.
// OUT (compile time, so that unused projections cost no issue slots on the chain warp):
// bit 0 = write ym = C ms, bit 1 = write yv = diag(C Vs C^T).
template <typename T, int N, int P, int AM, int OP, int OUT, typename Cfg>
__global__ void __launch_bounds__(Cfg::THREADS)
rts_backward_pipe_kernel(const T* filt, const T* __restrict__ C, const T* __restrict__ A, const T* __restrict__ Q,
                         int shared_mask, long B, long Tn_full, T* smth, T* __restrict__ ym, T* __restrict__ yv,
                         int* __restrict__ status, int tpc, const __grid_constant__ RoleMap roles,
                         const __grid_constant__ VTiles vt) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, REC = Rec<N>::BWD;
  constexpr int H = Cfg::H, F = Cfg::F, NST = Cfg::NST;
  const int TPC = tpc;
  using SM = PipeSmem<T, SG, REC, Cfg>;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile_in_cta = roles.tile[warp];
  if (tile_in_cta < 0) return;
  const bool is_chain = roles.h[warp] < 0;
  const int h = is_chain ? 0 : roles.h[warp];
  // virtual tiles (eksb_scan.cuh): slot v = (tile g, chunk c) smooths frames [t_lo, t_hi] of tile g downwards; below
  // the track's last chunk its first frame t_hi = w_c is GIVEN (exact smoothed state from the warm-up kernel, in vinit)
  const bool virt = vt.nc > 1;
  if (virt && *vt.gate != 0) return;
  const long v = (long)blockIdx.x * TPC + tile_in_cta;
  if (v >= num_tiles_dev(B) * (virt ? vt.nc : 1)) return;
  long g = v, t_lo = 0, t_hi = Tn_full - 1;
  const T* vinit = nullptr;
  if (virt) {
    g = v / vt.nc;
    const long c = v - g * vt.nc;
    if (c > 0) t_lo = c * vt.L - vt.cut[g * vt.nc + c - 1] + 1;
    if (c + 1 < vt.nc) {
      t_hi = (c + 1) * vt.L - vt.cut[g * vt.nc + c];
      vinit = static_cast<const T*>(vt.init) + ((g * vt.nc + c) * SG) * kLanes + lane;
    }
  }
  const long Tn = t_hi - t_lo + 1;                         // frames of THIS (virtual) tile
  unsigned char* tile_smem = smem_raw + (size_t)tile_in_cta * SM::kTile;
  T* raw = reinterpret_cast<T*>(tile_smem);
  T* recb = reinterpret_cast<T*>(tile_smem + SM::kRaw);
  uint64_t* full = reinterpret_cast<uint64_t*>(tile_smem + SM::kRaw + SM::kRec);
  const int bar_id = tile_in_cta + 1, bar_cnt = (H + roles.chain_warps) * kLanes;
  const long b = g * kLanes + lane;
  const long bp = b < B ? b : B - 1;
  Params<T, N, P, AM> prm;
  load_track_params<T, N, P, AM>(prm, C, A, Q, shared_mask, bp);
  const long nb = (Tn + F - 1) / F;
  const T* gsrc = filt + ((g * Tn_full + t_lo) * SG) * kLanes;
  constexpr uint32_t kFrameBytes = SG * kLanes * sizeof(T);
  bool ok = true;

  if (!is_chain) {
    const bool issuer = (h == 0 && lane == 0);
    BtdConst<T, N> bc;
    if (OP == 1) btd_make_const<T, N, P, AM>(prm, bc, ok);
    if (issuer) {
      for (int s = 0; s < NST; ++s) mbar_init(&full[s], 1);
      mbar_fence_init();
    }
    tile_barrier(bar_id, bar_cnt);
    if (issuer) {
      for (int s = 0; s < NST; ++s) {
        if (s < nb) {
          const long rem = Tn - (long)s * F;
          const long nfr = rem < F ? rem : F;
          const long t_lo = rem - nfr;
          mbar_expect_tx(&full[s], (uint32_t)nfr * kFrameBytes);
          bulk_g2s(raw + (size_t)s * F * SG * kLanes, gsrc + t_lo * SG * kLanes, (uint32_t)nfr * kFrameBytes, &full[s]);
        }
      }
    }
    for (long k = 0; k <= nb; ++k) {
      const int st = (int)(k % NST);
      if (k < nb) {
        mbar_wait(&full[st], (uint32_t)((k / NST) & 1));
        const long rem = Tn - k * F;
        const int nfr = (int)(rem < F ? rem : F);
        const T* src = raw + (size_t)st * F * SG * kLanes + lane;
        T* dst = recb + (k & 1) * (F * REC * kLanes) + lane;
        // one frame: filtered record at run offset `off` -> smoothing element j of this batch
        auto prep = [&](int j, int off, bool terminal) {
          T cur[SG], rec[REC];
#pragma unroll
          for (int r = 0; r < SG; ++r) cur[r] = src[(off * SG + r) * kLanes];
          if constexpr (OP == 1) {
            btd_bwd_prep<T, N>(bc, cur, cur + N, terminal, rec);
          } else {
            if (terminal) {                                // last frame: ms = mf, Vs = Vf  (:152-153); or, in a
                                                           // virtual tile below the last chunk, the given state
#pragma unroll
              for (int r = 0; r < N * N; ++r) rec[r] = T(0);
#pragma unroll
              for (int r = 0; r < NS; ++r) rec[N * N + r] = vinit ? vinit[(N + r) * kLanes] : cur[N + r];
#pragma unroll
              for (int r = 0; r < N; ++r) rec[N * N + NS + r] = vinit ? vinit[r * kLanes] : cur[r];
            } else {
              bwd_prep<T, N, P, AM>(prm, cur, cur + N, rec, ok);
            }
          }
#pragma unroll
          for (int r = 0; r < REC; ++r) dst[(j * REC + r) * kLanes] = rec[r];
        };
        if (k > 0 && nfr == F) {                           // interior batch: no per-frame guards
#pragma unroll
          for (int jj = 0; jj < (F + H - 1) / H; ++jj) {
            const int j = jj * H + h;
            if ((jj + 1) * H <= F || j < F) prep(j, F - 1 - j, false);
          }
        } else {
#pragma unroll
          for (int jj = 0; jj < (F + H - 1) / H; ++jj) {
            const int j = jj * H + h;
            if (j < nfr) prep(j, nfr - 1 - j, k == 0 && j == 0);
          }
        }
      }
      tile_barrier(bar_id, bar_cnt);
      if (issuer && k + NST < nb) {
        const long rem = Tn - (k + NST) * F;
        const long nfr = rem < F ? rem : F;
        const long t_lo = rem - nfr;
        mbar_expect_tx(&full[st], (uint32_t)nfr * kFrameBytes);
        bulk_g2s(raw + (size_t)st * F * SG * kLanes, gsrc + t_lo * SG * kLanes, (uint32_t)nfr * kFrameBytes, &full[st]);
      }
    }
    if (b < B && status) set_status(status + b, ok, false);
    return;
  }

  T ms[N], Vs[NS];
#pragma unroll
  for (int i = 0; i < N; ++i) ms[i] = T(0);
#pragma unroll
  for (int i = 0; i < NS; ++i) Vs[i] = T(0);
  constexpr bool kYm = (OUT & 1) != 0, kYv = (OUT & 2) != 0;
  // running output pointers at frame t = T-1 (the first frame the chain emits); they walk down
  T* so = smth + ((g * Tn_full + t_hi) * SG) * kLanes + lane;
  T* ymo = kYm ? ym + ((g * Tn_full + t_hi) * P) * kLanes + lane : nullptr;
  T* yvo = kYv ? yv + ((g * Tn_full + t_hi) * P) * kLanes + lane : nullptr;
  // one frame of the chain: record j of the current batch -> frame (base pointers) - j
  auto frame = [&](const T* src, int j) {
    T rec[REC];
#pragma unroll
    for (int r = 0; r < REC; ++r) rec[r] = src[(j * REC + r) * kLanes];
    bwd_chain<T, N>(rec, ms, Vs);
    T* o = so - (long)j * (SG * kLanes);
#pragma unroll
    for (int i = 0; i < N; ++i) o[i * kLanes] = ms[i];
#pragma unroll
    for (int i = 0; i < NS; ++i) o[(N + i) * kLanes] = Vs[i];
    if (kYm || kYv) {
#pragma unroll
      for (int jj = 0; jj < P; ++jj) {
        T s = T(0), q = T(0);
#pragma unroll
        for (int i = 0; i < N; ++i) {
          s += prm.C[jj][i] * ms[i];
          if (kYv) {
            T a = T(0);
#pragma unroll
            for (int k2 = 0; k2 < N; ++k2) a += Vs[sym_idx<N>(i, k2)] * prm.C[jj][k2];
            q += prm.C[jj][i] * a;
          }
        }
        if (kYm) ymo[(jj - j * P) * kLanes] = s;
        if (kYv) yvo[(jj - j * P) * kLanes] = q;
      }
    }
  };
  tile_barrier(bar_id, bar_cnt);                           // (barrier init)
  tile_barrier(bar_id, bar_cnt);                           // batch 0 ready
  const long nfull = Tn / F;                               // batches with all F frames: no per-frame guards
  const int role = roles.h[warp];
  if (role == kRoleBwdCov) {                               // covariance recursion (+ yv) only
    auto cframe = [&](const T* src, int j) {
      T rec[N * N + NS];
#pragma unroll
      for (int r = 0; r < N * N + NS; ++r) rec[r] = src[(j * REC + r) * kLanes];
      bwd_chain_cov<T, N>(rec, Vs);
      T* o = so - (long)j * (SG * kLanes);
#pragma unroll
      for (int i = 0; i < NS; ++i) o[(N + i) * kLanes] = Vs[i];
      if (kYv) {
#pragma unroll
        for (int jj = 0; jj < P; ++jj) {
          T q = T(0);
#pragma unroll
          for (int i = 0; i < N; ++i) {
            T a = T(0);
#pragma unroll
            for (int k2 = 0; k2 < N; ++k2) a += Vs[sym_idx<N>(i, k2)] * prm.C[jj][k2];
            q += prm.C[jj][i] * a;
          }
          yvo[(jj - j * P) * kLanes] = q;
        }
      }
    };
    for (long k = 0; k < nb; ++k) {
      const T* src = recb + (k & 1) * (F * REC * kLanes) + lane;
      const int nfr = k < nfull ? F : (int)(Tn - k * F);
#pragma unroll
      for (int j = 0; j < F; ++j)
        if (j < nfr) cframe(src, j);
      so -= F * SG * kLanes;
      if (kYv) yvo -= F * P * kLanes;
      tile_barrier(bar_id, bar_cnt);
    }
    return;
  }
  if (role == kRoleBwdMean) {                              // mean recursion (+ ym) only
    auto mframe = [&](const T* src, int j) {
      T rec[REC];
#pragma unroll
      for (int r = 0; r < N * N; ++r) rec[r] = src[(j * REC + r) * kLanes];
#pragma unroll
      for (int r = 0; r < N; ++r) rec[N * N + NS + r] = src[(j * REC + N * N + NS + r) * kLanes];
      bwd_chain_mean<T, N>(rec, ms);
      T* o = so - (long)j * (SG * kLanes);
#pragma unroll
      for (int i = 0; i < N; ++i) o[i * kLanes] = ms[i];
      if (kYm) {
#pragma unroll
        for (int jj = 0; jj < P; ++jj) {
          T s2 = T(0);
#pragma unroll
          for (int i = 0; i < N; ++i) s2 += prm.C[jj][i] * ms[i];
          ymo[(jj - j * P) * kLanes] = s2;
        }
      }
    };
    for (long k = 0; k < nb; ++k) {
      const T* src = recb + (k & 1) * (F * REC * kLanes) + lane;
      const int nfr = k < nfull ? F : (int)(Tn - k * F);
#pragma unroll
      for (int j = 0; j < F; ++j)
        if (j < nfr) mframe(src, j);
      so -= F * SG * kLanes;
      if (kYm) ymo -= F * P * kLanes;
      tile_barrier(bar_id, bar_cnt);
    }
    return;
  }
  for (long k = 0; k < nb; ++k) {
    const T* src = recb + (k & 1) * (F * REC * kLanes) + lane;
    if (k < nfull) {
#pragma unroll
      for (int j = 0; j < F; ++j) frame(src, j);
    } else {
      const int nfr = (int)(Tn - k * F);
#pragma unroll
      for (int j = 0; j < F; ++j)
        if (j < nfr) frame(src, j);
    }
    so -= F * SG * kLanes;
    if (kYm) ymo -= F * P * kLanes;
    if (kYv) yvo -= F * P * kLanes;
    tile_barrier(bar_id, bar_cnt);
  }
}

}  // namespace eksb

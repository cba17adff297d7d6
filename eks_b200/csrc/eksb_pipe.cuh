// eksb_pipe.cuh -- warp-specialised forward-filter / RTS kernels for the latency-bound regime
// (few thousand tracks x very long T, e.g. 16 384 x 100 000: only ~3.5 track tiles per SM).
//
// A sequential recursion exposes one warp per 32 tracks; a single warp issues ~1 dependent
// instruction every 4 cycles, so the plain kernels (eksb_kf.cuh) run at the *instruction
// latency* of the whole frame program.  Here each track tile gets 1 "chain" warp + H "helper"
// warps.  Helpers stream the frame records from HBM (coalesced, register prefetch ring), do
// every state-independent part of the frame (eksb_math.cuh: fwd_prep / bwd_prep) and hand a
// compact record to the chain warp through a double-buffered shared-memory ring; the chain warp
// executes only the true dependency chain (fwd_chain / bwd_chain) and writes the results.
// Hand-off is one named barrier (bar.sync id, (H+1)*32) per batch of F frames, per tile, so
// tiles sharing a CTA do not couple.  TPC tiles per CTA puts one chain warp on each of the
// SM's 4 warp schedulers.
#pragma once
#include "eksb_common.cuh"

namespace eksb {

__device__ __forceinline__ void tile_barrier(int id, int count) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}

template <int N_, int P_, int H_, int PFH_, int ROUNDS_, int TPC_> struct PipeCfg {
  static constexpr int H = H_, PFH = PFH_, ROUNDS = ROUNDS_, TPC = TPC_;
  static constexpr int F = H * PFH * ROUNDS;              // frames per hand-off batch
  static constexpr int THREADS = TPC * (H + 1) * kLanes;
};

// ------------------------------------------------------------------------------------------
// forward:  obs records -> filt records
// ------------------------------------------------------------------------------------------
template <typename T, int N, int P, int AM, typename Cfg>
__global__ void __launch_bounds__(Cfg::THREADS)
kf_forward_pipe_kernel(const T* __restrict__ obs, const T* __restrict__ m0, const T* __restrict__ S0,
                       const T* __restrict__ C, const T* __restrict__ A, const T* __restrict__ Q, int shared_mask,
                       long B, long Tn, int first, T* __restrict__ filt, int* __restrict__ status) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, R2 = 2 * P, REC = Rec<N>::FWD;
  constexpr int H = Cfg::H, PFH = Cfg::PFH, ROUNDS = Cfg::ROUNDS, TPC = Cfg::TPC, F = Cfg::F;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool is_chain = warp < TPC;
  const int tile_in_cta = is_chain ? warp : (warp - TPC) % TPC;
  const int h = is_chain ? 0 : (warp - TPC) / TPC;
  const long g = (long)blockIdx.x * TPC + tile_in_cta;
  if (g >= num_tiles_dev(B)) return;                       // chain + helpers of a tile leave together
  T* buf = reinterpret_cast<T*>(smem_raw) + (size_t)tile_in_cta * (2 * F * REC * kLanes) + lane;
  const int bar_id = tile_in_cta + 1, bar_cnt = (H + 1) * kLanes;
  const long b = g * kLanes + lane;
  const long bp = b < B ? b : B - 1;
  Params<T, N, P, AM> prm;
  load_track_params<T, N, P, AM>(prm, C, A, Q, shared_mask, bp);
  const long nb = (Tn + F - 1) / F;

  if (!is_chain) {
    // ---------------- helper: HBM -> registers -> info records in shared memory
    const T* o = obs + (g * Tn * R2) * kLanes + lane;
    T ring[PFH][R2];
#pragma unroll
    for (int i = 0; i < PFH; ++i) {
      long t = h + (long)i * H;
      if (t > Tn - 1) t = Tn - 1;
#pragma unroll
      for (int r = 0; r < R2; ++r) ring[i][r] = o[(t * R2 + r) * kLanes];
    }
    for (long k = 0; k <= nb; ++k) {
      if (k < nb) {
        T* dst = buf + (k & 1) * (F * REC * kLanes);
#pragma unroll
        for (int rd = 0; rd < ROUNDS; ++rd) {
#pragma unroll
          for (int i = 0; i < PFH; ++i) {
            const int j = (rd * PFH + i) * H + h;
            const long t = k * F + j;
            if (t < Tn) {
              T cur[R2];
#pragma unroll
              for (int r = 0; r < R2; ++r) cur[r] = ring[i][r];
              long tn = t + (long)PFH * H;
              if (tn > Tn - 1) tn = Tn - 1;
#pragma unroll
              for (int r = 0; r < R2; ++r) ring[i][r] = o[(tn * R2 + r) * kLanes];
              T rec[REC];
              fwd_prep<T, N, P, AM>(prm, cur, cur + P, rec);
#pragma unroll
              for (int r = 0; r < REC; ++r) dst[(j * REC + r) * kLanes] = rec[r];
            }
          }
        }
      }
      tile_barrier(bar_id, bar_cnt);                       // batch k published / batch k-1 released
    }
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
  bool ok = true;
  if (first == kPriorOnly) {
    T Vi[NS];
    sym_inv<T, N>(V, Vi, ok);
#pragma unroll
    for (int i = 0; i < NS; ++i) V[i] = Vi[i];
  }
  T* f = filt + (g * Tn * SG) * kLanes + lane;
  tile_barrier(bar_id, bar_cnt);                           // batch 0 ready
  for (long k = 0; k < nb; ++k) {
    const T* src = buf + (k & 1) * (F * REC * kLanes);
#pragma unroll
    for (int j = 0; j < F; ++j) {
      const long t = k * F + j;
      if (t < Tn) {
        T rec[REC];
#pragma unroll
        for (int r = 0; r < REC; ++r) rec[r] = src[(j * REC + r) * kLanes];
        if (!(t == 0 && first == kPriorOnly))
          fwd_chain<T, N, P, AM>(prm, rec, !(t == 0 && first == kPriorUpdate), m, V, ok);
        T* o = f + t * SG * kLanes;
#pragma unroll
        for (int i = 0; i < N; ++i) o[i * kLanes] = m[i];
#pragma unroll
        for (int i = 0; i < NS; ++i) o[(N + i) * kLanes] = V[i];
      }
    }
    tile_barrier(bar_id, bar_cnt);                         // batch k consumed / batch k+1 ready
  }
  if (b < B && status && !ok) status[b] = 1;
}

// ------------------------------------------------------------------------------------------
// backward:  filt records -> smth records (+ ym, yv); walks u = T-1-t upward
// ------------------------------------------------------------------------------------------
template <typename T, int N, int P, int AM, typename Cfg>
__global__ void __launch_bounds__(Cfg::THREADS)
rts_backward_pipe_kernel(const T* filt, const T* __restrict__ C, const T* __restrict__ A, const T* __restrict__ Q,
                         int shared_mask, long B, long Tn, T* smth, T* __restrict__ ym, T* __restrict__ yv,
                         int* __restrict__ status) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, REC = Rec<N>::BWD;
  constexpr int H = Cfg::H, PFH = Cfg::PFH, ROUNDS = Cfg::ROUNDS, TPC = Cfg::TPC, F = Cfg::F;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool is_chain = warp < TPC;
  const int tile_in_cta = is_chain ? warp : (warp - TPC) % TPC;
  const int h = is_chain ? 0 : (warp - TPC) / TPC;
  const long g = (long)blockIdx.x * TPC + tile_in_cta;
  if (g >= num_tiles_dev(B)) return;
  T* buf = reinterpret_cast<T*>(smem_raw) + (size_t)tile_in_cta * (2 * F * REC * kLanes) + lane;
  const int bar_id = tile_in_cta + 1, bar_cnt = (H + 1) * kLanes;
  const long b = g * kLanes + lane;
  const long bp = b < B ? b : B - 1;
  Params<T, N, P, AM> prm;
  load_track_params<T, N, P, AM>(prm, C, A, Q, shared_mask, bp);
  const long nb = (Tn + F - 1) / F;
  bool ok = true;

  if (!is_chain) {
    const T* fi = filt + (g * Tn * SG) * kLanes + lane;
    T ring[PFH][SG];
#pragma unroll
    for (int i = 0; i < PFH; ++i) {
      long u = h + (long)i * H;
      if (u > Tn - 1) u = Tn - 1;
      const long t = Tn - 1 - u;
#pragma unroll
      for (int r = 0; r < SG; ++r) ring[i][r] = fi[(t * SG + r) * kLanes];
    }
    for (long k = 0; k <= nb; ++k) {
      if (k < nb) {
        T* dst = buf + (k & 1) * (F * REC * kLanes);
#pragma unroll
        for (int rd = 0; rd < ROUNDS; ++rd) {
#pragma unroll
          for (int i = 0; i < PFH; ++i) {
            const int j = (rd * PFH + i) * H + h;
            const long u = k * F + j;
            if (u < Tn) {
              T cur[SG];
#pragma unroll
              for (int r = 0; r < SG; ++r) cur[r] = ring[i][r];
              long un = u + (long)PFH * H;
              if (un > Tn - 1) un = Tn - 1;
              const long tn = Tn - 1 - un;
#pragma unroll
              for (int r = 0; r < SG; ++r) ring[i][r] = fi[(tn * SG + r) * kLanes];
              T rec[REC];
              if (u == 0) {                                // last frame: ms = mf, Vs = Vf  (:152-153)
#pragma unroll
                for (int r = 0; r < N * N; ++r) rec[r] = T(0);
#pragma unroll
                for (int r = 0; r < NS; ++r) rec[N * N + r] = cur[N + r];
#pragma unroll
                for (int r = 0; r < N; ++r) rec[N * N + NS + r] = cur[r];
              } else {
                bwd_prep<T, N, P, AM>(prm, cur, cur + N, rec, ok);
              }
#pragma unroll
              for (int r = 0; r < REC; ++r) dst[(j * REC + r) * kLanes] = rec[r];
            }
          }
        }
      }
      tile_barrier(bar_id, bar_cnt);
    }
    if (b < B && status && !ok) status[b] = 1;
    return;
  }

  T ms[N], Vs[NS];
#pragma unroll
  for (int i = 0; i < N; ++i) ms[i] = T(0);
#pragma unroll
  for (int i = 0; i < NS; ++i) Vs[i] = T(0);
  T* so = smth + (g * Tn * SG) * kLanes + lane;
  T* ymo = ym ? ym + (g * Tn * P) * kLanes + lane : nullptr;
  T* yvo = yv ? yv + (g * Tn * P) * kLanes + lane : nullptr;
  tile_barrier(bar_id, bar_cnt);
  for (long k = 0; k < nb; ++k) {
    const T* src = buf + (k & 1) * (F * REC * kLanes);
#pragma unroll
    for (int j = 0; j < F; ++j) {
      const long u = k * F + j;
      if (u < Tn) {
        const long t = Tn - 1 - u;
        T rec[REC];
#pragma unroll
        for (int r = 0; r < REC; ++r) rec[r] = src[(j * REC + r) * kLanes];
        bwd_chain<T, N>(rec, ms, Vs);
        T* o = so + t * SG * kLanes;
#pragma unroll
        for (int i = 0; i < N; ++i) o[i * kLanes] = ms[i];
#pragma unroll
        for (int i = 0; i < NS; ++i) o[(N + i) * kLanes] = Vs[i];
        if (ymo || yvo) {
#pragma unroll
          for (int jj = 0; jj < P; ++jj) {
            T s = T(0), q = T(0);
#pragma unroll
            for (int i = 0; i < N; ++i) {
              s += prm.C[jj][i] * ms[i];
              if (yvo) {
                T a = T(0);
#pragma unroll
                for (int k2 = 0; k2 < N; ++k2) a += Vs[sym_idx<N>(i, k2)] * prm.C[jj][k2];
                q += prm.C[jj][i] * a;
              }
            }
            if (ymo) ymo[(t * P + jj) * kLanes] = s;
            if (yvo) yvo[(t * P + jj) * kLanes] = q;
          }
        }
      }
    }
    tile_barrier(bar_id, bar_cnt);
  }
}

}  // namespace eksb

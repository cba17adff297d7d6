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
//
// Time-split form (round 2, kf_nll_split below): few track tiles and long tracks -- one keypoint of a session under an
// optimiser, a grid search on a handful of tracks -- leave the kernel above with one CTA per tile and a whole track
// per chain warp (100 000 frames = 20 ms per evaluation, whatever the batch).  As in the time-split scan
// (eksb_scan.cuh) each tile is cut into chunks of L frames; scan_fwd_warm_kernel, once per candidate (Q differs), runs
// K_c frames from every chunk start c L until the filter has forgotten where it started (K_c may exceed L) and leaves
// the exact filtered state there; CTA (tile, chunk) starts each candidate's recursion at c L + K_c and sums its terms
// up to the next chunk's start -- every frame once -- into part[candidate][tile][chunk]; a small kernel adds the
// chunks in order.  A candidate that has not forgotten after 4 L frames (tiny process noise) raises its gate: its
// chain warps sit the time-split launch out and a second, plain launch of the same kernel evaluates just those whole.
#pragma once
#include "eksb_pipe.cuh"
#include "eksb_scan.cuh"

namespace eksb {

constexpr int kNllChains = 8, kNllHelpers = 2, kNllF = 4, kNllStages = 3;

// time-split launch: nc > 1 chunks of L frames per tile, blockIdx.x = tile * nc + chunk.  gate (optional, either form):
// candidate s takes part iff (gate[s] != 0) == (gate_want != 0).
template <typename T> struct NllSplit {
  long L = 0, nc = 1;
  const long* fcut = nullptr;      // [n_scale][tiles * nc] warm-up frames of each chunk
  const T* finit = nullptr;        // [n_scale][tiles * nc][SIGMA][32] filtered state of frame c L + K_c - 1
  const int* gate = nullptr;       // [n_scale]
  int gate_want = 0;
  double* part = nullptr;          // [n_scale][tiles * nc][32]
};

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
                   const T* __restrict__ yoff, int sc, NllSplit<T> sp) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, R2 = 2 * P, REC = RecNll<N>::FWD;
  constexpr int H = kNllHelpers, F = kNllF, NST = kNllStages;
  using SM = NllSmem<T, R2, REC>;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  T* raw = reinterpret_cast<T*>(smem_raw);
  T* recb = reinterpret_cast<T*>(smem_raw + SM::kRaw);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + SM::kRaw + SM::kRec);
  const int bar_cnt = (H + sc) * kLanes;
  const long nc = sp.nc > 1 ? sp.nc : 1;
  const long g = blockIdx.x / nc, c = blockIdx.x - g * nc;
  const long tiles = gridDim.x / nc;
  const long b = g * kLanes + lane, bp = b < B ? b : B - 1;
  // Candidate i of this CTA starts its recursion at run_i = c L + K_c (the warm-up kernel left the exact filtered state
  // of the frame before; K_c may exceed L) and sums the terms of [acc_i, te_i): acc_i = the running maximum of the
  // starts of chunks 0..c (the starts need not be monotone; the frames before it belong to an earlier chunk), te_i =
  // the next chunk's acc.  The CTA streams [ts, tE), the union over its candidates.
  long ts = 0, tE = Tn, my_run = 0, my_acc = 0, my_te = Tn;
  bool my_active = true;
  if (nc > 1 || sp.gate) {
    ts = Tn; tE = 0;
    bool any = false;
    for (int i = 0; i < sc; ++i) {
      const long si = (long)blockIdx.y * sc + i;
      const bool valid_i = (n_scale <= 0) ? (si == 0) : (si < n_scale);
      bool act = valid_i && !(sp.gate && ((sp.gate[si] != 0) != (sp.gate_want != 0)));
      long run_i = 0, acc_i = 0, te_i = Tn;
      if (act && nc > 1) {
        const long* cuts = sp.fcut + (si * tiles + g) * nc;
        long pm = 0;
        for (long k = 1; k <= c; ++k) {
          const long v = k * sp.L + cuts[k];
          pm = v > pm ? v : pm;
        }
        run_i = (c == 0) ? 0 : c * sp.L + cuts[c];
        acc_i = pm;
        if (c < nc - 1) {
          const long v = (c + 1) * sp.L + cuts[c + 1];
          te_i = v > pm ? v : pm;
        }
        act = acc_i < te_i;                                // (else nothing of the track is this chunk's to sum)
      }
      if (act) {
        ts = run_i < ts ? run_i : ts;
        tE = te_i > tE ? te_i : tE;
        any = true;
      }
      if (i == warp - H) { my_active = act; my_run = run_i; my_acc = acc_i; my_te = te_i; }
    }
    if (!any) return;                                      // (the whole CTA: none of its candidates is on this launch)
  }
  const long Tl = tE - ts;                                 // frames of this CTA
  Params<T, N, P, 0> prm;
  load_track_params<T, N, P, 0>(prm, C, A, Q, shared_mask, bp);
  const long nb = (Tl + F - 1) / F;
  constexpr uint32_t kFrameBytes = R2 * kLanes * sizeof(T);

  if (warp < H) {                                          // ---------------- helpers: obs frames -> (Lam, iv, yy, ldr)
    const int h = warp;
    const T* gsrc = obs + ((g * Tn + ts) * R2) * kLanes;
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
          const uint32_t bytes = (uint32_t)((Tl - t0 < F ? Tl - t0 : F)) * kFrameBytes;
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
        const long rem = Tl - k * F;
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
            nonpos = nonpos || (any_nonpos<T, P>(cur + P) && !(first == kPriorOnly && ts == 0 && k == 0 && j == 0));
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
        const uint32_t bytes = (uint32_t)((Tl - t0 < F ? Tl - t0 : F)) * kFrameBytes;
        mbar_expect_tx(&full[st], bytes);
        bulk_g2s(raw + (size_t)st * F * R2 * kLanes, gsrc + t0 * R2 * kLanes, bytes, &full[st]);
      }
    }
    if (blockIdx.y == 0 && b < B && status) set_status(status + b, true, nonpos);
    return;
  }

  // ---------------- chain warp: one candidate scale
  long s = (long)blockIdx.y * sc + (warp - H);
  const bool valid = my_active && ((n_scale <= 0) ? (s == 0) : (s < n_scale));
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
    if (valid && my_run > 0) {                             // exact filtered state of frame my_run - 1 (warm-up kernel)
      const T* s0 = sp.finit + (((s * tiles + g) * nc + c) * SG) * kLanes + lane;
#pragma unroll
      for (int i = 0; i < N; ++i) m[i] = s0[i * kLanes];
#pragma unroll
      for (int i = 0; i < NS; ++i) Nm[i] = s0[(N + i) * kLanes];
    } else if (first == kPriorOnly) {                      // eks/newton_eks.py:132 : P starts at inv(S0), frame 0 has no update
      T Vi[NS];
      sym_inv<T, N>(Nm, Vi, ok);
#pragma unroll
      for (int i = 0; i < NS; ++i) Nm[i] = Vi[i];
    }
  }
  double acc = 0.0;
  T xprod = T(1);
  // this candidate's frames as indices into the CTA's: recursion from lo_run, terms summed over [lo_acc, hi)
  const long lo_run = my_run - ts, lo_acc = my_acc - ts, hi = my_te - ts;
  auto frame = [&](const T* src, int j, bool is_first, bool summed) {
    if (is_first && first == kPriorOnly) return;
    T rec[REC], xdet;
#pragma unroll
    for (int r = 0; r < REC; ++r) rec[r] = src[(j * REC + r) * kLanes];
    const T e = fwd_nll_scaled<T, N, P>(prm, rec, !(is_first && first == kPriorUpdate), m, Nm, rr, xdet, ok);
    if (summed) {
      acc += (double)e + (double)rec[NS + N + 1];
      xprod *= xdet;
    }
  };
  tile_barrier(1, bar_cnt);                                // (barrier init)
  tile_barrier(1, bar_cnt);                                // batch 0 ready
  for (long k = 0; k < nb; ++k) {
    const T* src = recb + (k & 1) * (F * REC * kLanes) + lane;
    if (valid && k * F > lo_acc && (k + 1) * F < hi) {
#pragma unroll
      for (int j = 0; j < F; ++j) frame(src, j, false, true);
    } else if (valid) {
#pragma unroll
      for (int j = 0; j < F; ++j) {
        const long f = k * F + j;
        if (f >= lo_run && f < hi) frame(src, j, ts + f == 0, f >= lo_acc);
      }
    }
    acc -= log((double)xprod);                             // one log per F = 4 frames: log det(Id + Lam S) = -log xdet
    xprod = T(1);
    tile_barrier(1, bar_cnt);                              // batch k consumed / batch k+1 ready
  }
  if (valid && b < B) {
    if (nc > 1) sp.part[((s * tiles + g) * nc + c) * kLanes + lane] = acc;
    else nll[s * B + b] = 0.5 * (acc + (double)Tn * P * 1.8378770664093453 /* log 2pi */);
    if (status) set_status(status + b, ok, false);
  }
}

// nll[s][b] = the chunks' sums in order + the constant, for the candidates of the time-split launch
static __global__ void nll_split_reduce_kernel(const double* __restrict__ part, const int* __restrict__ gate, long B, long nc,
                                        int n_scale, double konst, double* __restrict__ nll) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long)n_scale * B) return;
  const long s = i / B, b = i - s * B;
  if (gate[s] != 0) return;
  const long g = b / kLanes, lane = b - g * kLanes, tiles = (B + kLanes - 1) / kLanes;
  const double* pp = part + ((s * tiles + g) * nc) * kLanes + lane;
  double acc = 0.0;
  for (long c = 0; c < nc; ++c) acc += pp[c * kLanes];
  nll[i] = 0.5 * (acc + konst);
}

template <typename T, int N, int P>
static int launch_nll_pipe(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int shared_mask, long B,
                           long Tn, int first, const T* q_scale, int n_scale, double* nll, int* status, const T* yoff,
                           cudaStream_t st, NllSplit<T> sp = NllSplit<T>()) {
  using SM = NllSmem<T, 2 * P, RecNll<N>::FWD>;
  const int ns = (q_scale && n_scale > 0) ? n_scale : 1;
  const int sc = ns < kNllChains ? ns : kNllChains;
  auto kern = kf_nll_pipe_kernel<T, N, P>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SM::kTotal);
  dim3 grid((unsigned)(num_tiles(B) * (sp.nc > 1 ? sp.nc : 1)), (unsigned)((ns + sc - 1) / sc));
  kern<<<grid, (kNllHelpers + sc) * kLanes, SM::kTotal, st>>>(obs, m0, S0, C, A, Q, shared_mask, B, Tn, first, q_scale,
                                                            q_scale ? n_scale : 0, nll, status, yoff, sc, sp);
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

// ---- time-split sweep ------------------------------------------------------------------------------------------
inline bool nll_split_dims_ok(int n, int p) { return (n == 2 && p == 4) || (n == 3 && (p == 4 || p == 6 || p == 8)); }

// Chunks per tile (0 = the plain launch is the route).  A CTA carries up to 8 chain warps, two per FP64 pipe -- about
// what one SM's FP64 issue rate feeds (measured: 16 candidates x 64 tiles, 1 vs 3.5 CTAs per SM: 41 vs 46 ms) -- so
// the target is sm_count * clamp(8 / chains per CTA, 1, 4) CTAs in one wave, chunks of at least 4096 frames (the
// warm-up runs at ~1 us per frame per chunk; K is a few hundred to a few thousand frames), tracks of >= 8192 frames.
// EKSB_NLL_SPLIT=0 switches it off, =N forces N chunks (tests).
inline long nll_split_chunks(long B, long Tn, int n, int p, int n_scale, long* L_out) {
  if (B <= 0 || !nll_split_dims_ok(n, p)) return 0;
  const char* e = getenv("EKSB_NLL_SPLIT");
  long forced = 0;
  if (e && e[0]) {
    forced = atol(e);
    if (forced <= 0) return 0;
  }
  const int ns = n_scale > 0 ? n_scale : 1;
  const int sc = ns < kNllChains ? ns : kNllChains;
  const long ctas = num_tiles(B) * ((ns + sc - 1) / sc);
  long nc = forced;
  if (!forced) {
    if (Tn < 8192) return 0;
    int per_sm = kNllChains / sc;
    per_sm = per_sm < 1 ? 1 : (per_sm > 4 ? 4 : per_sm);
    nc = ((long)sm_count() * per_sm) / ctas;
    const long nc_max = Tn / 4096;
    if (nc > nc_max) nc = nc_max;
  }
  if (nc < 2) return 0;
  long L = (Tn + nc - 1) / nc;
  L = (L + kNllF - 1) / kNllF * kNllF;
  if (L < 8) L = 8;
  nc = (Tn + L - 1) / L;
  if (nc < 2) return 0;
  if (L_out) *L_out = L;
  return nc;
}

struct NllWs {                     // byte offsets into the workspace (256-byte aligned parts)
  size_t gate, fcut, part, finit, total;
};
inline NllWs nll_ws_layout(long B, int n, int n_scale, long nc, int elem_size) {
  const size_t ns = (size_t)(n_scale > 0 ? n_scale : 1), vt = (size_t)num_tiles(B) * (size_t)nc;
  const size_t sg = (size_t)(n + n * (n + 1) / 2);
  auto up = [](size_t x) { return (x + 255) / 256 * 256; };
  NllWs w;
  w.gate = 0;
  w.fcut = up(ns * sizeof(int));
  w.part = w.fcut + up(ns * vt * sizeof(long));
  w.finit = w.part + up(ns * vt * kLanes * sizeof(double));
  w.total = w.finit + up(ns * vt * sg * kLanes * (size_t)elem_size);
  return w;
}
inline size_t nll_workspace_bytes(long B, long Tn, int n, int p, int n_scale, int elem_size) {
  long L = 0;
  const long nc = nll_split_chunks(B, Tn, n, p, n_scale, &L);
  return nc ? nll_ws_layout(B, n, n_scale, nc, elem_size).total : 0;
}

template <typename T, int N, int P>
static int launch_nll_split(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int mask, long B,
                            long Tn, int first, const T* q_scale, int n_scale, double* nll, int* status, const T* yoff,
                            unsigned char* ws, long L, long nc, cudaStream_t st) {
  const int ns = (q_scale && n_scale > 0) ? n_scale : 1;
  const long tiles = num_tiles(B);
  const NllWs w = nll_ws_layout(B, N, ns, nc, (int)sizeof(T));
  NllSplit<T> sp;
  sp.L = L; sp.nc = nc;
  int* gate = reinterpret_cast<int*>(ws + w.gate);
  long* fcut = reinterpret_cast<long*>(ws + w.fcut);
  T* finit = reinterpret_cast<T*>(ws + w.finit);
  sp.fcut = fcut; sp.finit = finit; sp.gate = gate; sp.gate_want = 0;
  sp.part = reinterpret_cast<double*>(ws + w.part);
  cudaMemsetAsync(gate, 0, (size_t)ns * sizeof(int), st);
  cudaMemsetAsync(sp.part, 0, (size_t)ns * tiles * nc * kLanes * sizeof(double), st);   // (a chunk with nothing to sum writes nothing)
  const long kcap = 4 * L;                                 // warm-up frames before a candidate is given up (evaluated whole)
  const char* w1 = getenv("EKSB_SCAN_WARM1");              // A/B switch: the one-warp warm-up kernel
  if ((w1 && w1[0] == '1') || sizeof(T) != 8 || tiles * (nc - 1) * ns > 3L * sm_count()) {   // (float, many chunks: see scan_run_fwd_off)
    dim3 wgrid((unsigned)((tiles * (nc - 1) + kScanWarps - 1) / kScanWarps), (unsigned)ns);
    if (yoff)
      scan_fwd_warm_kernel<T, N, P, 0, true><<<wgrid, kScanWarps * kLanes, 0, st>>>(obs, C, A, Q, mask, B, Tn, L, nc, yoff, fcut, gate,
                                                                                    (T*)nullptr, q_scale, finit, kcap);
    else
      scan_fwd_warm_kernel<T, N, P, 0, false><<<wgrid, kScanWarps * kLanes, 0, st>>>(obs, C, A, Q, mask, B, Tn, L, nc, yoff, fcut, gate,
                                                                                     (T*)nullptr, q_scale, finit, kcap);
  } else {
    dim3 wgrid((unsigned)(tiles * (nc - 1)), (unsigned)ns);  // three-role warm-up: one CTA per (tile, chunk) and candidate
    constexpr size_t wsm = Warm3Smem<T, N>::kTotal;
    if (yoff) {
      cudaFuncSetAttribute(scan_fwd_warm3_kernel<T, N, P, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wsm);
      scan_fwd_warm3_kernel<T, N, P, true><<<wgrid, 3 * kLanes, wsm, st>>>(obs, C, A, Q, mask, B, Tn, L, nc, yoff, fcut, gate, (T*)nullptr, q_scale, finit, kcap);
    } else {
      cudaFuncSetAttribute(scan_fwd_warm3_kernel<T, N, P, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wsm);
      scan_fwd_warm3_kernel<T, N, P, false><<<wgrid, 3 * kLanes, wsm, st>>>(obs, C, A, Q, mask, B, Tn, L, nc, yoff, fcut, gate, (T*)nullptr, q_scale, finit, kcap);
    }
  }
  int rc = launch_nll_pipe<T, N, P>(obs, m0, S0, C, A, Q, mask, B, Tn, first, q_scale, n_scale, nll, status, yoff, st, sp);
  if (rc != EKSB_OK) return rc;
  const long items = (long)ns * B;
  nll_split_reduce_kernel<<<(unsigned)((items + 255) / 256), 256, 0, st>>>(sp.part, gate, B, nc, ns,
                                                                          (double)Tn * P * 1.8378770664093453, nll);
  NllSplit<T> fb;                                          // candidates some chunk of which never forgot: whole tracks
  fb.gate = gate; fb.gate_want = 1;
  return launch_nll_pipe<T, N, P>(obs, m0, S0, C, A, Q, mask, B, Tn, first, q_scale, n_scale, nll, status, yoff, st, fb);
}

// NLL sweep with a workspace: time-split when nll_split_chunks says so and the workspace holds it, else the plain launch.
template <typename T>
int kf_nll_split(const T* obs, const T* yoff, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int mask, long B,
                 long Tn, int n, int p, int first, const T* q_scale, int n_scale, double* nll, int* status, void* ws,
                 size_t ws_bytes, cudaStream_t st) {
  long L = 0;
  const long nc = ws ? nll_split_chunks(B, Tn, n, p, q_scale ? n_scale : 0, &L) : 0;
  if (nc && nll_ws_layout(B, n, q_scale ? n_scale : 0, nc, (int)sizeof(T)).total > ws_bytes) {
    set_error("kf_nll: workspace too small (%zu bytes needed)", nll_ws_layout(B, n, q_scale ? n_scale : 0, nc, (int)sizeof(T)).total);
    return EKSB_EARG;
  }
  if (!nc) return kf_nll_pipe<T>(obs, m0, S0, C, A, Q, mask, B, Tn, n, p, first, q_scale, n_scale, nll, status, yoff, st);
#define X(N_, P_) \
  if (n == N_ && p == P_) return launch_nll_split<T, N_, P_>(obs, m0, S0, C, A, Q, mask, B, Tn, first, q_scale, n_scale, nll, status, yoff, (unsigned char*)ws, L, nc, st);
  X(2, 4) X(3, 4) X(3, 6) X(3, 8)
#undef X
  return EKSB_EDIM;
}

}  // namespace eksb

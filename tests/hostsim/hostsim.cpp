// Host build of the per-thread programs in eks_b200/csrc/eksb_math.cuh.
// TEST INFRASTRUCTURE ONLY: lets the CPU-only build box check the exact code the
// CUDA kernels run (same templates, lane stride 1) against the numpy oracle.
// Never loaded by the eks_b200 package.
#include <vector>
#include <cstring>
#include "../../eks_b200/csrc/eksb_math.cuh"
#include "../../eks_b200/csrc/eksb_dims.h"

using namespace eksb;

template <typename T, int N, int P, int AM>
static int run_case(long Tn, const T* y, const T* ev, const T* m0, const T* S0, const T* C, const T* A, const T* Q,
                    int first, T* filt, T* smth, T* ym, T* yv, T* cv, double* nll) {
  constexpr int NS = Dim<N>::NS;
  Params<T, N, P, AM> prm;
  load_params<T, N, P, AM>(prm, C, A, Q);
  std::vector<T> obs(Tn * 2 * P);
  for (long t = 0; t < Tn; ++t)
    for (int j = 0; j < P; ++j) { obs[t * 2 * P + j] = y[t * P + j]; obs[t * 2 * P + P + j] = ev[t * P + j]; }
  T m[N], V[NS];
  for (int i = 0; i < N; ++i) m[i] = m0[i];
  load_sym<T, N>(V, S0);
  double acc = 0.0;
  bool ok = kf_forward_range<T, N, P, AM, 1, 3, true>(prm, obs.data(), filt, Tn, first, m, V, acc);
  *nll = acc;
  if (smth) {
    T ms[N], Vs[NS];
    ok = rts_backward_range<T, N, P, AM, 1, 3>(prm, filt, smth, ym, yv, cv, Tn, true, ms, Vs) && ok;
  }
  return ok ? 0 : 1;
}

template <typename T>
static int dispatch(int n, int p, int am, long Tn, const T* y, const T* ev, const T* m0, const T* S0, const T* C,
                    const T* A, const T* Q, int first, T* filt, T* smth, T* ym, T* yv, T* cv, double* nll) {
#define X(N_, P_)                                                                                             \
  if (n == N_ && p == P_)                                                                                     \
    return am ? run_case<T, N_, P_, 1>(Tn, y, ev, m0, S0, C, A, Q, first, filt, smth, ym, yv, cv, nll)        \
              : run_case<T, N_, P_, 0>(Tn, y, ev, m0, S0, C, A, Q, first, filt, smth, ym, yv, cv, nll);
  EKSB_FOR_EACH_DIM(X)
#undef X
  return -1;
}

extern "C" int hs_kf_rts_f64(int n, int p, int am, long Tn, const double* y, const double* ev, const double* m0,
                             const double* S0, const double* C, const double* A, const double* Q, int first,
                             double* filt, double* smth, double* ym, double* yv, double* cv, double* nll) {
  return dispatch<double>(n, p, am, Tn, y, ev, m0, S0, C, A, Q, first, filt, smth, ym, yv, cv, nll);
}
extern "C" int hs_kf_rts_f32(int n, int p, int am, long Tn, const float* y, const float* ev, const float* m0,
                             const float* S0, const float* C, const float* A, const float* Q, int first,
                             float* filt, float* smth, float* ym, float* yv, float* cv, double* nll) {
  return dispatch<float>(n, p, am, Tn, y, ev, m0, S0, C, A, Q, first, filt, smth, ym, yv, cv, nll);
}

// ---- NLL in the split form (kf_nll_pipe_kernel): helper record with yy / ldr, scaled chain, one log per 4 frames ----
template <typename T, int N, int P>
static int run_nll_split(long Tn, const T* y, const T* ev, const T* m0, const T* S0, const T* C, const T* A, const T* Q,
                         int first, double* nll_terms) {
  constexpr int NS = Dim<N>::NS;
  Params<T, N, P, 0> prm;
  load_params<T, N, P, 0>(prm, C, A, Q);
  T m[N], Nm[NS], rr = T(1);
  for (int i = 0; i < N; ++i) m[i] = m0[i];
  load_sym<T, N>(Nm, S0);
  bool ok = true;
  if (first == kPriorOnly) {
    T Vi[NS];
    sym_inv<T, N>(Nm, Vi, ok);
    for (int i = 0; i < NS; ++i) Nm[i] = Vi[i];
  }
  T CC[P][NS];
  make_outer<T, N, P>(prm.C, CC);
  double acc = 0.0;
  T xprod = T(1);
  for (long t = 0; t < Tn; ++t) {
    if (t == 0 && first == kPriorOnly) continue;
    T rec[RecNll<N>::FWD], xdet;
    fwd_prep_outer_nll<T, N, P>(prm.C, CC, y + t * P, ev + t * P, rec);
    const T e = fwd_nll_scaled<T, N, P>(prm, rec, !(t == 0 && first == kPriorUpdate), m, Nm, rr, xdet, ok);
    acc += (double)e + (double)rec[NS + N + 1];
    xprod *= xdet;
    if ((t & 3) == 3 || t == Tn - 1) { acc -= log((double)xprod); xprod = T(1); }
  }
  *nll_terms = acc;
  return ok ? 0 : 1;
}

extern "C" int hs_nll_split_f64(int n, int p, long Tn, const double* y, const double* ev, const double* m0, const double* S0,
                                const double* C, const double* A, const double* Q, int first, double* nll_terms) {
#define X(N_, P_) if (n == N_ && p == P_) return run_nll_split<double, N_, P_>(Tn, y, ev, m0, S0, C, A, Q, first, nll_terms);
  X(2, 4) X(3, 4) X(3, 6) X(3, 8)
#undef X
  return -1;
}
extern "C" int hs_nll_split_f32(int n, int p, long Tn, const float* y, const float* ev, const float* m0, const float* S0,
                                const float* C, const float* A, const float* Q, int first, double* nll_terms) {
#define X(N_, P_) if (n == N_ && p == P_) return run_nll_split<float, N_, P_>(Tn, y, ev, m0, S0, C, A, Q, first, nll_terms);
  X(2, 4) X(3, 4) X(3, 6) X(3, 8)
#undef X
  return -1;
}

// ---- producer/consumer split form (what the warp-specialised kernels run) ----
template <typename T, int N, int P, int AM>
static int run_split(long Tn, const T* y, const T* ev, const T* m0, const T* S0, const T* C, const T* A, const T* Q,
                     int first, T* filt, T* smth) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA;
  Params<T, N, P, AM> prm;
  load_params<T, N, P, AM>(prm, C, A, Q);
  T m[N], V[NS], Nm[NS], rr = T(1);
  for (int i = 0; i < N; ++i) m[i] = m0[i];
  load_sym<T, N>(V, S0);
  for (int i = 0; i < NS; ++i) Nm[i] = V[i];
  bool ok = true;
  for (long t = 0; t < Tn; ++t) {
    T rec[Rec<N>::FWD];
    if constexpr (P * NS <= 36) {            // same selection as the kernel's helper warps (kOuter)
      T CC[P][NS];
      make_outer<T, N, P>(prm.C, CC);
      fwd_prep_outer<T, N, P>(prm.C, CC, y + t * P, ev + t * P, rec);
    } else {
      fwd_prep<T, N, P, AM>(prm, y + t * P, ev + t * P, rec);
    }
    if (!(t == 0 && first == kPriorOnly)) {
      if constexpr (AM == 0 && N <= 3) {     // same selection as kf_forward_pipe_kernel (kScaled)
        if (t == 0) { for (int i = 0; i < NS; ++i) Nm[i] = V[i]; rr = T(1); }
        fwd_chain_scaled<T, N, P>(prm, rec, !(t == 0 && first == kPriorUpdate), m, Nm, rr, V, ok);
      } else {
        fwd_chain<T, N, P, AM>(prm, rec, !(t == 0 && first == kPriorUpdate), m, V, ok);
      }
    } else {
      for (int i = 0; i < NS; ++i) Nm[i] = V[i];
      rr = T(1);
    }
    for (int i = 0; i < N; ++i) filt[t * SG + i] = m[i];
    for (int i = 0; i < NS; ++i) filt[t * SG + N + i] = V[i];
  }
  T ms[N], Vs[NS];
  for (long t = Tn - 1; t >= 0; --t) {
    if (t == Tn - 1) {
      for (int i = 0; i < N; ++i) ms[i] = filt[t * SG + i];
      for (int i = 0; i < NS; ++i) Vs[i] = filt[t * SG + N + i];
    } else {
      T rec[Rec<N>::BWD];
      bwd_prep<T, N, P, AM>(prm, filt + t * SG, filt + t * SG + N, rec, ok);
      bwd_chain<T, N>(rec, ms, Vs);
    }
    for (int i = 0; i < N; ++i) smth[t * SG + i] = ms[i];
    for (int i = 0; i < NS; ++i) smth[t * SG + N + i] = Vs[i];
  }
  return ok ? 0 : 1;
}

template <typename T>
static int dispatch_split(int n, int p, int am, long Tn, const T* y, const T* ev, const T* m0, const T* S0, const T* C,
                          const T* A, const T* Q, int first, T* filt, T* smth) {
#define X(N_, P_)                                                                       \
  if (n == N_ && p == P_)                                                               \
    return am ? run_split<T, N_, P_, 1>(Tn, y, ev, m0, S0, C, A, Q, first, filt, smth)  \
              : run_split<T, N_, P_, 0>(Tn, y, ev, m0, S0, C, A, Q, first, filt, smth);
  EKSB_FOR_EACH_DIM(X)
#undef X
  return -1;
}

extern "C" int hs_split_f64(int n, int p, int am, long Tn, const double* y, const double* ev, const double* m0,
                            const double* S0, const double* C, const double* A, const double* Q, int first,
                            double* filt, double* smth) {
  return dispatch_split<double>(n, p, am, Tn, y, ev, m0, S0, C, A, Q, first, filt, smth);
}
extern "C" int hs_split_f32(int n, int p, int am, long Tn, const float* y, const float* ev, const float* m0,
                            const float* S0, const float* C, const float* A, const float* Q, int first, float* filt,
                            float* smth) {
  return dispatch_split<float>(n, p, am, Tn, y, ev, m0, S0, C, A, Q, first, filt, smth);
}

// ---- block-tridiagonal solve through the same prep/chain pieces the kernels use ----
template <typename T, int N, int P, int AM>
static int run_btd(long Tn, const T* y, const T* ev, const T* m0, const T* S0, const T* C, const T* A, const T* Q,
                   T* work, T* sol) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA;
  Params<T, N, P, AM> prm;
  load_params<T, N, P, AM>(prm, C, A, Q);
  bool ok = true;
  BtdConst<T, N> c;
  btd_make_const<T, N, P, AM>(prm, c, ok);
  T S0p[NS], iS0[NS], iS0m0[N];
  load_sym<T, N>(S0p, S0);
  sym_inv<T, N>(S0p, iS0, ok);
  sym_matvec<T, N>(iS0, m0, iS0m0);
  T z[N], iM[NS];
  for (int i = 0; i < N; ++i) z[i] = T(0);
  for (int i = 0; i < NS; ++i) iM[i] = T(0);
  for (long t = 0; t < Tn; ++t) {
    T rec[Rec<N>::FWD];
    fwd_prep<T, N, P, AM>(prm, y + t * P, ev + t * P, rec);
    btd_fwd_chain<T, N>(c, rec, t == 0, t == Tn - 1, iS0, iS0m0, z, iM, ok);
    for (int i = 0; i < N; ++i) work[t * SG + i] = z[i];
    for (int i = 0; i < NS; ++i) work[t * SG + N + i] = iM[i];
  }
  T x[N], Sg[NS];
  for (int i = 0; i < N; ++i) x[i] = T(0);
  for (int i = 0; i < NS; ++i) Sg[i] = T(0);
  for (long t = Tn - 1; t >= 0; --t) {
    T rec[Rec<N>::BWD];
    btd_bwd_prep<T, N>(c, work + t * SG, work + t * SG + N, t == Tn - 1, rec);
    bwd_chain<T, N>(rec, x, Sg);
    for (int i = 0; i < N; ++i) sol[t * SG + i] = x[i];
    for (int i = 0; i < NS; ++i) sol[t * SG + N + i] = Sg[i];
  }
  return ok ? 0 : 1;
}

template <typename T>
static int dispatch_btd(int n, int p, int am, long Tn, const T* y, const T* ev, const T* m0, const T* S0, const T* C,
                        const T* A, const T* Q, T* work, T* sol) {
#define X(N_, P_)                                                                  \
  if (n == N_ && p == P_)                                                          \
    return am ? run_btd<T, N_, P_, 1>(Tn, y, ev, m0, S0, C, A, Q, work, sol)        \
              : run_btd<T, N_, P_, 0>(Tn, y, ev, m0, S0, C, A, Q, work, sol);
  EKSB_FOR_EACH_DIM(X)
#undef X
  return -1;
}

extern "C" int hs_btd_f64(int n, int p, int am, long Tn, const double* y, const double* ev, const double* m0,
                          const double* S0, const double* C, const double* A, const double* Q, double* work,
                          double* sol) {
  return dispatch_btd<double>(n, p, am, Tn, y, ev, m0, S0, C, A, Q, work, sol);
}
extern "C" int hs_btd_f32(int n, int p, int am, long Tn, const float* y, const float* ev, const float* m0,
                          const float* S0, const float* C, const float* A, const float* Q, float* work, float* sol) {
  return dispatch_btd<float>(n, p, am, Tn, y, ev, m0, S0, C, A, Q, work, sol);
}

// ---- chunked parallel-in-time scan, phases run one after another on the host ----
// Same chunk programs as the device kernels (scan_fwd_chunk / scan_bwd_chunk, lane stride 1), same
// hand-over through the cut positions; `early` = 0 switches the early write-out off (full re-run).
// cuts (optional, 2*nc longs) receives the forward and backward cut of every chunk.
template <typename T, int N, int P, int AM>
static int run_scan(long Tn, long L, int early, const T* y, const T* ev, const T* m0, const T* S0, const T* C,
                    const T* A, const T* Q, T* filt, T* smth, long* cuts) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, R2 = 2 * P;
  Params<T, N, P, AM> prm;
  load_params<T, N, P, AM>(prm, C, A, Q);
  bool ok = true;
  const long nc = (Tn + L - 1) / L;
  std::vector<T> obs(Tn * R2);
  for (long t = 0; t < Tn; ++t)
    for (int j = 0; j < P; ++j) { obs[t * R2 + j] = y[t * P + j]; obs[t * R2 + P + j] = ev[t * P + j]; }
  std::vector<T> fsum(nc * ScanRec<N>::FWD), finit(nc * SG), bsum(nc * ScanRec<N>::BWD), binit(nc * SG);
  std::vector<long> fcut(nc, 0), bcut(nc, 0);
  auto len = [&](long c) { return ((c + 1) * L < Tn ? (c + 1) * L : Tn) - c * L; };
  double nll = 0.0;
  // F1: chunk 0 from the prior; summaries (+ early write-out) of chunks 1..nc-1
  T m[N], V[NS];
  for (int i = 0; i < N; ++i) m[i] = m0[i];
  load_sym<T, N>(V, S0);
  ok = kf_forward_range<T, N, P, AM, 1, 2, false>(prm, obs.data(), filt, len(0), kPriorUpdate, m, V, nll) && ok;
  for (long c = 1; c < nc; ++c)
    fcut[c] = scan_fwd_chunk<T, N, P, AM, 1>(prm, obs.data() + c * L * R2, filt + c * L * SG, len(c), early != 0,
                                             &fsum[c * ScanRec<N>::FWD], ok, FwdTailPlain<T, N, P, AM, 1, 2>());
  // F2: fold -- skipped (as on the device) when every chunk wrote its own last frame
  bool f_alive = false;
  for (long c = 1; c < nc; ++c) f_alive = f_alive || fcut[c] >= len(c);
  if (f_alive)
    for (long c = 1; c < nc; ++c) {
      for (int i = 0; i < N; ++i) finit[c * SG + i] = m[i];
      for (int i = 0; i < NS; ++i) finit[c * SG + N + i] = V[i];
      fwd_apply_summary<T, N>(&fsum[c * ScanRec<N>::FWD], m, V, ok);
    }
  // F3: re-run of the frames before each chunk's cut, from the last frame of the previous chunk if that
  // chunk wrote it itself, else from the fold's boundary state
  for (long c = 1; c < nc; ++c) {
    if (fcut[c] <= 0) continue;
    const bool prev_final = (c == 1) || fcut[c - 1] < L;
    const T* src = prev_final ? filt + (c * L - 1) * SG : &finit[c * SG];
    T mm[N], VV[NS];
    for (int i = 0; i < N; ++i) mm[i] = src[i];
    for (int i = 0; i < NS; ++i) VV[i] = src[N + i];
    ok = kf_forward_range<T, N, P, AM, 1, 2, false>(prm, obs.data() + c * L * R2, filt + c * L * SG, fcut[c], kContinue,
                                                    mm, VV, nll) && ok;
  }
  // B1: terminal chunk directly; summaries (+ early write-out) of chunks 0..nc-2.  Out of place here
  // (smth != filt); the in-place aliasing is covered by the GPU tests.
  T ms[N], Vs[NS];
  ok = rts_backward_range<T, N, P, AM, 1, 2>(prm, filt + (nc - 1) * L * SG, smth + (nc - 1) * L * SG, nullptr, nullptr,
                                             nullptr, len(nc - 1), true, ms, Vs) && ok;
  for (long c = 0; c + 1 < nc; ++c)
    bcut[c] = scan_bwd_chunk<T, N, P, AM, 1>(prm, filt + c * L * SG, smth + c * L * SG, nullptr, L, early != 0,
                                             &bsum[c * ScanRec<N>::BWD], ok, BwdTailPlain<T, N, P, AM, 1, 2>());
  // B2: fold -- skipped when every chunk smoothed its own first frame
  bool b_alive = false;
  for (long c = 0; c + 1 < nc; ++c) b_alive = b_alive || bcut[c] <= 0;
  if (b_alive)
    for (long c = nc - 2; c >= 0; --c) {
      for (int i = 0; i < N; ++i) binit[c * SG + i] = ms[i];
      for (int i = 0; i < NS; ++i) binit[c * SG + N + i] = Vs[i];
      bwd_chain<T, N>(&bsum[c * ScanRec<N>::BWD], ms, Vs);
    }
  // B3: re-run of the frames from each chunk's cut on
  for (long c = nc - 2; c >= 0; --c) {
    const bool next_final = (c + 1 == nc - 1) || bcut[c + 1] > 0;
    const T* src = next_final ? smth + (c + 1) * L * SG : &binit[c * SG];
    T a[N], Bv[NS];
    for (int i = 0; i < N; ++i) a[i] = src[i];
    for (int i = 0; i < NS; ++i) Bv[i] = src[N + i];
    const long t0 = c * L + bcut[c];
    ok = rts_backward_range<T, N, P, AM, 1, 2>(prm, filt + t0 * SG, smth + t0 * SG, nullptr, nullptr, nullptr,
                                               L - bcut[c], false, a, Bv) && ok;
  }
  if (cuts)
    for (long c = 0; c < nc; ++c) { cuts[c] = fcut[c]; cuts[nc + c] = bcut[c]; }
  return ok ? 0 : 1;
}

template <typename T>
static int dispatch_scan(int n, int p, int am, long Tn, long L, int early, const T* y, const T* ev, const T* m0,
                         const T* S0, const T* C, const T* A, const T* Q, T* filt, T* smth, long* cuts) {
#define X(N_, P_)                                                                       \
  if (n == N_ && p == P_)                                                               \
    return am ? run_scan<T, N_, P_, 1>(Tn, L, early, y, ev, m0, S0, C, A, Q, filt, smth, cuts)       \
              : run_scan<T, N_, P_, 0>(Tn, L, early, y, ev, m0, S0, C, A, Q, filt, smth, cuts);
  EKSB_FOR_EACH_DIM(X)
#undef X
  return -1;
}
extern "C" int hs_scan_f64(int n, int p, int am, long Tn, long L, int early, const double* y, const double* ev,
                           const double* m0, const double* S0, const double* C, const double* A, const double* Q,
                           double* filt, double* smth, long* cuts) {
  return dispatch_scan<double>(n, p, am, Tn, L, early, y, ev, m0, S0, C, A, Q, filt, smth, cuts);
}
extern "C" int hs_scan_f32(int n, int p, int am, long Tn, long L, int early, const float* y, const float* ev,
                           const float* m0, const float* S0, const float* C, const float* A, const float* Q, float* filt,
                           float* smth, long* cuts) {
  return dispatch_scan<float>(n, p, am, Tn, L, early, y, ev, m0, S0, C, A, Q, filt, smth, cuts);
}

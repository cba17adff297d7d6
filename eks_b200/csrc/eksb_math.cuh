// eksb_math.cuh -- small-matrix arithmetic and the per-track EKS recursions.
//
// Everything here is a __host__ __device__ template so that the *same* per-thread
// program the sm_100a kernels run can also be compiled by g++ and unit-tested on
// the CPU-only build box (tests/hostsim -- test infrastructure; the product never
// calls the host build).  State dimension N, observation dimension P and the
// scalar type are template parameters; all loops have constant bounds and are
// fully unrolled so every matrix lives in registers.
//
// Symmetric N x N matrices are stored packed, upper triangle row-major:
//   sym_idx<N>(i,j), i<=j  ->  i*N - i*(i-1)/2 + (j-i)      (NS = N(N+1)/2 values)
#pragma once
#include <math.h>

#if defined(__CUDACC__)
#define EKSB_HD __host__ __device__ __forceinline__
#define EKSB_UNROLL _Pragma("unroll")
#else
#define EKSB_HD inline __attribute__((always_inline))
#define EKSB_UNROLL
#endif

namespace eksb {

constexpr int kLanes = 32;  // tracks per track-tile (one warp)

template <int N> struct Dim { static constexpr int NS = N * (N + 1) / 2; static constexpr int SIGMA = N + NS; };

template <int N> EKSB_HD constexpr int sym_idx(int i, int j) {
  return (i <= j) ? (i * N - i * (i - 1) / 2 + (j - i)) : (j * N - j * (j - 1) / 2 + (i - j));
}

// Reciprocal.  Device: MUFU approximation + Newton steps (no IEEE-division slow path, which
// costs a CALL and ~50 serialised cycles on the recursion's critical path); relative error
// <= ~1 ulp for normal inputs, 0/inf/NaN propagate like 1/x.  Host (tests): plain 1/x.
EKSB_HD float rcp(float x) {
#if defined(__CUDA_ARCH__)
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  float e = fmaf(-x, r, 1.0f);
  return (e == e) ? fmaf(r, e, r) : r;       // skip the refinement when x*r is NaN (x = 0 or inf)
#else
  return 1.0f / x;
#endif
}
EKSB_HD double rcp(double x) {
#if defined(__CUDA_ARCH__)
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double e = fma(-x, r, 1.0);
  r = fma(r, e, r);
  e = fma(-x, r, 1.0);
  double r2 = fma(r, e, r);
  return (e == e) ? r2 : r;
#else
  return 1.0 / x;
#endif
}

// Shorter-latency reciprocal for the forward chain's critical path: one Newton step on the MUFU seed
// (double: seed 2^-23 -> ~2^-45 relative, i.e. ~3e-14; float: the seed itself, ~1e-7).  The filter is
// contractive, so this error does not accumulate; parity bars are 1e-5 (f64) / 1e-3 (f32) relative.
EKSB_HD float rcp_chain(float x) {
#if defined(__CUDA_ARCH__)
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
#else
  return 1.0f / x;
#endif
}
EKSB_HD double rcp_chain(double x) {
#if defined(__CUDA_ARCH__)
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  const double e = fma(-x, r, 1.0);
  return fma(r, e, r);
#else
  return 1.0 / x;
#endif
}

// ---------------------------------------------------------------------------
// SPD inverse of a packed symmetric matrix.  Returns det(a) (used by the NLL).
// N<=3: closed-form adjugate (one reciprocal, short dependency chain);
// N>3 : unrolled LDL^T, inverse = L^-T D^-1 L^-1.
// ok is cleared when a pivot / determinant is <= 0 (the reference would raise
// numpy.linalg.LinAlgError or silently produce garbage).  The tests are written !(x <= 0) so that NaN
// does NOT clear it: the reference propagates NaN silently and so does this library.
// ---------------------------------------------------------------------------
template <typename T, int N> struct SymInv;

template <typename T> struct SymInv<T, 1> {
  static EKSB_HD T run(const T* a, T* inv, bool& ok) { ok = ok && !(a[0] <= T(0)); inv[0] = rcp(a[0]); return a[0]; }
};

template <typename T> struct SymInv<T, 2> {
  static EKSB_HD T run(const T* a, T* inv, bool& ok) {
    T det = a[0] * a[2] - a[1] * a[1];
    ok = ok && !(det <= T(0)) && !(a[0] <= T(0));
    T r = rcp(det);
    inv[0] = a[2] * r; inv[1] = -a[1] * r; inv[2] = a[0] * r;
    return det;
  }
};

template <typename T> struct SymInv<T, 3> {
  static EKSB_HD T run(const T* a, T* inv, bool& ok) {
    // a = [a00 a01 a02 a11 a12 a22]
    T c00 = a[3] * a[5] - a[4] * a[4];
    T c01 = a[2] * a[4] - a[1] * a[5];
    T c02 = a[1] * a[4] - a[2] * a[3];
    T c11 = a[0] * a[5] - a[2] * a[2];
    T c12 = a[1] * a[2] - a[0] * a[4];
    T c22 = a[0] * a[3] - a[1] * a[1];
    T det = a[0] * c00 + a[1] * c01 + a[2] * c02;
    ok = ok && !(det <= T(0)) && !(a[0] <= T(0)) && !(c22 <= T(0));
    T r = rcp(det);
    inv[0] = c00 * r; inv[1] = c01 * r; inv[2] = c02 * r;
    inv[3] = c11 * r; inv[4] = c12 * r; inv[5] = c22 * r;
    return det;
  }
};

template <typename T, int N> struct SymInv {
  static EKSB_HD T run(const T* a, T* inv, bool& ok) {
    T L[N][N];   // unit lower factor (strict lower part used)
    T d[N], rd[N];
    T det = T(1);
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) {
      T s = a[sym_idx<N>(j, j)];
      EKSB_UNROLL
      for (int k = 0; k < j; ++k) s -= L[j][k] * L[j][k] * d[k];
      d[j] = s; ok = ok && !(s <= T(0)); rd[j] = rcp(s); det *= s;
      EKSB_UNROLL
      for (int i = j + 1; i < N; ++i) {
        T v = a[sym_idx<N>(j, i)];
        EKSB_UNROLL
        for (int k = 0; k < j; ++k) v -= L[i][k] * L[j][k] * d[k];
        L[i][j] = v * rd[j];
      }
    }
    // Li = L^-1 (unit lower): Li[i][j] = -sum_{k=j}^{i-1} L[i][k] Li[k][j]
    T Li[N][N];
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) {
      Li[j][j] = T(1);
      EKSB_UNROLL
      for (int i = j + 1; i < N; ++i) {
        T s = T(0);
        EKSB_UNROLL
        for (int k = j; k < i; ++k) s -= L[i][k] * Li[k][j];
        Li[i][j] = s;
      }
    }
    // inv = Li^T D^-1 Li
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      EKSB_UNROLL
      for (int j = i; j < N; ++j) {
        T s = T(0);
        EKSB_UNROLL
        for (int k = j; k < N; ++k) s += Li[k][i] * Li[k][j] * rd[k];
        inv[sym_idx<N>(i, j)] = s;
      }
    }
    return det;
  }
};

template <typename T, int N> EKSB_HD T sym_inv(const T* a, T* inv, bool& ok) { return SymInv<T, N>::run(a, inv, ok); }

// Adjugate and determinant of a general (non-symmetric) N x N matrix, N <= 3:  M^-1 = adj / det.
template <typename T, int N> EKSB_HD void gen_adjugate(const T M[N][N], T adj[N][N], T& det) {
  if (N == 1) {
    adj[0][0] = T(1);
    det = M[0][0];
  } else if (N == 2) {
    adj[0][0] = M[1 % N][1 % N]; adj[0][1 % N] = -M[0][1 % N];
    adj[1 % N][0] = -M[1 % N][0]; adj[1 % N][1 % N] = M[0][0];
    det = M[0][0] * M[1 % N][1 % N] - M[0][1 % N] * M[1 % N][0];
  } else {
    constexpr int a = 1 % N, b = 2 % N;    // (indices written so that N < 3 instantiations still compile)
    adj[0][0] = M[a][a] * M[b][b] - M[a][b] * M[b][a];
    adj[0][a] = M[0][b] * M[b][a] - M[0][a] * M[b][b];
    adj[0][b] = M[0][a] * M[a][b] - M[0][b] * M[a][a];
    adj[a][0] = M[a][b] * M[b][0] - M[a][0] * M[b][b];
    adj[a][a] = M[0][0] * M[b][b] - M[0][b] * M[b][0];
    adj[a][b] = M[0][b] * M[a][0] - M[0][0] * M[a][b];
    adj[b][0] = M[a][0] * M[b][a] - M[a][a] * M[b][0];
    adj[b][a] = M[0][a] * M[b][0] - M[0][0] * M[b][a];
    adj[b][b] = M[0][0] * M[a][a] - M[0][a] * M[a][0];
    det = M[0][0] * adj[0][0] + M[0][a] * adj[a][0] + M[0][b] * adj[b][0];
  }
}

// y = S x  (S packed symmetric)
template <typename T, int N> EKSB_HD void sym_matvec(const T* S, const T* x, T* y) {
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    T s = T(0);
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) s += S[sym_idx<N>(i, j)] * x[j];
    y[i] = s;
  }
}

// out (packed sym) = J D J^T, D packed symmetric, J full row-major N x N
template <typename T, int N> EKSB_HD void sandwich(const T J[N][N], const T* D, T* out) {
  T W[N][N];
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) {
      T s = T(0);
      EKSB_UNROLL
      for (int k = 0; k < N; ++k) s += J[i][k] * D[sym_idx<N>(k, j)];
      W[i][j] = s;
    }
  }
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = i; j < N; ++j) {
      T s = T(0);
      EKSB_UNROLL
      for (int k = 0; k < N; ++k) s += W[i][k] * J[j][k];
      out[sym_idx<N>(i, j)] = s;
    }
  }
}

// ---------------------------------------------------------------------------
// Per-track model parameters, held in registers for the whole pass.
// AM = 0: diagonal state transition (covers A = I, every shipped config);
// AM = 1: dense A.
// ---------------------------------------------------------------------------
template <typename T, int N, int P, int AM> struct Params {
  static constexpr int NS = Dim<N>::NS;
  T C[P][N];
  T A[AM ? N * N : N];
  T AA[AM ? 1 : NS];   // a_i * a_j (diagonal mode only)
  T Q[NS];
};

// Load from dense row-major host-style arrays: C (P,N), A (N,N), Q (N,N).
template <typename T, int N, int P, int AM>
EKSB_HD void load_params(Params<T, N, P, AM>& prm, const T* C, const T* A, const T* Q) {
  EKSB_UNROLL
  for (int j = 0; j < P; ++j) {
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) prm.C[j][i] = C[j * N + i];
  }
  if (AM) {
    EKSB_UNROLL
    for (int i = 0; i < N * N; ++i) prm.A[i] = A[i];
  } else {
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) prm.A[i] = A[i * N + i];
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      EKSB_UNROLL
      for (int j = i; j < N; ++j) prm.AA[sym_idx<N>(i, j)] = A[i * N + i] * A[j * N + j];
    }
  }
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = i; j < N; ++j) prm.Q[sym_idx<N>(i, j)] = T(0.5) * (Q[i * N + j] + Q[j * N + i]);
  }
}

// packed symmetric from a dense row-major (N,N) array
template <typename T, int N> EKSB_HD void load_sym(T* S, const T* dense) {
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = i; j < N; ++j) S[sym_idx<N>(i, j)] = T(0.5) * (dense[i * N + j] + dense[j * N + i]);
  }
}

// mp = A m ; S = A V A^T + Q        (eks/ensemble_kalman.py:101-102)
template <typename T, int N, int P, int AM>
EKSB_HD void predict(const Params<T, N, P, AM>& prm, const T* m, const T* V, T* mp, T* S) {
  if (AM == 0) {
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) mp[i] = prm.A[i] * m[i];
    EKSB_UNROLL
    for (int k = 0; k < Dim<N>::NS; ++k) S[k] = prm.AA[k] * V[k] + prm.Q[k];
  } else {
    T G[N][N];
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      T s = T(0);
      EKSB_UNROLL
      for (int j = 0; j < N; ++j) s += prm.A[i * N + j] * m[j];
      mp[i] = s;
      EKSB_UNROLL
      for (int j = 0; j < N; ++j) {
        T g = T(0);
        EKSB_UNROLL
        for (int k = 0; k < N; ++k) g += prm.A[i * N + k] * V[sym_idx<N>(k, j)];
        G[i][j] = g;
      }
    }
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      EKSB_UNROLL
      for (int j = i; j < N; ++j) {
        T s = prm.Q[sym_idx<N>(i, j)];
        EKSB_UNROLL
        for (int k = 0; k < N; ++k) s += G[i][k] * prm.A[j * N + k];
        S[sym_idx<N>(i, j)] = s;
      }
    }
  }
}

// Measurement update with R = diag(r), in information form (N x N inverses only,
// independent of P):      Vf = (S^-1 + C^T R^-1 C)^-1 ,  mf = mp + Vf C^T R^-1 (y - C mp).
// Algebraically identical to eks/ensemble_kalman.py:104-105,110-117 (Woodbury) and to
// the update at eks/newton_eks.py:137-138.  When NLL is set, returns this frame's
//   log det(R + C S C^T) + nu^T (R + C S C^T)^-1 nu      (determinant lemma).
// The update is split in two so that a sweep over several process-noise scales (SURVEY A8 grid search) can
// share the state-independent half between the scales:
//   update_shared: w = 1/r, cw = C^T R^-1 (per observation), Lam0 = C^T R^-1 C, detR = prod r
//   update_state : everything that depends on the predicted state (mp, S)
template <typename T, int N, int P, int AM, bool NLL>
EKSB_HD void update_shared(const Params<T, N, P, AM>& prm, const T* r, T* w, T (*cw)[N], T* Lam0, double& detR) {
  constexpr int NS = Dim<N>::NS;
  EKSB_UNROLL
  for (int j = 0; j < P; ++j) w[j] = rcp(r[j]);
  EKSB_UNROLL
  for (int k = 0; k < NS; ++k) Lam0[k] = T(0);   // C^T R^-1 C is accumulated apart from S^-1 (state-independent)
  EKSB_UNROLL
  for (int j = 0; j < P; ++j) {
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) cw[j][i] = prm.C[j][i] * w[j];
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      EKSB_UNROLL
      for (int k = i; k < N; ++k) Lam0[sym_idx<N>(i, k)] += cw[j][i] * prm.C[j][k];
    }
  }
  detR = 1.0;
  if (NLL) {
    EKSB_UNROLL
    for (int j = 0; j < P; ++j) detR *= (double)r[j];
  }
}

template <typename T, int N, int P, int AM, bool NLL>
EKSB_HD double update_state(const Params<T, N, P, AM>& prm, const T* y, const T* w, const T (*cw)[N], const T* Lam0,
                            double detR, const T* mp, const T* S, T* m, T* V, bool& ok) {
  constexpr int NS = Dim<N>::NS;
  T nu[P];
  EKSB_UNROLL
  for (int j = 0; j < P; ++j) {
    T s = y[j];
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) s -= prm.C[j][i] * mp[i];
    nu[j] = s;
  }
  T Sinv[NS], Lam[NS];
  T detS = sym_inv<T, N>(S, Sinv, ok);
  T u[N];
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) u[i] = T(0);
  EKSB_UNROLL
  for (int j = 0; j < P; ++j) {
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) u[i] += cw[j][i] * nu[j];
  }
  EKSB_UNROLL
  for (int k = 0; k < NS; ++k) Lam[k] = Lam0[k] + Sinv[k];
  T detL = sym_inv<T, N>(Lam, V, ok);
  T Vu[N];
  sym_matvec<T, N>(V, u, Vu);
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) m[i] = mp[i] + Vu[i];
  if (NLL) {
    double quad = 0.0, uVu = 0.0;
    EKSB_UNROLL
    for (int j = 0; j < P; ++j) quad += (double)nu[j] * (double)nu[j] * (double)w[j];
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) uVu += (double)u[i] * (double)Vu[i];
    return log(detR * (double)detS * (double)detL) + (quad - uVu);
  }
  return 0.0;
}

template <typename T, int N, int P, int AM, bool NLL>
EKSB_HD double update(const Params<T, N, P, AM>& prm, const T* y, const T* r, const T* mp, const T* S,
                      T* m, T* V, bool& ok) {
  T w[P], cw[P][N], Lam0[Dim<N>::NS];
  double detR;
  update_shared<T, N, P, AM, NLL>(prm, r, w, cw, Lam0, detR);
  return update_state<T, N, P, AM, NLL>(prm, y, w, cw, Lam0, detR, mp, S, m, V, ok);
}

// ---------------------------------------------------------------------------
// Track-tile memory layout (DESIGN.md "HBM layout").  A tile is kLanes tracks;
// a frame record is ROWS scalars per track stored as ROWS rows of kLanes lanes:
//     addr(tile g, frame t, row r, lane l) = ((g*T + t)*ROWS + r)*kLanes + l
// so a warp (lane = track) reads/writes one full 128 B (f32) / 256 B (f64) line per
// row.  The range functions below take a pointer already offset to (g, t0, row 0,
// lane) and a compile-time lane stride LS (kLanes on the device, 1 in host tests).
//   obs record   : rows [0,P) = y, [P,2P) = ensemble variance          (2P rows)
//   state record : rows [0,N) = mean, [N,N+NS) = packed covariance      (SIGMA rows)
// ---------------------------------------------------------------------------
enum FirstFrame { kContinue = 0, kPriorUpdate = 1, kPriorOnly = 2 };

// Per-track status bits (include/eksb.h): 1 = a non-positive pivot (the reference's LinAlgError), 2 = a frame with a
// non-positive ensemble variance.  The information-form update divides by every variance, so such a track comes out
// NaN where the reference (which only needs R + C S C^T to be invertible) stays finite: the fast kernels flag it and
// the caller re-runs the track through the literal p x p LU kernel (eksb_kf_forward_generic).  NaN variances are NOT
// flagged: the reference propagates them and so do the kernels.
constexpr int kStatusPivot = 1, kStatusNonPosVar = 2;
template <typename T, int P> EKSB_HD bool any_nonpos(const T* r) {
  bool bad = false;
  EKSB_UNROLL
  for (int j = 0; j < P; ++j) bad = bad || (r[j] <= T(0));
  return bad;
}

// Forward filter over `n` frames.  (m,V) carries the state in and out:
//   first == kPriorUpdate : (m,V) = (m0,S0), frame 0 is an update with no predict
//                           (eks/ensemble_kalman.py:94-96);
//   first == kPriorOnly   : frame 0 just emits (m,V) (eks/newton_eks.py:128,132);
//   first == kContinue    : (m,V) is the filtered state of the frame before this range.
// filt may be null (NLL-only sweeps).  PF = software prefetch depth in frames.
template <typename T, int N, int P, int AM, int LS, int PF, bool NLL>
EKSB_HD bool kf_forward_range(const Params<T, N, P, AM>& prm, const T* __restrict__ obs, T* __restrict__ filt,
                              long n, int first, T* m, T* V, double& nll, const T* mu = nullptr, bool* nonpos = nullptr) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, R2 = 2 * P;
  bool ok = true;
  T buf[PF][R2];
  EKSB_UNROLL
  for (int k = 0; k < PF; ++k) {
    long f = (k < n) ? k : (n - 1);
    EKSB_UNROLL
    for (int r = 0; r < R2; ++r) buf[k][r] = obs[(f * R2 + r) * LS];
  }
  for (long t = 0; t < n; t += PF) {
    EKSB_UNROLL
    for (int k = 0; k < PF; ++k) {
      if (t + k < n) {
        T cur[R2];
        EKSB_UNROLL
        for (int r = 0; r < R2; ++r) cur[r] = buf[k][r];
        long f = t + k + PF;
        if (f > n - 1) f = n - 1;
        EKSB_UNROLL
        for (int r = 0; r < R2; ++r) buf[k][r] = obs[(f * R2 + r) * LS];
        const bool is_first = (t + k == 0);
        if (mu) {
          EKSB_UNROLL
          for (int j = 0; j < P; ++j) cur[j] -= mu[j];
        }
        if (nonpos && !(is_first && first == kPriorOnly)) *nonpos = *nonpos || any_nonpos<T, P>(cur + P);
        if (!(is_first && first == kPriorOnly)) {
          T mp[N], S[NS];
          if (is_first && first == kPriorUpdate) {
            EKSB_UNROLL
            for (int i = 0; i < N; ++i) mp[i] = m[i];
            EKSB_UNROLL
            for (int i = 0; i < NS; ++i) S[i] = V[i];
          } else {
            predict<T, N, P, AM>(prm, m, V, mp, S);
          }
          double e = update<T, N, P, AM, NLL>(prm, cur, cur + P, mp, S, m, V, ok);
          if (NLL) nll += e;
        }
        if (filt) {
          T* o = filt + (t + k) * SG * LS;
          EKSB_UNROLL
          for (int i = 0; i < N; ++i) o[i * LS] = m[i];
          EKSB_UNROLL
          for (int i = 0; i < NS; ++i) o[(N + i) * LS] = V[i];
        }
      }
    }
  }
  return ok;
}

// One RTS step (eks/ensemble_kalman.py:158-162): from filtered (m,V) at t and smoothed
// (ms1,Vs1) at t+1 to smoothed (ms,Vs) at t.  S = A V A^T + Q is recomputed, not stored.
// Jout (optional, full N x N) receives J so the caller can form CV = Vs1 J^T.
template <typename T, int N, int P, int AM>
EKSB_HD void rts_step(const Params<T, N, P, AM>& prm, const T* m, const T* V, const T* ms1, const T* Vs1,
                      T* ms, T* Vs, T (*Jout)[N], bool& ok) {
  constexpr int NS = Dim<N>::NS;
  T mp[N], S[NS], Sinv[NS];
  predict<T, N, P, AM>(prm, m, V, mp, S);
  sym_inv<T, N>(S, Sinv, ok);
  T G[N][N], J[N][N];
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) {
      if (AM == 0) {
        G[i][j] = V[sym_idx<N>(i, j)] * prm.A[j];
      } else {
        T s = T(0);
        EKSB_UNROLL
        for (int k = 0; k < N; ++k) s += V[sym_idx<N>(i, k)] * prm.A[j * N + k];
        G[i][j] = s;
      }
    }
  }
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) {
      T s = T(0);
      EKSB_UNROLL
      for (int k = 0; k < N; ++k) s += G[i][k] * Sinv[sym_idx<N>(k, j)];
      J[i][j] = s;
    }
  }
  T d[N], D[NS], JDJ[NS];
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) d[i] = ms1[i] - mp[i];
  EKSB_UNROLL
  for (int i = 0; i < NS; ++i) D[i] = Vs1[i] - S[i];
  sandwich<T, N>(J, D, JDJ);
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    T s = m[i];
    EKSB_UNROLL
    for (int k = 0; k < N; ++k) s += J[i][k] * d[k];
    ms[i] = s;
  }
  EKSB_UNROLL
  for (int i = 0; i < NS; ++i) Vs[i] = V[i] + JDJ[i];
  if (Jout) {
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      EKSB_UNROLL
      for (int j = 0; j < N; ++j) Jout[i][j] = J[i][j];
    }
  }
}

// Backward pass over `n` frames, last frame first.  (ms,Vs) carries the smoothed state
// of the frame after this range in, and of the first frame of the range out; with
// terminal set, the last frame of the range is the end of the track (ms=mf, Vs=Vf,
// eks/ensemble_kalman.py:152-153).  smth may alias filt (in-place smoothing: a frame
// is always read before it is overwritten).  Optional outputs, null to skip:
//   ym (P rows)  = C ms                      (multiview_pca_smoother.py:745)
//   yv (P rows)  = diag(C Vs C^T)            (:746, diagonal only)
//   cv (N*N rows)= Vs[t+1] J^T  at frame t   (eks/ensemble_kalman.py:162; frame n-1 of a
//                  terminal range is left untouched)
template <typename T, int N, int P, int AM, int LS, int PF>
EKSB_HD bool rts_backward_range(const Params<T, N, P, AM>& prm, const T* filt, T* smth, T* ym, T* yv, T* cv,
                                long n, bool terminal, T* ms, T* Vs) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA;
  bool ok = true;
  T buf[PF][SG];
  EKSB_UNROLL
  for (int k = 0; k < PF; ++k) {
    long f = n - 1 - k;
    if (f < 0) f = 0;
    EKSB_UNROLL
    for (int r = 0; r < SG; ++r) buf[k][r] = filt[(f * SG + r) * LS];
  }
  for (long t = n - 1; t >= 0; t -= PF) {
    EKSB_UNROLL
    for (int k = 0; k < PF; ++k) {
      if (t - k >= 0) {
        T cur[SG];
        EKSB_UNROLL
        for (int r = 0; r < SG; ++r) cur[r] = buf[k][r];
        long f = t - k - PF;
        if (f < 0) f = 0;
        EKSB_UNROLL
        for (int r = 0; r < SG; ++r) buf[k][r] = filt[(f * SG + r) * LS];
        const long tt = t - k;
        if (terminal && tt == n - 1) {
          EKSB_UNROLL
          for (int i = 0; i < N; ++i) ms[i] = cur[i];
          EKSB_UNROLL
          for (int i = 0; i < NS; ++i) Vs[i] = cur[N + i];
        } else {
          T ms1[N], Vs1[NS], J[N][N];
          EKSB_UNROLL
          for (int i = 0; i < N; ++i) ms1[i] = ms[i];
          EKSB_UNROLL
          for (int i = 0; i < NS; ++i) Vs1[i] = Vs[i];
          rts_step<T, N, P, AM>(prm, cur, cur + N, ms1, Vs1, ms, Vs, cv ? J : nullptr, ok);
          if (cv) {
            T* o = cv + tt * (N * N) * LS;
            EKSB_UNROLL
            for (int i = 0; i < N; ++i) {
              EKSB_UNROLL
              for (int j = 0; j < N; ++j) {
                T s = T(0);
                EKSB_UNROLL
                for (int k2 = 0; k2 < N; ++k2) s += Vs1[sym_idx<N>(i, k2)] * J[j][k2];
                o[(i * N + j) * LS] = s;
              }
            }
          }
        }
        T* o = smth + tt * SG * LS;
        EKSB_UNROLL
        for (int i = 0; i < N; ++i) o[i * LS] = ms[i];
        EKSB_UNROLL
        for (int i = 0; i < NS; ++i) o[(N + i) * LS] = Vs[i];
        if (ym || yv) {
          EKSB_UNROLL
          for (int j = 0; j < P; ++j) {
            T s = T(0), q = T(0);
            EKSB_UNROLL
            for (int i = 0; i < N; ++i) {
              s += prm.C[j][i] * ms[i];
              T a = T(0);
              EKSB_UNROLL
              for (int k2 = 0; k2 < N; ++k2) a += Vs[sym_idx<N>(i, k2)] * prm.C[j][k2];
              q += prm.C[j][i] * a;
            }
            if (ym) ym[(tt * P + j) * LS] = s;
            if (yv) yv[(tt * P + j) * LS] = q;
          }
        }
      }
    }
  }
  return ok;
}

// ---------------------------------------------------------------------------
// Producer / consumer split of the two recursions (used by the warp-specialised kernels,
// eksb_pipe.cuh).  Everything that does not depend on the running state is done by helper
// warps ("prep"), so the single warp that carries the sequential dependency ("chain") only
// executes the short critical path.
//   forward  prep : obs frame (y, r) -> info record  [ I = C^T R^-1 C (NS) | iv = C^T R^-1 y (N) ]
//   forward  chain: S = A V A^T + Q; V = (S^-1 + I)^-1; m = mp + V (iv - I mp)
//   backward prep : filtered (m, V)  -> smoothing element [ J (N*N) | L = V - J A V (NS) | g = m - J A m (N) ]
//                   with J = V A^T (A V A^T + Q)^-1   (eks/ensemble_kalman.py:158; the (E,g,L)
//                   element of Sarkka & Garcia-Fernandez 2021)
//   backward chain: ms = g + J ms1 ; Vs = L + J Vs1 J^T      (== :160-161)
// ---------------------------------------------------------------------------
template <int N> struct Rec {
  static constexpr int FWD = Dim<N>::NS + N;
  static constexpr int BWD = N * N + Dim<N>::NS + N;
};

template <typename T, int N, int P, int AM>
EKSB_HD void fwd_prep(const Params<T, N, P, AM>& prm, const T* y, const T* r, T* rec) {
  constexpr int NS = Dim<N>::NS;
  EKSB_UNROLL
  for (int k = 0; k < NS + N; ++k) rec[k] = T(0);
  EKSB_UNROLL
  for (int j = 0; j < P; ++j) {
    const T w = rcp(r[j]);
    T cw[N];
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) cw[i] = prm.C[j][i] * w;
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      rec[NS + i] += cw[i] * y[j];
      EKSB_UNROLL
      for (int k = i; k < N; ++k) rec[sym_idx<N>(i, k)] += cw[i] * prm.C[j][k];
    }
  }
}

// fwd_prep with the outer products C_j C_j^T precomputed once per track (helper warps of the pipelined
// kernel): per observation 1 reciprocal + 1 + N + NS multiply-adds instead of 1 + 2N + NS, and the
// one-Newton-step reciprocal (3e-14 relative in f64).  Same record as fwd_prep up to rounding.
template <typename T, int N, int P> EKSB_HD void make_outer(const T C[P][N], T CC[P][Dim<N>::NS]) {
  EKSB_UNROLL
  for (int j = 0; j < P; ++j) {
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      EKSB_UNROLL
      for (int k = i; k < N; ++k) CC[j][sym_idx<N>(i, k)] = C[j][i] * C[j][k];
    }
  }
}
template <typename T, int N, int P>
EKSB_HD void fwd_prep_outer(const T C[P][N], const T CC[P][Dim<N>::NS], const T* y, const T* r, T* rec) {
  constexpr int NS = Dim<N>::NS;
  EKSB_UNROLL
  for (int k = 0; k < NS + N; ++k) rec[k] = T(0);
  EKSB_UNROLL
  for (int j = 0; j < P; ++j) {
    const T w = rcp_chain(r[j]);
    const T wy = w * y[j];
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) rec[NS + i] += C[j][i] * wy;
    EKSB_UNROLL
    for (int k = 0; k < NS; ++k) rec[k] += w * CC[j][k];
  }
}

template <typename T, int N, int P, int AM>
EKSB_HD void fwd_chain(const Params<T, N, P, AM>& prm, const T* rec, bool do_predict, T* m, T* V, bool& ok) {
  constexpr int NS = Dim<N>::NS;
  T mp[N], S[NS], Sinv[NS], Lam[NS], u[N], Vu[N];
  if (do_predict) {
    predict<T, N, P, AM>(prm, m, V, mp, S);
  } else {
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) mp[i] = m[i];
    EKSB_UNROLL
    for (int i = 0; i < NS; ++i) S[i] = V[i];
  }
  sym_matvec<T, N>(rec, mp, u);                   // I mp  (off the covariance chain)
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) u[i] = rec[NS + i] - u[i];
  if (N <= 3) {
    // V = (S^-1 + Lam)^-1 = S (Id + Lam S)^-1 : ONE reciprocal on the frame-to-frame dependency
    // path instead of two (the two nested symmetric inverses cost ~27 dependent DP ops, this ~17;
    // the forward chain warp is bound by exactly that latency).  det(Id + Lam S) >= 1 for SPD S.
    T M[N][N];
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      EKSB_UNROLL
      for (int j = 0; j < N; ++j) {
        T a = (i == j) ? T(1) : T(0);
        EKSB_UNROLL
        for (int k = 0; k < N; ++k) a += rec[sym_idx<N>(i, k)] * S[sym_idx<N>(k, j)];
        M[i][j] = a;
      }
    }
    T adj[N][N], det;
    gen_adjugate<T, N>(M, adj, det);
    ok = ok && !(det <= T(0)) && !(S[0] <= T(0));
    const T r = rcp(det);
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      EKSB_UNROLL
      for (int j = i; j < N; ++j) {
        T a = T(0);
        EKSB_UNROLL
        for (int k = 0; k < N; ++k) a += S[sym_idx<N>(i, k)] * adj[k][j];
        V[sym_idx<N>(i, j)] = a * r;
      }
    }
  } else {
    sym_inv<T, N>(S, Sinv, ok);
    EKSB_UNROLL
    for (int k = 0; k < NS; ++k) Lam[k] = Sinv[k] + rec[k];
    sym_inv<T, N>(Lam, V, ok);
  }
  sym_matvec<T, N>(V, u, Vu);
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) m[i] = mp[i] + Vu[i];
}

// Adjugate (packed symmetric) and determinant of a packed symmetric N x N matrix, N <= 3.
template <typename T, int N> EKSB_HD void sym_adjugate(const T* a, T* adj, T& det) {
  if (N == 1) {
    adj[0] = T(1);
    det = a[0];
  } else if (N == 2) {
    adj[0] = a[2 % Dim<N>::NS]; adj[1 % Dim<N>::NS] = -a[1 % Dim<N>::NS]; adj[2 % Dim<N>::NS] = a[0];
    det = a[0] * a[2 % Dim<N>::NS] - a[1 % Dim<N>::NS] * a[1 % Dim<N>::NS];
  } else {
    constexpr int M = Dim<N>::NS;          // (indices mod M so that N < 3 instantiations still compile)
    // a = [a00 a01 a02 a11 a12 a22]
    adj[0] = a[3 % M] * a[5 % M] - a[4 % M] * a[4 % M];
    adj[1 % M] = a[2 % M] * a[4 % M] - a[1 % M] * a[5 % M];
    adj[2 % M] = a[1 % M] * a[4 % M] - a[2 % M] * a[3 % M];
    adj[3 % M] = a[0] * a[5 % M] - a[2 % M] * a[2 % M];
    adj[4 % M] = a[1 % M] * a[2 % M] - a[0] * a[4 % M];
    adj[5 % M] = a[0] * a[3 % M] - a[1 % M] * a[1 % M];
    det = a[0] * adj[0] + a[1 % M] * adj[1 % M] + a[2 % M] * adj[2 % M];
  }
}

// Forward chain with the covariance carried between frames as V = Nm * r (a symmetric numerator and a
// scalar), so that no matrix is ever divided on the frame-to-frame path and only ONE reciprocal is taken:
//   S  = AA o (Nm r) + Q                      = fma(AA o Nm, r, Q)
//   Lt = det(S) (S^-1 + Lam)                  = adj(S) + det(S) Lam            (Lam = C^T R^-1 C from the helpers)
//   V  = (S^-1 + Lam)^-1 = det(S) adj(Lt) / det(Lt)    ->   Nm = adj(Lt),  r = det(S) / det(Lt)
// ~82 double-precision instructions per frame on the chain warp (the two nested symmetric inverses of
// fwd_chain need two dependent reciprocals; the Id + Lam S form needs ~114 instructions).  The chain warp has a
// scheduler to itself and each DP instruction occupies its FP64 pipe for 2 cycles, so instruction count, not
// dependency depth, sets the frame time.  Diagonal A and N <= 3 only (every shipped config); Vout receives the
// normalised covariance (off the path).  First use: Nm = V, r = 1.  Magnitudes: det(Lt) ~ (det S * Lam)^N,
// comfortably inside both float ranges for ensemble variances in [1e-4, 1e3].
// The covariance half: (Nm, r) of frame t-1 and Lam = C^T R^-1 C of frame t  ->  (Nm, r) of frame t.
// Independent of the mean, so the three-role forward kernel runs it on a warp of its own.
template <typename T, int N, int P>
EKSB_HD void fwd_cov_scaled(const Params<T, N, P, 0>& prm, const T* lam, bool do_predict, T* Nm, T& r, bool& ok) {
  constexpr int NS = Dim<N>::NS;
  T S[NS];
  if (do_predict) {
    EKSB_UNROLL
    for (int k = 0; k < NS; ++k) S[k] = (prm.AA[k] * Nm[k]) * r + prm.Q[k];
  } else {
    EKSB_UNROLL
    for (int k = 0; k < NS; ++k) S[k] = Nm[k] * r;
  }
  T adjS[NS], detS, Lt[NS], detL;
  sym_adjugate<T, N>(S, adjS, detS);
  EKSB_UNROLL
  for (int k = 0; k < NS; ++k) Lt[k] = detS * lam[k] + adjS[k];
  sym_adjugate<T, N>(Lt, Nm, detL);
  ok = ok && !(detS <= T(0)) && !(detL <= T(0)) && !(S[0] <= T(0));
  r = detS * rcp_chain(detL);
}
// The mean half: m of frame t-1, the record of frame t and (Nm, r) of frame t  ->  m and the normalised V of frame t.
template <typename T, int N, int P>
EKSB_HD void fwd_mean_scaled(const Params<T, N, P, 0>& prm, const T* rec, bool do_predict, const T* Nm, T r, T* m,
                             T* Vout) {
  constexpr int NS = Dim<N>::NS;
  T mp[N], u[N], Nu[N];
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) mp[i] = do_predict ? prm.A[i] * m[i] : m[i];
  sym_matvec<T, N>(rec, mp, u);
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) u[i] = rec[NS + i] - u[i];
  sym_matvec<T, N>(Nm, u, Nu);
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) m[i] = mp[i] + Nu[i] * r;
  EKSB_UNROLL
  for (int k = 0; k < NS; ++k) Vout[k] = Nm[k] * r;
}
template <typename T, int N, int P>
EKSB_HD void fwd_chain_scaled(const Params<T, N, P, 0>& prm, const T* rec, bool do_predict, T* m, T* Nm, T& r, T* Vout,
                              bool& ok) {
  fwd_cov_scaled<T, N, P>(prm, rec, do_predict, Nm, r, ok);
  fwd_mean_scaled<T, N, P>(prm, rec, do_predict, Nm, r, m, Vout);
}

// ---------------------------------------------------------------------------
// NLL in the producer / consumer split (kf_nll_pipe_kernel, eksb_pipe.cuh; diagonal A, N <= 3 -- the scaled chain).
// Per frame  log det Sig + nu^T Sig^-1 nu,  Sig = R + C S C^T (== update_state<NLL>), from
//   log det Sig = log prod_j r_j + log det(Id + Lam S),    det(Id + Lam S) = det(Lt) / det(S)^(N-1) = 1 / (r det(S)^(N-2))
//   nu^T Sig^-1 nu = quad - u^T V u,   quad = sum_j w_j nu_j^2 = yy - mp.(iv + u),   u = iv - Lam mp,   u^T V u = r u.(Nm u)
// with Lam, iv from the info record, yy = sum_j w_j y_j^2 and ldr = log prod_j r_j two more values of it: nothing in the
// record depends on Q, so ONE helper stream feeds the chains of every smoothing-parameter candidate.
// ---------------------------------------------------------------------------
template <int N> struct RecNll { static constexpr int FWD = Rec<N>::FWD + 2; };   // Lam | iv | yy | ldr

template <typename T, int N, int P>
EKSB_HD void fwd_prep_outer_nll(const T C[P][N], const T CC[P][Dim<N>::NS], const T* y, const T* r, T* rec) {
  constexpr int NS = Dim<N>::NS;
  EKSB_UNROLL
  for (int k = 0; k < NS + N; ++k) rec[k] = T(0);
  T yy = T(0), prod = T(1);
  EKSB_UNROLL
  for (int j = 0; j < P; ++j) {
    const T w = rcp_chain(r[j]);
    const T wy = w * y[j];
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) rec[NS + i] += C[j][i] * wy;
    EKSB_UNROLL
    for (int k = 0; k < NS; ++k) rec[k] += w * CC[j][k];
    yy += wy * y[j];
    prod *= r[j];
  }
  rec[NS + N] = yy;
  rec[NS + N + 1] = (T)log((double)prod);
}

// the same record without the precomputed outer products (observation dims whose C_j C_j^T do not fit in registers)
template <typename T, int N, int P>
EKSB_HD void fwd_prep_nll(const T C[P][N], const T* y, const T* r, T* rec) {
  constexpr int NS = Dim<N>::NS;
  EKSB_UNROLL
  for (int k = 0; k < NS + N; ++k) rec[k] = T(0);
  T yy = T(0), prod = T(1);
  EKSB_UNROLL
  for (int j = 0; j < P; ++j) {
    const T w = rcp_chain(r[j]);
    const T wy = w * y[j];
    T cw[N];
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) cw[i] = C[j][i] * w;
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      rec[NS + i] += C[j][i] * wy;
      EKSB_UNROLL
      for (int k = i; k < N; ++k) rec[sym_idx<N>(i, k)] += cw[i] * C[j][k];
    }
    yy += wy * y[j];
    prod *= r[j];
  }
  rec[NS + N] = yy;
  rec[NS + N + 1] = (T)log((double)prod);
}

// One frame of an NLL chain: (m, Nm, r) carried like fwd_chain_scaled (V = Nm r), no normalised covariance formed.
// xdet receives 1 / det(Id + Lam S) (in (0, 1]; the caller multiplies a few of them before taking a log),
// returns yy - mp.(iv + u) - r u.(Nm u).
template <typename T, int N, int P>
EKSB_HD T fwd_nll_scaled(const Params<T, N, P, 0>& prm, const T* rec, bool do_predict, T* m, T* Nm, T& r, T& xdet, bool& ok) {
  constexpr int NS = Dim<N>::NS;
  T S[NS];
  if (do_predict) {
    EKSB_UNROLL
    for (int k = 0; k < NS; ++k) S[k] = (prm.AA[k] * Nm[k]) * r + prm.Q[k];
  } else {
    EKSB_UNROLL
    for (int k = 0; k < NS; ++k) S[k] = Nm[k] * r;
  }
  T adjS[NS], detS, Lt[NS], detL;
  sym_adjugate<T, N>(S, adjS, detS);
  EKSB_UNROLL
  for (int k = 0; k < NS; ++k) Lt[k] = detS * rec[k] + adjS[k];
  sym_adjugate<T, N>(Lt, Nm, detL);
  ok = ok && !(detS <= T(0)) && !(detL <= T(0)) && !(S[0] <= T(0));
  r = detS * rcp(detL);
  xdet = (N == 3) ? r * detS : (N == 2 ? r : r * rcp(detS));
  T mp[N], u[N], Nu[N];
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) mp[i] = do_predict ? prm.A[i] * m[i] : m[i];
  sym_matvec<T, N>(rec, mp, u);
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) u[i] = rec[NS + i] - u[i];
  sym_matvec<T, N>(Nm, u, Nu);
  T e = rec[NS + N], q = T(0);
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    e -= mp[i] * (rec[NS + i] + u[i]);
    q += u[i] * Nu[i];
    m[i] = mp[i] + Nu[i] * r;
  }
  return e - r * q;
}

template <typename T, int N, int P, int AM>
EKSB_HD void bwd_prep(const Params<T, N, P, AM>& prm, const T* m, const T* V, T* rec, bool& ok) {
  constexpr int NS = Dim<N>::NS;
  T mp[N], S[NS], Sinv[NS];
  predict<T, N, P, AM>(prm, m, V, mp, S);
  sym_inv<T, N>(S, Sinv, ok);
  T G[N][N], J[N][N];                             // G = V A^T
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) {
      if (AM == 0) {
        G[i][j] = V[sym_idx<N>(i, j)] * prm.A[j];
      } else {
        T s = T(0);
        EKSB_UNROLL
        for (int k = 0; k < N; ++k) s += V[sym_idx<N>(i, k)] * prm.A[j * N + k];
        G[i][j] = s;
      }
    }
  }
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) {
      T s = T(0);
      EKSB_UNROLL
      for (int k = 0; k < N; ++k) s += G[i][k] * Sinv[sym_idx<N>(k, j)];
      J[i][j] = s;
      rec[i * N + j] = s;
    }
  }
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = i; j < N; ++j) {                 // L = V - J (A V),  (A V)_{kj} = G[j][k]
      T s = V[sym_idx<N>(i, j)];
      EKSB_UNROLL
      for (int k = 0; k < N; ++k) s -= J[i][k] * G[j][k];
      rec[N * N + sym_idx<N>(i, j)] = s;
    }
    T gi = m[i];
    EKSB_UNROLL
    for (int k = 0; k < N; ++k) gi -= J[i][k] * mp[k];
    rec[N * N + NS + i] = gi;
  }
}

// The two halves of bwd_chain are independent recursions (the smoothed mean does not enter the smoothed
// covariance and vice versa); the backward kernel runs them on two warps when an SM holds a single tile.
template <typename T, int N>
EKSB_HD void bwd_chain_mean(const T* rec, T* ms) {
  constexpr int NS = Dim<N>::NS;
  T ms1[N];
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) ms1[i] = ms[i];
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    T s = rec[N * N + NS + i];
    EKSB_UNROLL
    for (int k = 0; k < N; ++k) s += rec[i * N + k] * ms1[k];
    ms[i] = s;
  }
}
template <typename T, int N>
EKSB_HD void bwd_chain_cov(const T* rec, T* Vs) {
  constexpr int NS = Dim<N>::NS;
  T J[N][N], Vs1[NS], JVJ[NS];
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) J[i][j] = rec[i * N + j];
  }
  EKSB_UNROLL
  for (int i = 0; i < NS; ++i) Vs1[i] = Vs[i];
  sandwich<T, N>(J, Vs1, JVJ);
  EKSB_UNROLL
  for (int i = 0; i < NS; ++i) Vs[i] = rec[N * N + i] + JVJ[i];
}
template <typename T, int N>
EKSB_HD void bwd_chain(const T* rec, T* ms, T* Vs) {
  constexpr int NS = Dim<N>::NS;
  T J[N][N], ms1[N], Vs1[NS], JVJ[NS];
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    ms1[i] = ms[i];
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) J[i][j] = rec[i * N + j];
  }
  EKSB_UNROLL
  for (int i = 0; i < NS; ++i) Vs1[i] = Vs[i];
  sandwich<T, N>(J, Vs1, JVJ);
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    T s = rec[N * N + NS + i];
    EKSB_UNROLL
    for (int k = 0; k < N; ++k) s += J[i][k] * ms1[k];
    ms[i] = s;
  }
  EKSB_UNROLL
  for (int i = 0; i < NS; ++i) Vs[i] = rec[N * N + i] + JVJ[i];
}

// ---------------------------------------------------------------------------
// A7: block-tridiagonal Newton system  H x = g  of the EKS posterior (spec:
// ref/Kalman_smoother.pdf p.1, Simulations.ipynb[16-20], eks/block_tridiag.py:15-24):
//   H_tt = B^T D_t^-1 B + [S0^-1 | E^-1] + [A^T E^-1 A | 0 at t=T-1],  H_{t,t+1} = U = -A^T E^-1,
//   g_t  = B^T D_t^-1 y_t (+ S0^-1 m0 at t=0).
// Block-Thomas:  M_t = H_tt - U^T M_{t-1}^-1 U,  z_t = g_t - U^T M_{t-1}^-1 z_{t-1}   (forward chain)
//                x_t = M_t^-1 (z_t - U x_{t+1})                                       (back-substitution)
// and the diagonal blocks of H^-1:  Sig_t = M_t^-1 + G_t Sig_{t+1} G_t^T,  G_t = -M_t^-1 U.
// B^T D^-1 B and B^T D^-1 y are exactly the info record fwd_prep() builds, and the backward
// recursion has the shape of bwd_chain() with (J,L,g) = (G_t, M_t^-1, M_t^-1 z_t), so the solver
// reuses both pipelines.  x equals the RTS mean and Sig the RTS covariance.
// ---------------------------------------------------------------------------
template <typename T, int N> struct BtdConst {
  T iE[Dim<N>::NS];     // E^-1
  T K[Dim<N>::NS];      // A^T E^-1 A
  T Ut[N][N];           // U^T = -E^-1 A
};

template <typename T, int N, int P, int AM>
EKSB_HD void btd_make_const(const Params<T, N, P, AM>& prm, BtdConst<T, N>& c, bool& ok) {
  sym_inv<T, N>(prm.Q, c.iE, ok);
  T EA[N][N];
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) {
      if (AM == 0) {
        EA[i][j] = c.iE[sym_idx<N>(i, j)] * prm.A[j];
      } else {
        T s = T(0);
        EKSB_UNROLL
        for (int k = 0; k < N; ++k) s += c.iE[sym_idx<N>(i, k)] * prm.A[k * N + j];
        EA[i][j] = s;
      }
      c.Ut[i][j] = -EA[i][j];
    }
  }
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = i; j < N; ++j) {
      T s = T(0);
      if (AM == 0) {
        s = prm.A[i] * EA[i][j];
      } else {
        EKSB_UNROLL
        for (int k = 0; k < N; ++k) s += prm.A[k * N + i] * EA[k][j];
      }
      c.K[sym_idx<N>(i, j)] = s;
    }
  }
}

// one forward-elimination step; (z, iM) carry z_{t-1}, M_{t-1}^-1 in and z_t, M_t^-1 out
template <typename T, int N>
EKSB_HD void btd_fwd_chain(const BtdConst<T, N>& c, const T* rec, bool first, bool last, const T* iS0,
                           const T* iS0m0, T* z, T* iM, bool& ok) {
  constexpr int NS = Dim<N>::NS;
  T M[NS], g[N];
  EKSB_UNROLL
  for (int k = 0; k < NS; ++k) M[k] = rec[k] + (first ? iS0[k] : c.iE[k]) + (last ? T(0) : c.K[k]);
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) g[i] = rec[NS + i] + (first ? iS0m0[i] : T(0));
  if (!first) {
    T W[N][N];
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      EKSB_UNROLL
      for (int j = 0; j < N; ++j) {
        T s = T(0);
        EKSB_UNROLL
        for (int k = 0; k < N; ++k) s += c.Ut[i][k] * iM[sym_idx<N>(k, j)];
        W[i][j] = s;
      }
    }
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      EKSB_UNROLL
      for (int j = i; j < N; ++j) {
        T s = T(0);
        EKSB_UNROLL
        for (int k = 0; k < N; ++k) s += W[i][k] * c.Ut[j][k];
        M[sym_idx<N>(i, j)] -= s;
      }
      T s = T(0);
      EKSB_UNROLL
      for (int k = 0; k < N; ++k) s += W[i][k] * z[k];
      g[i] -= s;
    }
  }
  sym_inv<T, N>(M, iM, ok);
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) z[i] = g[i];
}

// (z_t, M_t^-1) -> backward record [ J = G_t | L = M_t^-1 | g = M_t^-1 z_t ]  (J = 0 at the last frame)
template <typename T, int N>
EKSB_HD void btd_bwd_prep(const BtdConst<T, N>& c, const T* z, const T* iM, bool terminal, T* rec) {
  constexpr int NS = Dim<N>::NS;
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) {
      T s = T(0);
      EKSB_UNROLL
      for (int k = 0; k < N; ++k) s -= iM[sym_idx<N>(i, k)] * c.Ut[j][k];
      rec[i * N + j] = terminal ? T(0) : s;
    }
    T gi = T(0);
    EKSB_UNROLL
    for (int k = 0; k < N; ++k) gi += iM[sym_idx<N>(i, k)] * z[k];
    rec[N * N + NS + i] = gi;
  }
  EKSB_UNROLL
  for (int k = 0; k < NS; ++k) rec[N * N + k] = iM[k];
}

// ---------------------------------------------------------------------------
// Parallel-in-time (K4): exact chunked scan for long tracks.  Time is cut into chunks; a chunk's
// effect on the filter is summarised by the element (Phi, b, C, eta, J) of Sarkka &
// Garcia-Fernandez (IEEE TAC 2021): given the state x_s just before the chunk,
//     x_end | x_s, y_chunk ~ N(Phi x_s + b, C),     p(y_chunk | x_s)  ~ exp(-1/2 x_s^T J x_s + eta^T x_s).
// Instead of the generic element combine (a 3x3 inverse + ~8 products per frame) the summary is
// accumulated sequentially inside the chunk by running the ordinary filter from an exactly known
// start (mean 0, covariance 0) and carrying the sensitivity Phi of its mean to x_s:
//     G  = V S^-1            (V, S^-1 from the ordinary step;  I - V I_t = V S^-1)
//     X  = S^-1 - S^-1 V S^-1 = H^T (R + H S H^T)^-1 H          (innovation information)
//     Psi = A Phi;  J += Psi^T X Psi;  eta += Psi^T S^-1 (m_new - A b_old);  Phi = G Psi
// which costs ~2.4x a filter step.  Summaries of consecutive chunks are then folded per track
// (fwd_apply_summary), and every chunk is re-run from its now exact start state.
// The smoother's chunk summary is the (E, g, L) element: ms[t0] = E ms[t1] + g, Vs[t0] = E Vs[t1] E^T + L.
// ---------------------------------------------------------------------------
template <int N> struct ScanRec {
  static constexpr int NS = Dim<N>::NS;
  static constexpr int FWD = N * N + N + NS + N + NS;   // Phi | b | C | eta | J
  static constexpr int BWD = N * N + NS + N;            // E | L | g   (same order as Rec<N>::BWD)
};

// One frame of the forward chunk summary.  (b, C) are the running filter state started at (0, 0);
// Phi starts at I, eta and J at 0.
template <typename T, int N, int P, int AM>
EKSB_HD void fwd_summary_step(const Params<T, N, P, AM>& prm, const T* rec, T (*Phi)[N], T* b, T* C, T* eta, T* J,
                              bool& ok) {
  constexpr int NS = Dim<N>::NS;
  T mp[N], S[NS], u[N], Vu[N];
  predict<T, N, P, AM>(prm, b, C, mp, S);
  sym_matvec<T, N>(rec, mp, u);
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) u[i] = rec[NS + i] - u[i];
  // Psi = A Phi
  T Psi[N][N];
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) {
      if (AM == 0) {
        Psi[i][j] = prm.A[i] * Phi[i][j];
      } else {
        T s = T(0);
        EKSB_UNROLL
        for (int k = 0; k < N; ++k) s += prm.A[i * N + k] * Phi[k][j];
        Psi[i][j] = s;
      }
    }
  }
  // Everything below hangs off ONE matrix, Mi = (Id + Lam S)^-1 (Lam = rec[0..NS)):
  //   V = S Mi,   S^-1 V = Mi,   G = V S^-1 = Mi^T,   X = S^-1 - S^-1 V S^-1 = Mi Lam   (symmetric)
  // so the step needs no S^-1 and a single reciprocal (N <= 3: adjugate; else via two symmetric inverses).
  T Mi[N][N], V[NS], X[NS];
  if (N <= 3) {
    T M[N][N], adj[N][N], det;
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      EKSB_UNROLL
      for (int j = 0; j < N; ++j) {
        T a = (i == j) ? T(1) : T(0);
        EKSB_UNROLL
        for (int k = 0; k < N; ++k) a += rec[sym_idx<N>(i, k)] * S[sym_idx<N>(k, j)];
        M[i][j] = a;
      }
    }
    gen_adjugate<T, N>(M, adj, det);
    ok = ok && !(det <= T(0)) && !(S[0] <= T(0));
    const T r = rcp(det);
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      EKSB_UNROLL
      for (int j = 0; j < N; ++j) Mi[i][j] = adj[i][j] * r;
    }
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      EKSB_UNROLL
      for (int j = i; j < N; ++j) {
        T a = T(0), x = T(0);
        EKSB_UNROLL
        for (int k = 0; k < N; ++k) {
          a += S[sym_idx<N>(i, k)] * Mi[k][j];
          x += Mi[i][k] * rec[sym_idx<N>(k, j)];
        }
        V[sym_idx<N>(i, j)] = a;
        X[sym_idx<N>(i, j)] = x;
      }
    }
  } else {
    T Sinv[NS], Lam[NS];
    sym_inv<T, N>(S, Sinv, ok);
    EKSB_UNROLL
    for (int k = 0; k < NS; ++k) Lam[k] = Sinv[k] + rec[k];
    sym_inv<T, N>(Lam, V, ok);
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      EKSB_UNROLL
      for (int j = 0; j < N; ++j) {
        T a = T(0);
        EKSB_UNROLL
        for (int k = 0; k < N; ++k) a += Sinv[sym_idx<N>(i, k)] * V[sym_idx<N>(k, j)];
        Mi[i][j] = a;
      }
    }
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      EKSB_UNROLL
      for (int j = i; j < N; ++j) {
        T a = Sinv[sym_idx<N>(i, j)];
        EKSB_UNROLL
        for (int k = 0; k < N; ++k) a -= Mi[i][k] * Sinv[sym_idx<N>(k, j)];
        X[sym_idx<N>(i, j)] = a;
      }
    }
  }
  sym_matvec<T, N>(V, u, Vu);
  // J += Psi^T X Psi ; eta += Psi^T (S^-1 V u) = Psi^T (Mi u)
  T XP[N][N], w[N];
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    T wi = T(0);
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) {
      T a = T(0);
      EKSB_UNROLL
      for (int k = 0; k < N; ++k) a += X[sym_idx<N>(i, k)] * Psi[k][j];
      XP[i][j] = a;
      wi += Mi[i][j] * u[j];
    }
    w[i] = wi;
  }
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = i; j < N; ++j) {
      T a = T(0);
      EKSB_UNROLL
      for (int k = 0; k < N; ++k) a += Psi[k][i] * XP[k][j];
      J[sym_idx<N>(i, j)] += a;
    }
    T e = T(0);
    EKSB_UNROLL
    for (int k = 0; k < N; ++k) e += Psi[k][i] * w[k];
    eta[i] += e;
  }
  // Phi = G Psi = Mi^T Psi ;  state update
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) {
      T a = T(0);
      EKSB_UNROLL
      for (int k = 0; k < N; ++k) a += Mi[k][i] * Psi[k][j];
      Phi[i][j] = a;
    }
    b[i] = mp[i] + Vu[i];
  }
  EKSB_UNROLL
  for (int k = 0; k < NS; ++k) C[k] = V[k];
}

// The sensitivity-only part of fwd_summary_step: the ordinary filter step on (b, C) plus Phi <- G A Phi, without
// the likelihood terms (J, eta).  Enough to find the frame after which a chunk's running state no longer depends on
// the unknown state before the chunk (the warm-up kernels of the time-split pipelined scan, eksb_scan.cuh).
template <typename T, int N, int P, int AM>
EKSB_HD void fwd_phi_step(const Params<T, N, P, AM>& prm, const T* rec, T (*Phi)[N], T* b, T* C, bool& ok) {
  constexpr int NS = Dim<N>::NS;
  T mp[N], S[NS], u[N], Vu[N];
  predict<T, N, P, AM>(prm, b, C, mp, S);
  sym_matvec<T, N>(rec, mp, u);
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) u[i] = rec[NS + i] - u[i];
  T Psi[N][N];
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) {
      if (AM == 0) {
        Psi[i][j] = prm.A[i] * Phi[i][j];
      } else {
        T s = T(0);
        EKSB_UNROLL
        for (int k = 0; k < N; ++k) s += prm.A[i * N + k] * Phi[k][j];
        Psi[i][j] = s;
      }
    }
  }
  T Mi[N][N], V[NS];                    // Mi = (Id + Lam S)^-1 ;  V = S Mi ;  G = V S^-1 = Mi^T
  if (N <= 3) {
    T M[N][N], adj[N][N], det;
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      EKSB_UNROLL
      for (int j = 0; j < N; ++j) {
        T a = (i == j) ? T(1) : T(0);
        EKSB_UNROLL
        for (int k = 0; k < N; ++k) a += rec[sym_idx<N>(i, k)] * S[sym_idx<N>(k, j)];
        M[i][j] = a;
      }
    }
    gen_adjugate<T, N>(M, adj, det);
    ok = ok && !(det <= T(0)) && !(S[0] <= T(0));
    const T r = rcp(det);
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      EKSB_UNROLL
      for (int j = 0; j < N; ++j) Mi[i][j] = adj[i][j] * r;
    }
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      EKSB_UNROLL
      for (int j = i; j < N; ++j) {
        T a = T(0);
        EKSB_UNROLL
        for (int k = 0; k < N; ++k) a += S[sym_idx<N>(i, k)] * Mi[k][j];
        V[sym_idx<N>(i, j)] = a;
      }
    }
  } else {
    T Sinv[NS], Lam[NS];
    sym_inv<T, N>(S, Sinv, ok);
    EKSB_UNROLL
    for (int k = 0; k < NS; ++k) Lam[k] = Sinv[k] + rec[k];
    sym_inv<T, N>(Lam, V, ok);
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      EKSB_UNROLL
      for (int j = 0; j < N; ++j) {
        T a = T(0);
        EKSB_UNROLL
        for (int k = 0; k < N; ++k) a += Sinv[sym_idx<N>(i, k)] * V[sym_idx<N>(k, j)];
        Mi[i][j] = a;
      }
    }
  }
  sym_matvec<T, N>(V, u, Vu);
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) {
      T a = T(0);
      EKSB_UNROLL
      for (int k = 0; k < N; ++k) a += Mi[k][i] * Psi[k][j];
      Phi[i][j] = a;
    }
    b[i] = mp[i] + Vu[i];
  }
  EKSB_UNROLL
  for (int k = 0; k < NS; ++k) C[k] = V[k];
}

// Fold a chunk summary into the filtered state just before the chunk: (m,P) -> state at the chunk's end.
template <typename T, int N>
EKSB_HD void fwd_apply_summary(const T* sum, T* m, T* P, bool& ok) {
  constexpr int NS = Dim<N>::NS;
  const T* Phi = sum;                 // N*N row-major
  const T* b = sum + N * N;
  const T* C = b + N;
  const T* eta = C + NS;
  const T* J = eta + N;
  T Pinv[NS], Lam[NS], P1[NS], r[N], m1[N], Jm[N];
  sym_inv<T, N>(P, Pinv, ok);
  EKSB_UNROLL
  for (int k = 0; k < NS; ++k) Lam[k] = Pinv[k] + J[k];
  sym_inv<T, N>(Lam, P1, ok);         // cov of x_s given the chunk's data
  sym_matvec<T, N>(J, m, Jm);
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) r[i] = eta[i] - Jm[i];
  sym_matvec<T, N>(P1, r, m1);
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) m1[i] += m[i];
  T Phim[N][N];
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    T s = b[i];
    EKSB_UNROLL
    for (int k = 0; k < N; ++k) { s += Phi[i * N + k] * m1[k]; Phim[i][k] = Phi[i * N + k]; }
    m[i] = s;
  }
  T PPP[NS];
  sandwich<T, N>(Phim, P1, PPP);
  EKSB_UNROLL
  for (int k = 0; k < NS; ++k) P[k] = PPP[k] + C[k];
}

// Compose one more (earlier) frame element (J,L,g) onto the chunk's smoothing summary (E,L,g).
template <typename T, int N>
EKSB_HD void bwd_summary_step(const T* rec, T* acc) {
  constexpr int NS = Dim<N>::NS;
  T Jm[N][N], En[N][N];
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) Jm[i][j] = rec[i * N + j];
  }
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) {
      T s = T(0);
      EKSB_UNROLL
      for (int k = 0; k < N; ++k) s += Jm[i][k] * acc[k * N + j];
      En[i][j] = s;
    }
  }
  bwd_chain<T, N>(rec, acc + N * N + NS, acc + N * N);     // g = g_t + J g_acc ; L = L_t + J L_acc J^T
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) acc[i * N + j] = En[i][j];
  }
}

// ---------------------------------------------------------------------------
// Chunk programs of the scan (one warp per (track tile, chunk) on the device, one call per chunk in
// the host tests).  Both stop paying for the summary as soon as it has become irrelevant:
// the filter forgets its start exponentially, so after K frames the sensitivity Phi of the chunk's
// running state to the unknown start state is numerically zero (|Phi| < ScanCut, checked every 16
// frames for the whole warp).  From there on (b, C) IS the filtered state, so the rest of the chunk
// is run as the ordinary filter and its records are written out at once; only the first K frames of
// the chunk have to be re-run after the fold.  The same holds for the smoother's E (product of the
// RTS gains).  HBM traffic drops from (2 reads of the input + 1 write) to (1 + K/L reads + 1 write).
// ---------------------------------------------------------------------------
template <typename T> struct ScanCut;
// State error after the cut <= ScanCut * |x_start| (x_start = the unknown state before the chunk, of the same magnitude
// as the states that follow it): a tenth of an ulp of the state in either type.  Round 1 used 1e-22 / 1e-11; every
// decade costs ~20 warm-up frames per chunk on the shapes of BASELINE.json (K = 480 -> 370 in f64).
template <> struct ScanCut<double> { static constexpr double value = 1e-17; };
template <> struct ScanCut<float> { static constexpr float value = 1e-8f; };

#if defined(__CUDA_ARCH__)
#define EKSB_WARP_ALL(pred) __all_sync(0xffffffffu, (pred))
#else
#define EKSB_WARP_ALL(pred) (pred)
#endif

// Forward summary of one chunk of `n` frames.  obs / filt point at the chunk's first frame.
// sum (ScanRec<N>::FWD values, local) receives the element.  Returns the cut k: frames [k, n) of the
// chunk have been written to filt (k == n: nothing written, the whole chunk must be re-run).
// Tail = the ordinary filter over the rest of the chunk: bool tail(prm, obs, filt_or_null, n, m, V).  The default
// is kf_forward_range with a register prefetch (host tests, re-run kernels); the summary kernels pass a variant that
// stages the frames through shared memory (eksb_scan.cuh FwdTailRing) -- same arithmetic.
template <typename T, int N, int P, int AM, int LS, int PF> struct FwdTailPlain {
  EKSB_HD bool operator()(const Params<T, N, P, AM>& prm, const T* obs, T* filt, long n, T* m, T* V, const T* mu,
                          bool* nonpos) const {
    double acc = 0.0;
    return kf_forward_range<T, N, P, AM, LS, PF, false>(prm, obs, filt, n, kContinue, m, V, acc, mu, nonpos);
  }
};
template <typename T, int N, int P, int AM, int LS, int PF> struct BwdTailPlain {
  EKSB_HD bool operator()(const Params<T, N, P, AM>& prm, const T* filt, T* smth, T* ym, long n, T* ms, T* Vs) const {
    return rts_backward_range<T, N, P, AM, LS, PF>(prm, filt, smth, ym, nullptr, nullptr, n, false, ms, Vs);
  }
};

template <typename T, int N, int P, int AM, int LS, typename Tail>
EKSB_HD long scan_fwd_chunk(const Params<T, N, P, AM>& prm, const T* __restrict__ obs, T* __restrict__ filt, long n,
                            bool early_out, T* sum, bool& ok, const Tail& tail, const T* mu = nullptr,
                            bool* nonpos = nullptr) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, R2 = 2 * P;
  T Phi[N][N], bb[N], Cc[NS], eta[N], J[NS];
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    bb[i] = eta[i] = T(0);
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) Phi[i][j] = (i == j) ? T(1) : T(0);
  }
  EKSB_UNROLL
  for (int i = 0; i < NS; ++i) Cc[i] = J[i] = T(0);
  T nxt[R2];
  EKSB_UNROLL
  for (int r = 0; r < R2; ++r) nxt[r] = obs[r * LS];
  long t = 0;
  bool dead = false;
  while (t < n) {
    T cur[R2], rec[Rec<N>::FWD];
    EKSB_UNROLL
    for (int r = 0; r < R2; ++r) cur[r] = nxt[r];
    const long f = (t + 1 < n) ? t + 1 : n - 1;
    EKSB_UNROLL
    for (int r = 0; r < R2; ++r) nxt[r] = obs[(f * R2 + r) * LS];
    if (mu) {
      EKSB_UNROLL
      for (int j = 0; j < P; ++j) cur[j] -= mu[j];
    }
    if (nonpos) *nonpos = *nonpos || any_nonpos<T, P>(cur + P);
    fwd_prep<T, N, P, AM>(prm, cur, cur + P, rec);
    fwd_summary_step<T, N, P, AM>(prm, rec, Phi, bb, Cc, eta, J, ok);
    ++t;
    if ((t & 15) == 0) {
      T mx = T(0);
      EKSB_UNROLL
      for (int i = 0; i < N; ++i) {
        EKSB_UNROLL
        for (int j = 0; j < N; ++j) mx = fmax(mx, fabs(Phi[i][j]));
      }
      if (EKSB_WARP_ALL(mx < ScanCut<T>::value)) { dead = true; break; }
    }
  }
  long cut = n;
  if (dead) {
    EKSB_UNROLL
    for (int i = 0; i < N; ++i) {
      EKSB_UNROLL
      for (int j = 0; j < N; ++j) Phi[i][j] = T(0);
    }
    if (t < n) {
      if (early_out) cut = t;
      ok = tail(prm, obs + t * R2 * LS, early_out ? filt + t * SG * LS : nullptr, n - t, bb, Cc, mu, nonpos) && ok;
    }
  }
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) {
    EKSB_UNROLL
    for (int j = 0; j < N; ++j) sum[i * N + j] = Phi[i][j];
    sum[N * N + i] = bb[i];
    sum[N * N + N + NS + i] = eta[i];
  }
  EKSB_UNROLL
  for (int i = 0; i < NS; ++i) {
    sum[N * N + N + i] = Cc[i];
    sum[N * N + N + NS + N + i] = J[i];
  }
  return cut;
}

// Backward (smoother) summary of one chunk of `n` frames, last frame first; filt / smth / ym point at
// the chunk's first frame (smth may alias filt: a frame is read before it is overwritten, and frames
// >= the returned cut are left untouched for the re-run).  acc (ScanRec<N>::BWD values, local) receives
// (E, L, g).  Returns the cut k: frames [0, k) of the chunk have been smoothed and written (k == 0:
// nothing written).
template <typename T, int N, int P, int AM, int LS, typename Tail>
EKSB_HD long scan_bwd_chunk(const Params<T, N, P, AM>& prm, const T* filt, T* smth, T* ym, long n, bool early_out,
                            T* acc, bool& ok, const Tail& tail) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA, SR = ScanRec<N>::BWD;
  EKSB_UNROLL
  for (int i = 0; i < SR; ++i) acc[i] = T(0);
  EKSB_UNROLL
  for (int i = 0; i < N; ++i) acc[i * N + i] = T(1);
  T nxt[SG];
  EKSB_UNROLL
  for (int r = 0; r < SG; ++r) nxt[r] = filt[((n - 1) * SG + r) * LS];
  long t = n - 1, done = 0;
  bool dead = false;
  while (t >= 0) {
    T cur[SG], rec[Rec<N>::BWD];
    EKSB_UNROLL
    for (int r = 0; r < SG; ++r) cur[r] = nxt[r];
    const long f = (t - 1 >= 0) ? t - 1 : 0;
    EKSB_UNROLL
    for (int r = 0; r < SG; ++r) nxt[r] = filt[(f * SG + r) * LS];
    bwd_prep<T, N, P, AM>(prm, cur, cur + N, rec, ok);
    bwd_summary_step<T, N>(rec, acc);
    --t;
    ++done;
    if (early_out && (done & 15) == 0) {
      T mx = T(0);
      EKSB_UNROLL
      for (int i = 0; i < N * N; ++i) mx = fmax(mx, fabs(acc[i]));
      if (EKSB_WARP_ALL(mx < ScanCut<T>::value)) { dead = true; break; }
    }
  }
  long cut = 0;
  if (dead && t >= 0) {
    cut = t + 1;
    EKSB_UNROLL
    for (int i = 0; i < N * N; ++i) acc[i] = T(0);
    ok = tail(prm, filt, smth, ym, cut, acc + N * N + NS, acc + N * N) && ok;
  }
  return cut;
}

}  // namespace eksb

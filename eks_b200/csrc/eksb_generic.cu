// eksb_generic.cu -- runtime-dimension kernels that follow the reference's arithmetic
// literally (p x p LU solve with partial pivoting, like numpy.linalg.solve -> LAPACK gesv).
// They cover what the register-resident template kernels do not: a dense observation-noise
// matrix R with non-zero off-diagonals (the reference's signature allows it,
// eks/ensemble_kalman.py:112), (n,p) pairs that are not instantiated, and kalman_dot() itself.
// Same track-tile layouts as the fast kernels; one track per thread, small matrices in local
// memory.  Not a hot path: every shipped config has diagonal R and n=3.
#include "eksb_common.cuh"

namespace eksb {

constexpr int kNMax = 8, kPMax = 16;

// Solve M X = B in place (M: p x p, B: p x k, row-major, leading dims ldm / ldb) by Gaussian
// elimination with partial pivoting.  Returns false on an exactly zero pivot (LAPACK info > 0).
template <typename T>
__device__ bool lu_solve(T* M, int ldm, T* Bm, int ldb, int p, int k) {
  for (int c = 0; c < p; ++c) {
    int piv = c;
    T best = fabs(M[c * ldm + c]);
    for (int r = c + 1; r < p; ++r) {
      T v = fabs(M[r * ldm + c]);
      if (v > best) { best = v; piv = r; }
    }
    if (best == T(0)) return false;
    if (piv != c) {
      for (int j = 0; j < p; ++j) { T t = M[c * ldm + j]; M[c * ldm + j] = M[piv * ldm + j]; M[piv * ldm + j] = t; }
      for (int j = 0; j < k; ++j) { T t = Bm[c * ldb + j]; Bm[c * ldb + j] = Bm[piv * ldb + j]; Bm[piv * ldb + j] = t; }
    }
    const T inv = T(1) / M[c * ldm + c];
    for (int r = c + 1; r < p; ++r) {
      const T f = M[r * ldm + c] * inv;
      if (f != T(0)) {
        for (int j = c + 1; j < p; ++j) M[r * ldm + j] -= f * M[c * ldm + j];
        for (int j = 0; j < k; ++j) Bm[r * ldb + j] -= f * Bm[c * ldb + j];
      }
    }
  }
  for (int c = p - 1; c >= 0; --c) {
    const T inv = T(1) / M[c * ldm + c];
    for (int j = 0; j < k; ++j) {
      T s = Bm[c * ldb + j];
      for (int r = c + 1; r < p; ++r) s -= M[c * ldm + r] * Bm[r * ldb + j];
      Bm[c * ldb + j] = s * inv;
    }
  }
  return true;
}

// out (n x k) = V C^T (R + C V C^T)^-1 arr      (eks/ensemble_kalman.py:110-117)
template <typename T>
__device__ bool kalman_dot_dev(const T* arr, int k, const T* V, const T* C, const T* R, int n, int p, T* out) {
  T sig[kPMax * kPMax], rhs[kPMax * kNMax], cv[kPMax * kNMax];
  for (int i = 0; i < p; ++i)
    for (int j = 0; j < n; ++j) {
      T s = T(0);
      for (int l = 0; l < n; ++l) s += C[i * n + l] * V[l * n + j];
      cv[i * n + j] = s;                       // C V
    }
  for (int i = 0; i < p; ++i)
    for (int j = 0; j < p; ++j) {
      T s = R[i * p + j];
      for (int l = 0; l < n; ++l) s += cv[i * n + l] * C[j * n + l];
      sig[i * p + j] = s;                      // R + C V C^T
    }
  for (int i = 0; i < p * k; ++i) rhs[i] = arr[i];
  if (!lu_solve(sig, p, rhs, k, p, k)) return false;
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < k; ++j) {
      T s = T(0);
      for (int a = 0; a < n; ++a) {            // V (C^T rhs)
        T ct = T(0);
        for (int l = 0; l < p; ++l) ct += C[l * n + a] * rhs[l * k + j];
        s += V[i * n + a] * ct;
      }
      out[i * k + j] = s;
    }
  return true;
}

template <typename T>
__global__ void kalman_dot_kernel(const T* arr, int k, const T* V, const T* C, const T* R, int n, int p, T* out, int* status) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    bool ok = kalman_dot_dev(arr, k, V, C, R, n, p, out);
    if (status) *status = ok ? 0 : 1;
  }
}

// Forward filter, literal form.  Rfull: (B,p,p) or shared (p,p) per r_shared; its diagonal is
// replaced by the frame's ensemble variance (eks/ensemble_kalman.py:88-89,99-100).
template <typename T>
__global__ void __launch_bounds__(kLanes)
kf_forward_generic_kernel(const T* __restrict__ obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q,
                          const T* Rfull, int r_shared, int shared_mask, long B, long Tn, int n, int p, int first,
                          T* __restrict__ filt, int* status) {
  const long g = blockIdx.x;
  const int lane = threadIdx.x;
  const long b = g * kLanes + lane;
  const long bp = b < B ? b : B - 1;
  const int ns = n * (n + 1) / 2, sg = n + ns;
  const T* c = C + ((shared_mask & EKSB_SHARED_C) ? 0 : bp * p * n);
  const T* a = A + ((shared_mask & EKSB_SHARED_A) ? 0 : bp * n * n);
  const T* q = Q + ((shared_mask & EKSB_SHARED_Q) ? 0 : bp * n * n);
  const T* pm = m0 + ((shared_mask & EKSB_SHARED_M0) ? 0 : bp * n);
  const T* ps = S0 + ((shared_mask & EKSB_SHARED_S0) ? 0 : bp * n * n);
  T R[kPMax * kPMax], m[kNMax], V[kNMax * kNMax], S[kNMax * kNMax], tmp[kNMax * kNMax];
  T y[kPMax], arr[kPMax * kNMax], out[kNMax * kNMax], mp[kNMax];
  for (int i = 0; i < p * p; ++i) R[i] = Rfull ? Rfull[(r_shared ? 0 : bp * p * p) + i] : ((i / p == i % p) ? T(1) : T(0));
  for (int i = 0; i < n; ++i) m[i] = pm[i];
  for (int i = 0; i < n * n; ++i) V[i] = ps[i];
  bool ok = true;
  const T* o = obs + (g * Tn * (2 * p)) * kLanes + lane;
  T* f = filt + (g * Tn * sg) * kLanes + lane;
  for (long t = 0; t < Tn; ++t) {
    for (int j = 0; j < p; ++j) { y[j] = o[(t * 2 * p + j) * kLanes]; R[j * p + j] = o[(t * 2 * p + p + j) * kLanes]; }
    if (t == 0 && first == kPriorOnly) {
      // Newton variant: no update at frame 0; P = inv(S0) via the same LU
      for (int i = 0; i < n * n; ++i) { tmp[i] = V[i]; S[i] = (i / n == i % n) ? T(1) : T(0); }
      ok = lu_solve(tmp, n, S, n, n, n) && ok;
      for (int i = 0; i < n * n; ++i) V[i] = S[i];
    } else {
      if (t == 0) {
        for (int i = 0; i < n; ++i) mp[i] = m[i];
        for (int i = 0; i < n * n; ++i) S[i] = V[i];
      } else {
        for (int i = 0; i < n; ++i) { T s = T(0); for (int l = 0; l < n; ++l) s += a[i * n + l] * m[l]; mp[i] = s; }
        for (int i = 0; i < n; ++i)
          for (int j = 0; j < n; ++j) { T s = T(0); for (int l = 0; l < n; ++l) s += a[i * n + l] * V[l * n + j]; tmp[i * n + j] = s; }
        for (int i = 0; i < n; ++i)
          for (int j = 0; j < n; ++j) { T s = q[i * n + j]; for (int l = 0; l < n; ++l) s += tmp[i * n + l] * a[j * n + l]; S[i * n + j] = s; }
      }
      for (int j = 0; j < p; ++j) { T s = y[j]; for (int l = 0; l < n; ++l) s -= c[j * n + l] * mp[l]; arr[j] = s; }
      ok = kalman_dot_dev(arr, 1, S, c, R, n, p, out) && ok;
      for (int i = 0; i < n; ++i) m[i] = mp[i] + out[i];
      for (int j = 0; j < p; ++j)
        for (int i = 0; i < n; ++i) { T s = T(0); for (int l = 0; l < n; ++l) s += c[j * n + l] * S[l * n + i]; arr[j * n + i] = s; }
      ok = kalman_dot_dev(arr, n, S, c, R, n, p, out) && ok;
      for (int i = 0; i < n * n; ++i) V[i] = S[i] - out[i];
    }
    for (int i = 0; i < n; ++i) f[(t * sg + i) * kLanes] = m[i];
    int kk = 0;
    for (int i = 0; i < n; ++i)
      for (int j = i; j < n; ++j) f[(t * sg + n + kk++) * kLanes] = T(0.5) * (V[i * n + j] + V[j * n + i]);
  }
  if (b < B && status) set_status(status + b, ok, false);
}

// RTS smoother, literal form (eks/ensemble_kalman.py:152-162), any n <= 8, p <= 16, dense A:
//   J = solve(S[t], A Vf[t])^T (:158, LU with partial pivoting), Vs[t] = Vf[t] + J (Vs[t+1] - S[t]) J^T (:160),
//   ms[t] = mf[t] + J (ms[t+1] - A mf[t]) (:161), CV[t] = Vs[t+1] J^T (:162).
// pred (optional): NS-row packed records holding the caller's S[t] (the reference reads its S argument and never
// touches Q); NULL recomputes S[t] = A Vf[t] A^T + Q, which is what filtering_pass stored there (:101).
template <typename T>
__global__ void __launch_bounds__(kLanes)
rts_backward_generic_kernel(const T* filt, const T* __restrict__ pred, const T* C, const T* A, const T* Q, int shared_mask,
                            long B, long Tn, int n, int p, T* smth, T* __restrict__ ym, T* __restrict__ yv,
                            T* __restrict__ cv, int* status) {
  const long g = blockIdx.x;
  const int lane = threadIdx.x;
  const long b = g * kLanes + lane;
  const long bp = b < B ? b : B - 1;
  const int ns = n * (n + 1) / 2, sg = n + ns;
  const T* c = C + ((shared_mask & EKSB_SHARED_C) ? 0 : bp * p * n);
  const T* a = A + ((shared_mask & EKSB_SHARED_A) ? 0 : bp * n * n);
  const T* q = Q + ((shared_mask & EKSB_SHARED_Q) ? 0 : bp * n * n);
  T m[kNMax], V[kNMax * kNMax], S[kNMax * kNMax], AV[kNMax * kNMax], J[kNMax * kNMax], W[kNMax * kNMax];
  T ms[kNMax], Vs[kNMax * kNMax], ms1[kNMax], Vs1[kNMax * kNMax], mp[kNMax];
  bool ok = true;
  const T* f = filt + (g * Tn * sg) * kLanes + lane;
  T* o = smth + (g * Tn * sg) * kLanes + lane;
  auto unpack = [&](const T* rec, long t, int rows, int r0, T* full) {
    int kk = 0;
    for (int i = 0; i < n; ++i)
      for (int j = i; j < n; ++j) { const T v = rec[(t * rows + r0 + kk++) * kLanes]; full[i * n + j] = v; full[j * n + i] = v; }
  };
  for (long t = Tn - 1; t >= 0; --t) {
    for (int i = 0; i < n; ++i) m[i] = f[(t * sg + i) * kLanes];
    unpack(f, t, sg, n, V);
    if (t == Tn - 1) {                                       // :152-153
      for (int i = 0; i < n; ++i) ms[i] = m[i];
      for (int i = 0; i < n * n; ++i) Vs[i] = V[i];
    } else {
      for (int i = 0; i < n; ++i) { ms1[i] = ms[i]; T s = T(0); for (int l = 0; l < n; ++l) s += a[i * n + l] * m[l]; mp[i] = s; }
      for (int i = 0; i < n * n; ++i) Vs1[i] = Vs[i];
      for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) { T s = T(0); for (int l = 0; l < n; ++l) s += a[i * n + l] * V[l * n + j]; AV[i * n + j] = s; }
      if (pred) {
        unpack(pred + (g * Tn * ns) * kLanes + lane, t, ns, 0, S);
      } else {
        for (int i = 0; i < n; ++i)
          for (int j = 0; j < n; ++j) { T s = q[i * n + j]; for (int l = 0; l < n; ++l) s += AV[i * n + l] * a[j * n + l]; S[i * n + j] = s; }
      }
      for (int i = 0; i < n * n; ++i) { W[i] = S[i]; J[i] = AV[i]; }
      ok = lu_solve(W, n, J, n, n, n) && ok;                 // J <- S^-1 (A V);  the gain is its transpose
      // D = Vs1 - S ;  Vs = V + J^T' ... with G = J^T: Vs = V + G D G^T, ms = m + G (ms1 - mp), CV = Vs1 G^T
      for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) {                        // W = G D,  G[i][l] = J[l][i]
          T s = T(0);
          for (int l = 0; l < n; ++l) s += J[l * n + i] * (Vs1[l * n + j] - S[l * n + j]);
          W[i * n + j] = s;
        }
      for (int i = 0; i < n; ++i) {
        for (int j = 0; j < n; ++j) { T s = V[i * n + j]; for (int l = 0; l < n; ++l) s += W[i * n + l] * J[l * n + j]; Vs[i * n + j] = s; }
        T s = m[i];
        for (int l = 0; l < n; ++l) s += J[l * n + i] * (ms1[l] - mp[l]);
        ms[i] = s;
      }
      if (cv)
        for (int i = 0; i < n; ++i)
          for (int j = 0; j < n; ++j) {                      // CV = Vs1 G^T = Vs1 J
            T s = T(0);
            for (int l = 0; l < n; ++l) s += Vs1[i * n + l] * J[l * n + j];
            cv[((g * Tn + t) * (n * n) + i * n + j) * kLanes + lane] = s;
          }
    }
    for (int i = 0; i < n; ++i) o[(t * sg + i) * kLanes] = ms[i];
    int kk = 0;
    for (int i = 0; i < n; ++i)
      for (int j = i; j < n; ++j) o[(t * sg + n + kk++) * kLanes] = T(0.5) * (Vs[i * n + j] + Vs[j * n + i]);
    if (ym || yv)
      for (int j = 0; j < p; ++j) {
        T s = T(0), qv = T(0);
        for (int i = 0; i < n; ++i) {
          s += c[j * n + i] * ms[i];
          T acc = T(0);
          for (int l = 0; l < n; ++l) acc += Vs[i * n + l] * c[j * n + l];
          qv += c[j * n + i] * acc;
        }
        if (ym) ym[((g * Tn + t) * p + j) * kLanes + lane] = s;
        if (yv) yv[((g * Tn + t) * p + j) * kLanes + lane] = qv;
      }
  }
  if (b < B && status) set_status(status + b, ok, false);
}

template <typename T>
static int kalman_dot_api(const T* arr, int k, const T* V, const T* C, const T* R, int n, int p, T* out, int* status, void* stream) {
  if (!arr || !V || !C || !R || !out || n < 1 || p < 1 || k < 1 || n > kNMax || p > kPMax || k > kNMax) {
    set_error("kalman_dot: bad argument (n<=%d, p<=%d, k<=%d)", kNMax, kPMax, kNMax);
    return EKSB_EARG;
  }
  kalman_dot_kernel<T><<<1, 32, 0, (cudaStream_t)stream>>>(arr, k, V, C, R, n, p, out, status);
  return check_launch("kalman_dot_kernel");
}

template <typename T>
static int kf_forward_generic_api(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, const T* R,
                                  int r_shared, int shared_mask, long B, long Tn, int n, int p, int first, T* filt,
                                  int* status, void* stream) {
  if (!obs || !m0 || !S0 || !C || !A || !Q || !filt || B <= 0 || Tn <= 0) { set_error("kf_forward_generic: bad argument"); return EKSB_EARG; }
  if (n < 1 || p < 1 || n > kNMax || p > kPMax) { set_error("kf_forward_generic: need n<=%d, p<=%d", kNMax, kPMax); return EKSB_EDIM; }
  if (first != kPriorUpdate && first != kPriorOnly) { set_error("kf_forward_generic: bad first-frame mode"); return EKSB_EARG; }
  kf_forward_generic_kernel<T><<<(unsigned)num_tiles(B), kLanes, 0, (cudaStream_t)stream>>>(
      obs, m0, S0, C, A, Q, R, r_shared, shared_mask, B, Tn, n, p, first, filt, status);
  return check_launch("kf_forward_generic_kernel");
}

template <typename T>
static int rts_backward_generic_api(const T* filt, const T* pred, const T* C, const T* A, const T* Q, int shared_mask, long B,
                                    long Tn, int n, int p, T* smth, T* ym, T* yv, T* cv, int* status, void* stream) {
  if (!filt || !smth || !C || !A || !Q || B <= 0 || Tn <= 0) { set_error("rts_backward_generic: bad argument"); return EKSB_EARG; }
  if (n < 1 || p < 1 || n > kNMax || p > kPMax) { set_error("rts_backward_generic: need n<=%d, p<=%d", kNMax, kPMax); return EKSB_EDIM; }
  rts_backward_generic_kernel<T><<<(unsigned)num_tiles(B), kLanes, 0, (cudaStream_t)stream>>>(
      filt, pred, C, A, Q, shared_mask, B, Tn, n, p, smth, ym, yv, cv, status);
  return check_launch("rts_backward_generic_kernel");
}

}  // namespace eksb

using namespace eksb;

extern "C" int eksb_rts_backward_generic_f64(const double* filt, const double* pred, const double* C, const double* A,
                                             const double* Q, int shared_mask, long B, long T, int n, int p, double* smth,
                                             double* ym, double* yv, double* cv, int* status, void* stream) {
  return rts_backward_generic_api<double>(filt, pred, C, A, Q, shared_mask, B, T, n, p, smth, ym, yv, cv, status, stream);
}
extern "C" int eksb_rts_backward_generic_f32(const float* filt, const float* pred, const float* C, const float* A,
                                             const float* Q, int shared_mask, long B, long T, int n, int p, float* smth,
                                             float* ym, float* yv, float* cv, int* status, void* stream) {
  return rts_backward_generic_api<float>(filt, pred, C, A, Q, shared_mask, B, T, n, p, smth, ym, yv, cv, status, stream);
}

extern "C" int eksb_kalman_dot_f64(const double* array, int k, const double* V, const double* C, const double* R, int n,
                                   int p, double* out, int* status, void* stream) {
  return kalman_dot_api<double>(array, k, V, C, R, n, p, out, status, stream);
}
extern "C" int eksb_kalman_dot_f32(const float* array, int k, const float* V, const float* C, const float* R, int n, int p,
                                   float* out, int* status, void* stream) {
  return kalman_dot_api<float>(array, k, V, C, R, n, p, out, status, stream);
}
extern "C" int eksb_kf_forward_generic_f64(const double* obs, const double* m0, const double* S0, const double* C,
                                           const double* A, const double* Q, const double* R, int r_shared,
                                           int shared_mask, long B, long T, int n, int p, int first, double* filt,
                                           int* status, void* stream) {
  return kf_forward_generic_api<double>(obs, m0, S0, C, A, Q, R, r_shared, shared_mask, B, T, n, p, first, filt, status, stream);
}
extern "C" int eksb_kf_forward_generic_f32(const float* obs, const float* m0, const float* S0, const float* C,
                                           const float* A, const float* Q, const float* R, int r_shared, int shared_mask,
                                           long B, long T, int n, int p, int first, float* filt, int* status,
                                           void* stream) {
  return kf_forward_generic_api<float>(obs, m0, S0, C, A, Q, R, r_shared, shared_mask, B, T, n, p, first, filt, status, stream);
}

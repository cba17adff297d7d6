// eksb_setup.cu -- K8: per-track Kalman parameter setup on the device.
//
// Replaces, for B tracks at once and without leaving HBM, the reference's per-keypoint host block
// multiview_pca_smoother.py:684-731:
//   max_vars = max_j ev[t,j]; good = max_vars <= np.percentile(max_vars, q)      (:686-688)
//   means_j  = mean over good frames of preds[:,j]                               (:698-700)
//   PCA(3) of the centred good rows -> W = components_ (3,p)                     (:707-708, :29-31)
//   S0 = diag(var(good pcs)) [ddof 0]; Q = s * cov(diff of consecutive good pcs) [ddof 1]  (:722-728)
//   C = W^T                                                                      (:730)
// Everything downstream of the good-frame mask is linear in y, so ONE pass over the obs records
// accumulates the masked first/second moments of y and of dy (difference of consecutive good frames)
// in the observation space; W, S0 = diag(W Cov0 W^T) and Q = s W Cov_d W^T follow per track from a
// p x p Jacobi eigen-decomposition.  The percentile is an exact order statistic: an 8-bit MSB radix
// select over the IEEE bit patterns of max_vars (non-negative, so bit order == value order), then
// numpy's linear interpolation formula, evaluated without FMA contraction so the threshold (and
// with it the mask) is bit-identical to numpy's.
// Accumulation and the eigen-solve are double precision for both storage types.
#include "eksb_common.cuh"

namespace eksb {

constexpr int kSelWarps = 8;      // warps per tile in the select kernel
constexpr int kMomSlices = 16;    // time slices (= warps) per tile in the moments kernel
constexpr int kMomWarpsPerCta = 4;

template <typename T> struct KeyOf;
template <> struct KeyOf<double> {
  using type = unsigned long long;
  static constexpr int kBits = 64;
  static __device__ __forceinline__ type of(double v) { return (type)__double_as_longlong(v); }
  static __device__ __forceinline__ double back(type k) { return __longlong_as_double((long long)k); }
};
template <> struct KeyOf<float> {
  using type = unsigned int;
  static constexpr int kBits = 32;
  static __device__ __forceinline__ type of(float v) { return __float_as_uint(v); }
  static __device__ __forceinline__ float back(type k) { return __uint_as_float(k); }
};

// max over the P ensemble-variance rows of one frame (rows [P,2P) of the obs record); -0.0 -> +0.0
template <typename T, int P> __device__ __forceinline__ T frame_max_var(const T* rec_lane) {
  T m = rec_lane[P * kLanes];
#pragma unroll
  for (int j = 1; j < P; ++j) {
    const T v = rec_lane[(P + j) * kLanes];
    m = v > m ? v : m;
  }
  return m + T(0);
}

// numpy.percentile(method='linear') position and interpolation (numpy/lib/_function_base_impl.py
// _quantile / _lerp), in plain IEEE double operations.
__host__ __device__ inline void percentile_position(long T, double q, long& lo, double& gamma) {
  const double quant = q / 100.0;
  double pos = (double)(T - 1) * quant;
  if (pos < 0) pos = 0;
  if (pos > (double)(T - 1)) pos = (double)(T - 1);
  lo = (long)pos;                 // floor (pos >= 0)
  gamma = pos - (double)lo;
}

// ---- 1. per-track percentile threshold ------------------------------------------------------------
template <typename T, int P>
__global__ void __launch_bounds__(kSelWarps* kLanes)
setup_select_kernel(const T* __restrict__ obs, long B, long Tn, double q, double* __restrict__ thr) {
  using K = KeyOf<T>;
  using Key = typename K::type;
  __shared__ unsigned int hist[256][kLanes];
  __shared__ Key s_prefix[kLanes];
  __shared__ long s_rank[kLanes];
  __shared__ unsigned int s_cnt_le[kLanes];
  __shared__ Key s_next[kSelWarps][kLanes];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long g = blockIdx.x;
  const T* base = obs + (g * Tn * (2 * P)) * kLanes + lane;
  long lo;
  double gamma;
  percentile_position(Tn, q, lo, gamma);
  if (warp == 0) { s_prefix[lane] = 0; s_rank[lane] = lo; }
  for (int shift = K::kBits - 8; shift >= 0; shift -= 8) {
    for (int i = threadIdx.x; i < 256 * kLanes; i += blockDim.x) (&hist[0][0])[i] = 0;
    __syncthreads();
    const Key prefix = s_prefix[lane];
    const bool top = (shift == K::kBits - 8);
    for (long t = warp; t < Tn; t += kSelWarps) {
      const Key key = K::of(frame_max_var<T, P>(base + t * (2 * P) * kLanes));
      if (top || (key >> (shift + 8)) == (prefix >> (shift + 8)))
        atomicAdd(&hist[(unsigned)(key >> shift) & 255u][lane], 1u);
    }
    __syncthreads();
    if (warp == 0) {
      long k = s_rank[lane];
      int d = 0;
      for (; d < 255; ++d) {
        const unsigned c = hist[d][lane];
        if (k < (long)c) break;
        k -= c;
      }
      s_rank[lane] = k;
      s_prefix[lane] = prefix | ((Key)d << shift);
    }
    __syncthreads();
  }
  // v_lo = sorted[lo] is now known exactly; sorted[lo+1] = v_lo if enough copies, else the next larger value
  const Key klo = s_prefix[lane];
  unsigned cnt_le = 0;
  Key nxt = ~(Key)0;
  for (long t = warp; t < Tn; t += kSelWarps) {
    const Key key = K::of(frame_max_var<T, P>(base + t * (2 * P) * kLanes));
    if (key <= klo) ++cnt_le;
    else if (key < nxt) nxt = key;
  }
  if (warp == 0) s_cnt_le[lane] = 0;
  __syncthreads();
  atomicAdd(&s_cnt_le[lane], cnt_le);
  s_next[warp][lane] = nxt;
  __syncthreads();
  if (warp == 0) {
    for (int w = 1; w < kSelWarps; ++w) nxt = s_next[w][lane] < nxt ? s_next[w][lane] : nxt;
    const double a = (double)K::back(klo);
    const double b2 = ((long)s_cnt_le[lane] >= lo + 2 || lo + 1 >= Tn) ? a : (double)K::back(nxt);
    const double diff = __dsub_rn(b2, a);
    double v = __dadd_rn(a, __dmul_rn(diff, gamma));
    if (gamma >= 0.5) v = __dsub_rn(b2, __dmul_rn(diff, __dsub_rn(1.0, gamma)));
    const long b = g * kLanes + lane;
    if (b < B) thr[b] = v;
  }
}

// ---- 2. masked moments of y and of dy over the good frames ----------------------------------------
// One warp per (tile, time slice).  Partial record rows (doubles, [tile][slice][row][32 lanes]):
//   0: n_good | 1..P: sum(y-s) | sum (y-s)(y-s)^T packed | n_d | sum dy | sum dy dy^T packed |
//   first good y | last good y          with s = y[frame 0] (shift against cancellation)
template <int P> struct MomRows {
  static constexpr int PS = P * (P + 1) / 2;
  static constexpr int N0 = 0, SY = 1, SYY = SY + P, ND = SYY + PS, SD = ND + 1, SDD = SD + P, FIRST = SDD + PS,
                       LAST = FIRST + P, ROWS = LAST + P;
};

template <typename T, int P>
__global__ void __launch_bounds__(kMomWarpsPerCta* kLanes)
setup_moments_kernel(const T* __restrict__ obs, long B, long Tn, const double* __restrict__ thr,
                     double* __restrict__ part) {
  using MR = MomRows<P>;
  constexpr int PS = MR::PS;
  const int lane = threadIdx.x & 31;
  const long w = (long)blockIdx.x * kMomWarpsPerCta + (threadIdx.x >> 5);
  const long g = w / kMomSlices;
  const int sl = (int)(w - g * kMomSlices);
  if (g >= num_tiles_dev(B)) return;
  const long b = g * kLanes + lane;
  const double th = thr[b < B ? b : B - 1];
  const T* base = obs + (g * Tn * (2 * P)) * kLanes + lane;
  const long len = (Tn + kMomSlices - 1) / kMomSlices;
  const long t0 = sl * len, t1 = (t0 + len < Tn) ? t0 + len : Tn;
  double sh[P], sy[P], syy[PS], sd[P], sdd[PS], first[P], last[P];
  double n0 = 0, nd = 0;
#pragma unroll
  for (int j = 0; j < P; ++j) { sh[j] = (double)base[j * kLanes]; sy[j] = sd[j] = first[j] = last[j] = 0; }
#pragma unroll
  for (int j = 0; j < PS; ++j) syy[j] = sdd[j] = 0;
  bool have = false;
  for (long t = t0; t < t1; ++t) {
    const T* r = base + t * (2 * P) * kLanes;
    if ((double)frame_max_var<T, P>(r) <= th) {
      double y[P];
#pragma unroll
      for (int j = 0; j < P; ++j) y[j] = (double)r[j * kLanes];
      if (have) {
        double d[P];
#pragma unroll
        for (int j = 0; j < P; ++j) { d[j] = y[j] - last[j]; sd[j] += d[j]; }
        int k = 0;
#pragma unroll
        for (int i = 0; i < P; ++i)
#pragma unroll
          for (int j = i; j < P; ++j) sdd[k++] += d[i] * d[j];
        nd += 1;
      } else {
#pragma unroll
        for (int j = 0; j < P; ++j) first[j] = y[j];
        have = true;
      }
      double c[P];
#pragma unroll
      for (int j = 0; j < P; ++j) { last[j] = y[j]; c[j] = y[j] - sh[j]; sy[j] += c[j]; }
      int k = 0;
#pragma unroll
      for (int i = 0; i < P; ++i)
#pragma unroll
        for (int j = i; j < P; ++j) syy[k++] += c[i] * c[j];
      n0 += 1;
    }
  }
  double* o = part + ((g * kMomSlices + sl) * MR::ROWS) * kLanes + lane;
  o[MR::N0 * kLanes] = n0;
  o[MR::ND * kLanes] = nd;
#pragma unroll
  for (int j = 0; j < P; ++j) {
    o[(MR::SY + j) * kLanes] = sy[j];
    o[(MR::SD + j) * kLanes] = sd[j];
    o[(MR::FIRST + j) * kLanes] = first[j];
    o[(MR::LAST + j) * kLanes] = last[j];
  }
#pragma unroll
  for (int j = 0; j < PS; ++j) { o[(MR::SYY + j) * kLanes] = syy[j]; o[(MR::SDD + j) * kLanes] = sdd[j]; }
}

// ---- 3. per-track finalize: merge slices, eigen-decomposition, (means, C, S0, Q) ---------------------
// cyclic Jacobi (Rutishauser's update form) on a symmetric P x P matrix held in a flat array; on return
// d[] holds the eigenvalues and column c of v (v[i*P+c]) the c-th eigenvector.  Only the upper triangle of a
// is referenced.
template <int P> __device__ __noinline__ void jacobi_eig(double* a, double* d, double* v) {
  double bb[P], z[P];
  for (int i = 0; i < P; ++i) {
    for (int j = 0; j < P; ++j) v[i * P + j] = (i == j) ? 1.0 : 0.0;
    bb[i] = d[i] = a[i * P + i];
    z[i] = 0.0;
  }
  for (int sweep = 0; sweep < 50; ++sweep) {
    double off = 0.0;
    for (int i = 0; i < P - 1; ++i)
      for (int j = i + 1; j < P; ++j) off += fabs(a[i * P + j]);
    if (off == 0.0) break;
    for (int ip = 0; ip < P - 1; ++ip)
      for (int iq = ip + 1; iq < P; ++iq) {
        const double apq = a[ip * P + iq];
        const double g = 100.0 * fabs(apq);
        if (sweep > 3 && fabs(d[ip]) + g == fabs(d[ip]) && fabs(d[iq]) + g == fabs(d[iq])) {
          a[ip * P + iq] = 0.0;
        } else if (apq != 0.0) {
          const double h = d[iq] - d[ip];
          double t;
          if (fabs(h) + g == fabs(h)) {
            t = apq / h;
          } else {
            const double theta = 0.5 * h / apq;
            t = 1.0 / (fabs(theta) + sqrt(1.0 + theta * theta));
            if (theta < 0.0) t = -t;
          }
          const double c = 1.0 / sqrt(1.0 + t * t), sn = t * c, tau = sn / (1.0 + c), hh = t * apq;
          z[ip] -= hh; z[iq] += hh; d[ip] -= hh; d[iq] += hh;
          a[ip * P + iq] = 0.0;
          for (int j = 0; j < ip; ++j) {                       // rotate the upper triangle
            const double x = a[j * P + ip], y = a[j * P + iq];
            a[j * P + ip] = x - sn * (y + x * tau);
            a[j * P + iq] = y + sn * (x - y * tau);
          }
          for (int j = ip + 1; j < iq; ++j) {
            const double x = a[ip * P + j], y = a[j * P + iq];
            a[ip * P + j] = x - sn * (y + x * tau);
            a[j * P + iq] = y + sn * (x - y * tau);
          }
          for (int j = iq + 1; j < P; ++j) {
            const double x = a[ip * P + j], y = a[iq * P + j];
            a[ip * P + j] = x - sn * (y + x * tau);
            a[iq * P + j] = y + sn * (x - y * tau);
          }
          for (int j = 0; j < P; ++j) {
            const double x = v[j * P + ip], y = v[j * P + iq];
            v[j * P + ip] = x - sn * (y + x * tau);
            v[j * P + iq] = y + sn * (x - y * tau);
          }
        }
      }
    for (int i = 0; i < P; ++i) { bb[i] += z[i]; d[i] = bb[i]; z[i] = 0.0; }
  }
}

template <typename T, int P, int N>
__global__ void setup_finalize_kernel(const T* __restrict__ obs, const double* __restrict__ part, long B, long Tn,
                                      double smooth_param, T* __restrict__ means, T* __restrict__ Cout,
                                      T* __restrict__ S0, T* __restrict__ Q, int* __restrict__ status) {
  using MR = MomRows<P>;
  constexpr int PS = MR::PS;
  const long b = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const long g = b / kLanes;
  const int lane = (int)(b - g * kLanes);
  double n0 = 0, nd = 0, sy[P], syy[PS], sd[P], sdd[PS], last[P];
  for (int j = 0; j < P; ++j) sy[j] = sd[j] = last[j] = 0;
  for (int j = 0; j < PS; ++j) syy[j] = sdd[j] = 0;
  bool have = false;
  for (int sl = 0; sl < kMomSlices; ++sl) {
    const double* r = part + ((g * kMomSlices + sl) * MR::ROWS) * kLanes + lane;
    const double n_s = r[MR::N0 * kLanes];
    if (n_s == 0) continue;
    if (have) {                                 // stitch: first good frame of this slice follows `last`
      double d[P];
      for (int j = 0; j < P; ++j) { d[j] = r[(MR::FIRST + j) * kLanes] - last[j]; sd[j] += d[j]; }
      int k = 0;
      for (int i = 0; i < P; ++i)
        for (int j = i; j < P; ++j) sdd[k++] += d[i] * d[j];
      nd += 1;
    }
    have = true;
    n0 += n_s;
    nd += r[MR::ND * kLanes];
    for (int j = 0; j < P; ++j) {
      sy[j] += r[(MR::SY + j) * kLanes];
      sd[j] += r[(MR::SD + j) * kLanes];
      last[j] = r[(MR::LAST + j) * kLanes];
    }
    for (int j = 0; j < PS; ++j) { syy[j] += r[(MR::SYY + j) * kLanes]; sdd[j] += r[(MR::SDD + j) * kLanes]; }
  }
  bool ok = n0 >= 2 && nd >= 2;
  const T* base = obs + (g * Tn * (2 * P)) * kLanes + lane;
  double cov0[P][P], covd[P][P], a[P * P], v[P * P], ev[P];
  {
    int k = 0;
    for (int i = 0; i < P; ++i)
      for (int j = i; j < P; ++j, ++k) {
        const double c0 = (syy[k] - sy[i] * sy[j] / n0) / n0;              // ddof 0 (np.var, :722)
        const double cd = (sdd[k] - sd[i] * sd[j] / nd) / (nd - 1.0);      // ddof 1 (np.cov, :728)
        cov0[i][j] = cov0[j][i] = c0;
        covd[i][j] = covd[j][i] = cd;
        a[i * P + j] = a[j * P + i] = c0;
      }
  }
  for (int j = 0; j < P; ++j) means[b * P + j] = (T)((double)base[j * kLanes] + sy[j] / n0);   // (:698-700)
  jacobi_eig<P>(a, ev, v);
  int order[P];
  for (int i = 0; i < P; ++i) order[i] = i;
  for (int i = 0; i < P - 1; ++i)              // eigenvalues descending (sklearn PCA order)
    for (int j = i + 1; j < P; ++j)
      if (ev[order[j]] > ev[order[i]]) { const int tmp = order[i]; order[i] = order[j]; order[j] = tmp; }
  double W[N][P];
  for (int c = 0; c < N; ++c) {
    const int col = order[c];
    int jm = 0;                                 // sklearn svd_flip(u_based_decision=False): largest |entry| positive
    for (int j = 1; j < P; ++j)
      if (fabs(v[j * P + col]) > fabs(v[jm * P + col])) jm = j;
    const double sgn = v[jm * P + col] < 0 ? -1.0 : 1.0;
    for (int j = 0; j < P; ++j) W[c][j] = sgn * v[j * P + col];
  }
  for (int j = 0; j < P; ++j)
    for (int c = 0; c < N; ++c) Cout[(b * P + j) * N + c] = (T)W[c][j];     // C = W^T (:730)
  for (int c = 0; c < N; ++c)
    for (int e = 0; e < N; ++e) {
      double s0 = 0, q = 0;
      for (int i = 0; i < P; ++i)
        for (int j = 0; j < P; ++j) {
          s0 += W[c][i] * cov0[i][j] * W[e][j];
          q += W[c][i] * covd[i][j] * W[e][j];
        }
      S0[(b * N + c) * N + e] = (T)((c == e) ? s0 : 0.0);                  // np.diag(np.var(..)) (:722)
      Q[(b * N + c) * N + e] = (T)(smooth_param * q);
    }
  if (status && !ok) status[b] = 1;
}

// ---- 4. centre the observations in place: y[t,j] -= means[b,j]  (:702-705) ---------------------------
template <typename T>
__global__ void center_rows_kernel(T* __restrict__ rec, const T* __restrict__ means, long B, long Tn, int rows, int R) {
  const long line = (long)blockIdx.x * (blockDim.x / kLanes) + (threadIdx.x >> 5);   // (g, t, r<R)
  const int lane = threadIdx.x & 31;
  const long total = num_tiles_dev(B) * Tn * R;
  if (line >= total) return;
  const int r = (int)(line % R);
  const long gt = line / R;
  const long g = gt / Tn;
  const long b = g * kLanes + lane;
  if (b >= B) return;
  rec[(gt * rows + r) * kLanes + lane] -= means[b * R + r];
}

template <typename T, int P>
static int setup_run(T* obs, long B, long Tn, int n, double q, double s, int center, T* means, T* C, T* S0, T* Q,
                     double* ws, int* status, cudaStream_t st) {
  using MR = MomRows<P>;
  const long tiles = num_tiles(B);
  double* thr = ws;
  double* part = ws + tiles * kLanes;
  setup_select_kernel<T, P><<<(unsigned)tiles, kSelWarps * kLanes, 0, st>>>(obs, B, Tn, q, thr);
  const long warps = tiles * kMomSlices;
  setup_moments_kernel<T, P><<<(unsigned)((warps + kMomWarpsPerCta - 1) / kMomWarpsPerCta), kMomWarpsPerCta * kLanes, 0, st>>>(
      obs, B, Tn, thr, part);
  if (n != 3) { set_error("pca_setup: n=%d (the reference fits 3 components)", n); return EKSB_EDIM; }
  setup_finalize_kernel<T, P, 3><<<(unsigned)((B + 63) / 64), 64, 0, st>>>(obs, part, B, Tn, s, means, C, S0, Q, status);
  if (center) {
    const long lines = tiles * Tn * P;
    center_rows_kernel<T><<<(unsigned)((lines + 7) / 8), 8 * kLanes, 0, st>>>(obs, means, B, Tn, 2 * P, P);
  }
  (void)sizeof(MR);
  return check_launch("pca_setup kernels");
}

template <typename T>
static int setup_api(T* obs, long B, long Tn, int n, int p, double q, double s, int center, T* means, T* C, T* S0, T* Q,
                     double* ws, size_t ws_bytes, int* status, void* stream) {
  if (!obs || !means || !C || !S0 || !Q || !ws || B <= 0 || Tn <= 1) { set_error("pca_setup: null pointer or empty batch"); return EKSB_EARG; }
  if (!(q >= 0.0 && q <= 100.0)) { set_error("pca_setup: quantile %g outside [0,100]", q); return EKSB_EARG; }
  if (ws_bytes < eksb_pca_setup_workspace_bytes(B, p)) { set_error("pca_setup: workspace too small"); return EKSB_EARG; }
  cudaStream_t st = (cudaStream_t)stream;
  switch (p) {
    case 4: return setup_run<T, 4>(obs, B, Tn, n, q, s, center, means, C, S0, Q, ws, status, st);
    case 6: return setup_run<T, 6>(obs, B, Tn, n, q, s, center, means, C, S0, Q, ws, status, st);
    case 8: return setup_run<T, 8>(obs, B, Tn, n, q, s, center, means, C, S0, Q, ws, status, st);
  }
  set_error("pca_setup: p=%d not instantiated (4, 6, 8)", p);
  return EKSB_EDIM;
}

}  // namespace eksb

using namespace eksb;

extern "C" size_t eksb_pca_setup_workspace_bytes(long B, int p) {
  if (B <= 0 || p <= 0) return 0;
  const size_t rows = 3 + 4 * (size_t)p + (size_t)p * (p + 1);
  return (size_t)num_tiles(B) * kLanes * sizeof(double) * (1 + kMomSlices * rows);
}
extern "C" int eksb_pca_setup_f64(double* obs, long B, long T, int n, int p, double quantile, double smooth_param,
                                  int center, double* means, double* C, double* S0, double* Q, double* workspace,
                                  size_t workspace_bytes, int* status, void* stream) {
  return setup_api<double>(obs, B, T, n, p, quantile, smooth_param, center, means, C, S0, Q, workspace, workspace_bytes, status, stream);
}
extern "C" int eksb_pca_setup_f32(float* obs, long B, long T, int n, int p, double quantile, double smooth_param,
                                  int center, float* means, float* C, float* S0, float* Q, double* workspace,
                                  size_t workspace_bytes, int* status, void* stream) {
  return setup_api<float>(obs, B, T, n, p, quantile, smooth_param, center, means, C, S0, Q, workspace, workspace_bytes, status, stream);
}

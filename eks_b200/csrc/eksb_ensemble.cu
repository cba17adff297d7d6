// eksb_ensemble.cu -- A1: ensemble median/mean + variance of the sample mean
// (eks/ensemble_kalman.py:4-57), optionally fused with the multi-view restack into
// track-tile obs records (multiview_pca_smoother.py:645-682).
//
// HBM-bound streaming reduction: each thread owns one (frame, column) element, reads its M
// model values with fully coalesced loads (consecutive threads = consecutive columns of the
// [M][T][K] tensor), selects the median with an in-register compare-exchange network and
// accumulates mean / variance in numpy's order with contraction-free arithmetic so medians,
// means AND variances are bit-identical to numpy on the same inputs.
#include "eksb_common.cuh"

namespace eksb {

// Contraction-free IEEE ops (nvcc would otherwise fuse a*b+c into FMA and change the last bit).
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double div_rn(double a, double b) { return __ddiv_rn(a, b); }
__device__ __forceinline__ float div_rn(float a, float b) { return __fdiv_rn(a, b); }

template <typename T> __device__ __forceinline__ T quiet_nan();
template <> __device__ __forceinline__ double quiet_nan<double>() { return __longlong_as_double(0x7ff8000000000000LL); }
template <> __device__ __forceinline__ float quiet_nan<float>() { return __int_as_float(0x7fc00000); }

// avg / var of M values held in registers.
//   median: odd M -> middle order statistic; even M -> (lo+hi)/2; any NaN -> NaN (numpy.median)
//   mean  : ((v0+v1)+v2)+... / M                                   (numpy.mean over a strided axis)
//   var   : sum_m (v_m - mean)^2 / M, then / M again               (eks/ensemble_kalman.py:46)
template <typename T, int M>
__device__ __forceinline__ void reduce_models(const T (&v)[M], int mode, T& avg, T& var) {
  T sum = v[0];
#pragma unroll
  for (int m = 1; m < M; ++m) sum = add_rn(sum, v[m]);
  const T mean = div_rn(sum, T(M));
  T ss = T(0);
#pragma unroll
  for (int m = 0; m < M; ++m) {
    const T d = add_rn(v[m], -mean);
    ss = add_rn(ss, mul_rn(d, d));
  }
  var = div_rn(div_rn(ss, T(M)), T(M));
  if (mode == EKSB_MODE_MEAN) { avg = mean; return; }
  T s[M];
  bool has_nan = false;
#pragma unroll
  for (int m = 0; m < M; ++m) { s[m] = v[m]; has_nan = has_nan || (v[m] != v[m]); }
  // odd-even transposition sort: M rounds of independent compare-exchanges, all in registers
#pragma unroll
  for (int r = 0; r < M; ++r) {
#pragma unroll
    for (int i = (r & 1); i + 1 < M; i += 2) {
      const T a = s[i], b = s[i + 1];
      const bool sw = b < a;
      s[i] = sw ? b : a;
      s[i + 1] = sw ? a : b;
    }
  }
  T med = (M & 1) ? s[M / 2] : mul_rn(add_rn(s[(M - 1) / 2], s[M / 2]), T(0.5));
  avg = has_nan ? quiet_nan<T>() : med;
}

// Plain layout: avg/var (T,K).  One thread per element of the flattened (T*K) axis.
template <typename T, int M>
__global__ void __launch_bounds__(256) ensemble_kernel(const T* __restrict__ preds, long TK, int mode,
                                                       T* __restrict__ avg, T* __restrict__ var) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < TK; i += stride) {
    T v[M];
#pragma unroll
    for (int m = 0; m < M; ++m) v[m] = preds[(long)m * TK + i];
    T a, s;
    reduce_models<T, M>(v, mode, a, s);
    avg[i] = a;
    var[i] = s;
  }
}

// Any ensemble size up to kMaxModels (runtime M; values in local memory, insertion sort): the reference takes any
// number of models (eks/ensemble_kalman.py:41-46).  Same summation order and rounding as reduce_models.
constexpr int kMaxModels = 64;
template <typename T>
__global__ void __launch_bounds__(256) ensemble_anym_kernel(const T* __restrict__ preds, int M, long TK, int mode,
                                                            T* __restrict__ avg, T* __restrict__ var) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < TK; i += stride) {
    T v[kMaxModels];
    bool has_nan = false;
    for (int m = 0; m < M; ++m) { v[m] = preds[(long)m * TK + i]; has_nan = has_nan || (v[m] != v[m]); }
    T sum = v[0];
    for (int m = 1; m < M; ++m) sum = add_rn(sum, v[m]);
    const T mean = div_rn(sum, T(M));
    T ss = T(0);
    for (int m = 0; m < M; ++m) { const T d = add_rn(v[m], -mean); ss = add_rn(ss, mul_rn(d, d)); }
    var[i] = div_rn(div_rn(ss, T(M)), T(M));
    if (mode == EKSB_MODE_MEAN) { avg[i] = mean; continue; }
    for (int a = 1; a < M; ++a) {
      const T x = v[a];
      int c = a - 1;
      while (c >= 0 && x < v[c]) { v[c + 1] = v[c]; --c; }
      v[c + 1] = x;
    }
    const T med = (M & 1) ? v[M / 2] : mul_rn(add_rn(v[(M - 1) / 2], v[M / 2]), T(0.5));
    avg[i] = has_nan ? quiet_nan<T>() : med;
  }
}

// Fused restack: column k = b*P + j.  A CTA of 256 threads handles one frame-chunk of one
// track tile: thread (w, lane) walks frames w, w+8, ... and all P observation rows of track
// 32 g + lane... reads of one frame row are contiguous over (lane, j) = 32*P scalars, and each
// written row is a full 32-lane line.
template <typename T, int M, int P>
__global__ void __launch_bounds__(256) ensemble_tiled_kernel(const T* __restrict__ preds, long Tn, long B, int mode,
                                                             const T* __restrict__ offset, T* __restrict__ obs,
                                                             T* __restrict__ mxout /* [tile][t][32] or null */) {
  constexpr int kFramesPerCta = 64;
  const long nchunk = (Tn + kFramesPerCta - 1) / kFramesPerCta;
  const long g = blockIdx.x / nchunk;
  const long t0 = (blockIdx.x - g * nchunk) * kFramesPerCta;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long b = g * kLanes + lane;
  const long K = B * P, TK = Tn * K;
  T off[P];
#pragma unroll
  for (int j = 0; j < P; ++j) off[j] = (offset && b < B) ? offset[b * P + j] : T(0);
  for (int f = warp; f < kFramesPerCta; f += 8) {
    const long t = t0 + f;
    if (t >= Tn) break;
    T* rec = obs + ((g * Tn + t) * (2 * P)) * kLanes + lane;
    T mxv = T(1);                              // max over the P ensemble variances of this frame (padding lanes: ev = 1)
    if (b < B) {
      T v[P][M];
#pragma unroll
      for (int m = 0; m < M; ++m) {
        const T* row = preds + (long)m * TK + t * K + b * P;
#pragma unroll
        for (int j = 0; j < P; ++j) v[j][m] = row[j];
      }
#pragma unroll
      for (int j = 0; j < P; ++j) {
        T a, s;
        reduce_models<T, M>(v[j], mode, a, s);
        rec[j * kLanes] = a - off[j];
        rec[(P + j) * kLanes] = s;
        mxv = (j == 0) ? s : (s > mxv ? s : mxv);
      }
    } else {
#pragma unroll
      for (int j = 0; j < P; ++j) { rec[j * kLanes] = T(0); rec[(P + j) * kLanes] = T(1); }
    }
    // the parameter setup's first pass (eksb_setup.cu setup_mx_kernel: max_vars, -0.0 -> +0.0) done while the
    // variances are in registers: 1 row written instead of P rows read back
    if (mxout) mxout[(g * Tn + t) * kLanes + lane] = mxv + T(0);
  }
}

template <typename T, int M>
static int launch_ensemble(const T* preds, long TK, int mode, T* avg, T* var, cudaStream_t st) {
  long blocks = (TK + 255) / 256;
  const long cap = (long)sm_count() * 64;   // a few waves over the SMs; grid-stride covers the rest
  if (blocks > cap) blocks = cap;
  ensemble_kernel<T, M><<<(unsigned)blocks, 256, 0, st>>>(preds, TK, mode, avg, var);
  return check_launch("ensemble_kernel");
}

template <typename T>
static int ensemble_api(const T* preds, int M, long Tn, long K, int mode, T* avg, T* var, void* stream) {
  if (!preds || !avg || !var || Tn <= 0 || K <= 0) { set_error("ensemble: bad argument"); return EKSB_EARG; }
  if (mode != EKSB_MODE_MEDIAN && mode != EKSB_MODE_MEAN) { set_error("ensemble: bad mode %d", mode); return EKSB_EARG; }
  cudaStream_t st = (cudaStream_t)stream;
  switch (M) {
#define CASE(M_) case M_: return launch_ensemble<T, M_>(preds, Tn * K, mode, avg, var, st);
    CASE(1) CASE(2) CASE(3) CASE(4) CASE(5) CASE(6) CASE(7) CASE(8) CASE(9) CASE(10) CASE(11) CASE(12)
    CASE(13) CASE(14) CASE(15) CASE(16)
#undef CASE
    default: break;
  }
  if (M < 1 || M > kMaxModels) { set_error("ensemble: M=%d outside 1..%d", M, kMaxModels); return EKSB_EDIM; }
  long blocks = (Tn * K + 255) / 256;
  if (blocks > (long)sm_count() * 64) blocks = (long)sm_count() * 64;
  ensemble_anym_kernel<T><<<(unsigned)blocks, 256, 0, st>>>(preds, M, Tn * K, mode, avg, var);
  return check_launch("ensemble_anym_kernel");
}

template <typename T, int M>
static int launch_ensemble_tiled(const T* preds, long Tn, long B, int p, int mode, const T* offset, T* obs, T* mx, cudaStream_t st) {
  const unsigned grid = (unsigned)(((Tn + 63) / 64) * num_tiles(B));
  switch (p) {
    case 4: ensemble_tiled_kernel<T, M, 4><<<grid, 256, 0, st>>>(preds, Tn, B, mode, offset, obs, mx); break;
    case 6: ensemble_tiled_kernel<T, M, 6><<<grid, 256, 0, st>>>(preds, Tn, B, mode, offset, obs, mx); break;
    case 8: ensemble_tiled_kernel<T, M, 8><<<grid, 256, 0, st>>>(preds, Tn, B, mode, offset, obs, mx); break;
    default: set_error("ensemble_tiled: p=%d not in {4,6,8}", p); return EKSB_EDIM;
  }
  return check_launch("ensemble_tiled_kernel");
}

template <typename T>
static int ensemble_tiled_api(const T* preds, int M, long Tn, long B, int p, int mode, const T* offset, T* obs, void* stream,
                              T* mx = nullptr) {
  if (!preds || !obs || Tn <= 0 || B <= 0) { set_error("ensemble_tiled: bad argument"); return EKSB_EARG; }
  if (mode != EKSB_MODE_MEDIAN && mode != EKSB_MODE_MEAN) { set_error("ensemble_tiled: bad mode %d", mode); return EKSB_EARG; }
  cudaStream_t st = (cudaStream_t)stream;
  switch (M) {
#define CASE(M_) case M_: return launch_ensemble_tiled<T, M_>(preds, Tn, B, p, mode, offset, obs, mx, st);
    CASE(1) CASE(2) CASE(3) CASE(4) CASE(5) CASE(6) CASE(7) CASE(8)
#undef CASE
    default: set_error("ensemble_tiled: M=%d outside 1..8", M); return EKSB_EDIM;
  }
}

}  // namespace eksb

using namespace eksb;

extern "C" int eksb_ensemble_f64(const double* preds, int M, long T, long K, int mode, double* avg, double* var, void* stream) {
  return ensemble_api<double>(preds, M, T, K, mode, avg, var, stream);
}
extern "C" int eksb_ensemble_f32(const float* preds, int M, long T, long K, int mode, float* avg, float* var, void* stream) {
  return ensemble_api<float>(preds, M, T, K, mode, avg, var, stream);
}
extern "C" int eksb_ensemble_tiled_f64(const double* preds, int M, long T, long B, int p, int mode, const double* offset,
                                       double* obs, void* stream) {
  return ensemble_tiled_api<double>(preds, M, T, B, p, mode, offset, obs, stream);
}
extern "C" int eksb_ensemble_tiled_f32(const float* preds, int M, long T, long B, int p, int mode, const float* offset,
                                       float* obs, void* stream) {
  return ensemble_tiled_api<float>(preds, M, T, B, p, mode, offset, obs, stream);
}

extern "C" int eksb_ensemble_tiled_mx_f64(const double* preds, int M, long T, long B, int p, int mode, const double* offset,
                                          double* obs, double* max_var, void* stream) {
  return ensemble_tiled_api<double>(preds, M, T, B, p, mode, offset, obs, stream, max_var);
}
extern "C" int eksb_ensemble_tiled_mx_f32(const float* preds, int M, long T, long B, int p, int mode, const float* offset,
                                          float* obs, float* max_var, void* stream) {
  return ensemble_tiled_api<float>(preds, M, T, B, p, mode, offset, obs, stream, max_var);
}

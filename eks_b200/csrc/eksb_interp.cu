// eksb_interp.cu -- asynchronous-view alignment: batched linear interpolation of
// one camera's prediction series onto another camera's timestamps, with the optional mirror flip.
//
// Reference: multiview_pca_smoother.py:70-101 -- for every left-camera frame whose timestamp lies inside
// the right camera's range, scipy.interpolate.interp1d(kind='linear') over the two bracketing right-camera
// frames, then x -> img_width - x (:90-91).  interp1d's arithmetic is
//     slope = (y_hi - y_lo) / (x_hi - x_lo);   y = slope * (x - x_lo) + y_lo
// evaluated here operation by operation without FMA contraction, so results are bit-identical.
// One thread per (series, query frame); the K columns of a source frame are contiguous.
#include "eksb_common.cuh"

namespace eksb {

template <typename T> struct Rn;
template <> struct Rn<double> {
  static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
  static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
  static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
  static __device__ __forceinline__ double div(double a, double b) { return __ddiv_rn(a, b); }
};
template <> struct Rn<float> {
  static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
  static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
  static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
  static __device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
};

template <typename T>
__global__ void interp_rows_kernel(const T* __restrict__ ts_src, const T* __restrict__ src, long S, long Ts, int K,
                                   const T* __restrict__ ts_q, long Tq, const T* __restrict__ flip,
                                   T* __restrict__ out) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= S * Tq) return;
  const long s = i / Tq, tq = i - s * Tq;
  const T x = ts_q[tq];
  // hi = first index with ts_src[hi] >= x (numpy.searchsorted side='left'), clipped to [1, Ts-1]
  long lo_i = 0, hi_i = Ts;
  while (lo_i < hi_i) {
    const long mid = (lo_i + hi_i) >> 1;
    if (ts_src[mid] < x) lo_i = mid + 1; else hi_i = mid;
  }
  long hi = lo_i < 1 ? 1 : (lo_i > Ts - 1 ? Ts - 1 : lo_i);
  const long lo = hi - 1;
  const T x_lo = ts_src[lo], x_hi = ts_src[hi];
  const T dx = Rn<T>::sub(x_hi, x_lo), off = Rn<T>::sub(x, x_lo);
  const T* r_lo = src + (s * Ts + lo) * K;
  const T* r_hi = src + (s * Ts + hi) * K;
  T* o = out + (s * Tq + tq) * K;
  for (int j = 0; j < K; ++j) {
    const T slope = Rn<T>::div(Rn<T>::sub(r_hi[j], r_lo[j]), dx);
    T y = Rn<T>::add(Rn<T>::mul(slope, off), r_lo[j]);
    if (flip && flip[j] != T(0)) y = Rn<T>::sub(flip[j], y);
    o[j] = y;
  }
}

template <typename T>
static int interp_api(const T* ts_src, const T* src, long S, long Ts, int K, const T* ts_q, long Tq, const T* flip,
                      T* out, void* stream) {
  if (!ts_src || !src || !ts_q || !out || S <= 0 || Ts < 2 || K <= 0 || Tq <= 0) {
    set_error("interp_rows: null pointer, empty batch or fewer than 2 source frames");
    return EKSB_EARG;
  }
  const long n = S * Tq;
  interp_rows_kernel<T><<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(ts_src, src, S, Ts, K, ts_q, Tq, flip, out);
  return check_launch("interp_rows_kernel");
}

}  // namespace eksb

using namespace eksb;

extern "C" int eksb_interp_rows_f64(const double* ts_src, const double* src, long S, long Ts, int K, const double* ts_q,
                                    long Tq, const double* flip, double* out, void* stream) {
  return interp_api<double>(ts_src, src, S, Ts, K, ts_q, Tq, flip, out, stream);
}
extern "C" int eksb_interp_rows_f32(const float* ts_src, const float* src, long S, long Ts, int K, const float* ts_q,
                                    long Tq, const float* flip, float* out, void* stream) {
  return interp_api<float>(ts_src, src, S, Ts, K, ts_q, Tq, flip, out, stream);
}

// eksb_layout.cu -- library housekeeping and the layout kernels: (B,T,R) <-> track-tile
// transposes and the dense expansion of state records into the reference's return values.
#include <stdarg.h>
#include <stdlib.h>
#include "eksb_common.cuh"

namespace eksb {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int sm_count() {
  static int cached[64];                                   // 0 = not queried yet; benign race: every thread writes the same value
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    cached[dev] = (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0) ? n : 148;
  }
  return cached[dev];
}

bool use_pipe(long B) {
  const char* e = getenv("EKSB_PIPE");                    // read per call: tests switch kernel families in-process
  return !(e && e[0] == '0') && B > 0;
}

long pipe_groups(long tiles, int elem_size) {
  const char* m = getenv("EKSB_PIPE_MAX_TILES");
  const char* g = getenv("EKSB_PIPE_GROUP_TILES");
  // Measured (tools/gpu_groups.sh, 8 000 frames): f64 65 536 tracks 5.78 / 6.30 / 5.90 / 6.27 / 6.41 TB/s algorithmic in
  // groups of 148 / 256 / 293 / 410 / 512 tiles (plain kernels: 4.72); f32 131 072 tracks 5.72 / 6.21 / 5.85 / 5.56 in
  // groups of 293 / 512 / 586 / 820 (plain: 4.15).  512 tiles = 128 CTAs of 4 tiles, in both types.
  (void)elem_size;
  const long max_single = m ? atol(m) : 592;              // up to 4 tiles on every SM: one launch
  long group = g ? atol(g) : 512;
  if (group < 1) group = 1;
  return tiles <= max_single ? 1 : (tiles + group - 1) / group;
}

constexpr int kFT = 16;   // frames per transpose tile

// (B,T,R) row-major -> rows [row0,row0+R) of track-tile records.  One CTA moves a tile of
// 32 tracks x kFT frames through shared memory: global reads are contiguous runs of kFT*R
// scalars per track, global writes are full 32-lane lines.  grid = n_tiles * ceil(T/kFT).
template <typename T>
__global__ void __launch_bounds__(256) pack_rows_kernel(const T* __restrict__ src, long B, long Tn, int R,
                                                        T* __restrict__ dst, int rows, int row0, T pad) {
  extern __shared__ unsigned char smem_raw[];
  T* s = reinterpret_cast<T*>(smem_raw);
  const long nchunk = (Tn + kFT - 1) / kFT;      // 1-D grid: tile-major, frame-chunk minor
  const long g = blockIdx.x / nchunk;
  const long t0 = (blockIdx.x - g * nchunk) * kFT;
  const int nf = (int)((Tn - t0) < kFT ? (Tn - t0) : kFT);
  const int run = nf * R, pitch = kFT * R + 1;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = warp; i < kLanes; i += 8) {
    const long b = g * kLanes + i;
    for (int k = lane; k < run; k += 32) s[i * pitch + k] = (b < B) ? src[(b * Tn + t0) * R + k] : pad;
  }
  __syncthreads();
  for (int fr = warp; fr < nf * R; fr += 8) {
    const int f = fr / R, r = fr - f * R;
    dst[((g * Tn + t0 + f) * rows + row0 + r) * kLanes + lane] = s[lane * pitch + fr];
  }
}

template <typename T>
__global__ void __launch_bounds__(256) unpack_rows_kernel(const T* __restrict__ src, int rows, int row0, long B, long Tn,
                                                          int R, T* __restrict__ dst) {
  extern __shared__ unsigned char smem_raw[];
  T* s = reinterpret_cast<T*>(smem_raw);
  const long nchunk = (Tn + kFT - 1) / kFT;      // 1-D grid: tile-major, frame-chunk minor
  const long g = blockIdx.x / nchunk;
  const long t0 = (blockIdx.x - g * nchunk) * kFT;
  const int nf = (int)((Tn - t0) < kFT ? (Tn - t0) : kFT);
  const int run = nf * R, pitch = kFT * R + 1;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int fr = warp; fr < run; fr += 8) {
    const int f = fr / R, r = fr - f * R;
    s[lane * pitch + fr] = src[((g * Tn + t0 + f) * rows + row0 + r) * kLanes + lane];
  }
  __syncthreads();
  for (int i = warp; i < kLanes; i += 8) {
    const long b = g * kLanes + i;
    if (b < B)
      for (int k = lane; k < run; k += 32) dst[(b * Tn + t0) * R + k] = s[i * pitch + k];
  }
}

template <typename T>
static int pack_rows_api(const T* src, long B, long Tn, int R, T* dst, int rows, int row0, T pad, void* stream, bool unpack) {
  if (!src || !dst || B <= 0 || Tn <= 0 || R <= 0 || row0 < 0 || row0 + R > rows) { set_error("pack/unpack_rows: bad argument"); return EKSB_EARG; }
  if (R > 64) { set_error("pack/unpack_rows: R > 64"); return EKSB_EARG; }
  unsigned grid = (unsigned)(((Tn + kFT - 1) / kFT) * num_tiles(B));
  size_t smem = sizeof(T) * kLanes * (kFT * R + 1);
  if (smem > 48 * 1024) {
    cudaFuncSetAttribute(pack_rows_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(unpack_rows_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  }
  if (unpack) unpack_rows_kernel<T><<<grid, 256, smem, (cudaStream_t)stream>>>(src, rows, row0, B, Tn, R, dst);
  else        pack_rows_kernel<T><<<grid, 256, smem, (cudaStream_t)stream>>>(src, B, Tn, R, dst, rows, row0, pad);
  return check_launch(unpack ? "unpack_rows_kernel" : "pack_rows_kernel");
}

// One thread per (track, frame): state record -> dense mean / cov / predicted cov.
// Output formatting for the numpy-facing API (small T); reads are lane-coalesced.
template <typename T, int N>
__global__ void expand_state_kernel(const T* __restrict__ state, long B, long Tn, const T* __restrict__ A,
                                    const T* __restrict__ Q, const T* __restrict__ S0, int shared_mask,
                                    T* __restrict__ mean, T* __restrict__ cov, T* __restrict__ pred) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA;
  const int lane = threadIdx.x & 31;
  const long nchunk = (Tn + 7) / 8;
  const long g = blockIdx.x / nchunk;
  const long t = (blockIdx.x - g * nchunk) * 8 + (threadIdx.x >> 5);
  const long b = g * kLanes + lane;
  if (t >= Tn || b >= B) return;
  const T* rec = state + ((g * Tn + t) * SG) * kLanes + lane;
  T m[N], V[NS];
#pragma unroll
  for (int i = 0; i < N; ++i) m[i] = rec[i * kLanes];
#pragma unroll
  for (int i = 0; i < NS; ++i) V[i] = rec[(N + i) * kLanes];
  const long o = b * Tn + t;
#pragma unroll
  for (int i = 0; i < N; ++i) mean[o * N + i] = m[i];
#pragma unroll
  for (int i = 0; i < N; ++i)
#pragma unroll
    for (int j = 0; j < N; ++j) cov[(o * N + i) * N + j] = V[sym_idx<N>(i, j)];
  if (pred) {
    const T* a = A + ((shared_mask & EKSB_SHARED_A) ? 0 : b * (N * N));
    const T* q = Q + ((shared_mask & EKSB_SHARED_Q) ? 0 : b * (N * N));
    const T* s0 = S0 + ((shared_mask & EKSB_SHARED_S0) ? 0 : b * (N * N));
#pragma unroll
    for (int i = 0; i < N; ++i)
#pragma unroll
      for (int j = 0; j < N; ++j) {
        T v = T(0);
        if (Tn == 1) {
          v = s0[i * N + j];                       // eks/ensemble_kalman.py:96 with T == 1
        } else if (t < Tn - 1) {                   // :101 ; S[T-1] stays zero
          v = q[i * N + j];
          for (int k = 0; k < N; ++k)
            for (int l = 0; l < N; ++l) v += a[i * N + k] * V[sym_idx<N>(k, l)] * a[j * N + l];
        }
        pred[(o * N + i) * N + j] = v;
      }
  }
}

template <typename T, int N>
__global__ void pack_state_kernel(const T* __restrict__ mean, const T* __restrict__ cov, long B, long Tn,
                                  T* __restrict__ state) {
  constexpr int NS = Dim<N>::NS, SG = Dim<N>::SIGMA;
  const int lane = threadIdx.x & 31;
  const long nchunk = (Tn + 7) / 8;
  const long g = blockIdx.x / nchunk;
  const long t = (blockIdx.x - g * nchunk) * 8 + (threadIdx.x >> 5);
  const long b = g * kLanes + lane;
  if (t >= Tn) return;
  T* rec = state + ((g * Tn + t) * SG) * kLanes + lane;
  const long o = (b < B ? b : B - 1) * Tn + t;   // padding lanes replicate the last track
#pragma unroll
  for (int i = 0; i < N; ++i) rec[i * kLanes] = mean[o * N + i];
#pragma unroll
  for (int i = 0; i < N; ++i)
#pragma unroll
    for (int j = i; j < N; ++j)
      rec[(N + sym_idx<N>(i, j)) * kLanes] = T(0.5) * (cov[(o * N + i) * N + j] + cov[(o * N + j) * N + i]);
}

template <typename T>
static int expand_state_api(const T* state, long B, long Tn, int n, const T* A, const T* Q, const T* S0, int shared_mask,
                            T* mean, T* cov, T* pred, void* stream) {
  if (!state || !mean || !cov || B <= 0 || Tn <= 0 || (pred && (!A || !Q || !S0))) { set_error("expand_state: bad argument"); return EKSB_EARG; }
  unsigned grid = (unsigned)(((Tn + 7) / 8) * num_tiles(B));
  cudaStream_t st = (cudaStream_t)stream;
  switch (n) {
#define CASE(N_) case N_: expand_state_kernel<T, N_><<<grid, 256, 0, st>>>(state, B, Tn, A, Q, S0, shared_mask, mean, cov, pred); break;
    CASE(1) CASE(2) CASE(3) CASE(4) CASE(5) CASE(6) CASE(7) CASE(8)
#undef CASE
    default: set_error("expand_state: n=%d not instantiated", n); return EKSB_EDIM;
  }
  return check_launch("expand_state_kernel");
}

template <typename T>
static int pack_state_api(const T* mean, const T* cov, long B, long Tn, int n, T* state, void* stream) {
  if (!state || !mean || !cov || B <= 0 || Tn <= 0) { set_error("pack_state: bad argument"); return EKSB_EARG; }
  unsigned grid = (unsigned)(((Tn + 7) / 8) * num_tiles(B));
  cudaStream_t st = (cudaStream_t)stream;
  switch (n) {
#define CASE(N_) case N_: pack_state_kernel<T, N_><<<grid, 256, 0, st>>>(mean, cov, B, Tn, state); break;
    CASE(1) CASE(2) CASE(3) CASE(4) CASE(5) CASE(6) CASE(7) CASE(8)
#undef CASE
    default: set_error("pack_state: n=%d not instantiated", n); return EKSB_EDIM;
  }
  return check_launch("pack_state_kernel");
}

// ym = C * mean: rows [0,n) of a state record buffer -> p-row records (y_m = C ms, multiview_pca_smoother.py:745;
// y_m = B q for the Newton variant, :905).  One warp per (tile, frame); runtime n <= 8, p <= 16.
template <typename T>
__global__ void project_kernel(const T* __restrict__ state, int rows, const T* __restrict__ C, int shared_mask, long B,
                               long Tn, int n, int p, T* __restrict__ ym) {
  const int lane = threadIdx.x & 31;
  const long nchunk = (Tn + 7) / 8;
  const long g = blockIdx.x / nchunk;
  const long t = (blockIdx.x - g * nchunk) * 8 + (threadIdx.x >> 5);
  if (t >= Tn) return;
  const long b = g * kLanes + lane;
  const long bp = b < B ? b : B - 1;
  const T* c = C + ((shared_mask & EKSB_SHARED_C) ? 0 : bp * p * n);
  T m[8];
  for (int i = 0; i < n; ++i) m[i] = state[((g * Tn + t) * rows + i) * kLanes + lane];
  for (int j = 0; j < p; ++j) {
    T s = T(0);
    for (int i = 0; i < n; ++i) s += c[j * n + i] * m[i];
    ym[((g * Tn + t) * p + j) * kLanes + lane] = s;
  }
}

template <typename T>
static int project_api(const T* state, int rows, const T* C, int shared_mask, long B, long Tn, int n, int p, T* ym, void* stream) {
  if (!state || !C || !ym || B <= 0 || Tn <= 0 || n < 1 || n > 8 || p < 1 || p > 16 || rows < n) { set_error("project: bad argument"); return EKSB_EARG; }
  unsigned grid = (unsigned)(((Tn + 7) / 8) * num_tiles(B));
  project_kernel<T><<<grid, 256, 0, (cudaStream_t)stream>>>(state, rows, C, shared_mask, B, Tn, n, p, ym);
  return check_launch("project_kernel");
}

// z[t,j] = (y[t,j] - (C ms[t])_j) / sqrt((C Vs[t] C^T)_jj + add_ev * ev[t,j])   (build-side
// definition, the reference has no z-score).  One warp per (tile, frame); runtime n <= 8, p <= 16.
template <typename T>
__global__ void zscore_kernel(const T* __restrict__ obs, const T* __restrict__ smth, const T* __restrict__ C,
                              int shared_mask, long B, long Tn, int n, int p, int add_ev, T* __restrict__ z) {
  const int lane = threadIdx.x & 31;
  const long nchunk = (Tn + 7) / 8;
  const long g = blockIdx.x / nchunk;
  const long t = (blockIdx.x - g * nchunk) * 8 + (threadIdx.x >> 5);
  if (t >= Tn) return;
  const long b = g * kLanes + lane;
  const long bp = b < B ? b : B - 1;
  const T* c = C + ((shared_mask & EKSB_SHARED_C) ? 0 : bp * p * n);
  const int sg = n + n * (n + 1) / 2;
  T m[8], V[8][8];
  const T* rec = smth + ((g * Tn + t) * sg) * kLanes + lane;
  for (int i = 0; i < n; ++i) m[i] = rec[i * kLanes];
  int k = n;
  for (int i = 0; i < n; ++i)
    for (int j = i; j < n; ++j) { V[i][j] = rec[k * kLanes]; V[j][i] = V[i][j]; ++k; }
  const T* o = obs + ((g * Tn + t) * 2 * p) * kLanes + lane;
  for (int j = 0; j < p; ++j) {
    T mu = T(0), var = T(0);
    for (int i = 0; i < n; ++i) {
      mu += c[j * n + i] * m[i];
      T a = T(0);
      for (int l = 0; l < n; ++l) a += V[i][l] * c[j * n + l];
      var += c[j * n + i] * a;
    }
    if (add_ev) var += o[(p + j) * kLanes];
    z[((g * Tn + t) * p + j) * kLanes + lane] = (o[j * kLanes] - mu) / sqrt(var);
  }
}

// Compile-time (n, p): C and the frame records live in registers, one warp handles kZF consecutive frames of a
// tile with all their loads issued before the first use (the runtime-dimension kernel above keeps m, V in local
// memory and re-reads C for every frame: 2.1 TB/s on the bench shape).
template <int N, int P> struct ZFrames {   // frames per warp: bounded by the registers their records take
  static constexpr int value = (N + N * (N + 1) / 2 + 2 * P) > 20 ? ((N + N * (N + 1) / 2 + 2 * P) > 30 ? 1 : 2) : 4;
};
template <typename T, int N, int P>
__global__ void __launch_bounds__(256)
zscore_dim_kernel(const T* __restrict__ obs, const T* __restrict__ smth, const T* __restrict__ C, int shared_mask, long B,
                  long Tn, int add_ev, T* __restrict__ z) {
  constexpr int NS = N * (N + 1) / 2, SG = N + NS, kZF = ZFrames<N, P>::value;
  const int lane = threadIdx.x & 31;
  const long nblk = (Tn + kZF - 1) / kZF;
  const long w = (long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (w >= num_tiles_dev(B) * nblk) return;
  const long g = w / nblk;
  const long t0 = (w - g * nblk) * kZF;
  const long b = g * kLanes + lane;
  const long bp = b < B ? b : B - 1;
  const T* cp = C + ((shared_mask & EKSB_SHARED_C) ? 0 : bp * P * N);
  T c[P][N];
#pragma unroll
  for (int j = 0; j < P; ++j)
#pragma unroll
    for (int i = 0; i < N; ++i) c[j][i] = cp[j * N + i];
  const int nf = (int)(Tn - t0 < kZF ? Tn - t0 : kZF);
  T st[kZF][SG], ob[kZF][2 * P];
#pragma unroll
  for (int f = 0; f < kZF; ++f)
    if (f < nf) {
      const T* rec = smth + ((g * Tn + t0 + f) * SG) * kLanes + lane;
      const T* o = obs + ((g * Tn + t0 + f) * 2 * P) * kLanes + lane;
#pragma unroll
      for (int r = 0; r < SG; ++r) st[f][r] = rec[r * kLanes];
#pragma unroll
      for (int r = 0; r < 2 * P; ++r) ob[f][r] = o[r * kLanes];
    }
#pragma unroll
  for (int f = 0; f < kZF; ++f)
    if (f < nf) {
      T* out = z + ((g * Tn + t0 + f) * P) * kLanes + lane;
#pragma unroll
      for (int j = 0; j < P; ++j) {
        T mu = T(0), var = T(0);
#pragma unroll
        for (int i = 0; i < N; ++i) {
          mu += c[j][i] * st[f][i];
          T a = T(0);
#pragma unroll
          for (int l = 0; l < N; ++l) a += st[f][N + sym_idx<N>(i, l)] * c[j][l];
          var += c[j][i] * a;
        }
        if (add_ev) var += ob[f][P + j];
        out[j * kLanes] = (ob[f][j] - mu) * rsqrt(var);          // MUFU seed + Newton: ~1 ulp, a third of div + sqrt in f64
      }
    }
}

template <typename T>
static int zscore_api(const T* obs, const T* smth, const T* C, int shared_mask, long B, long Tn, int n, int p, int add_ev,
                      T* z, void* stream) {
  if (!obs || !smth || !C || !z || B <= 0 || Tn <= 0 || n < 1 || n > 8 || p < 1 || p > 16) { set_error("zscore: bad argument"); return EKSB_EARG; }
#define X(N_, P_)                                                                                                      \
  if (n == N_ && p == P_) {                                                                                            \
    const long warps = num_tiles(B) * ((Tn + ZFrames<N_, P_>::value - 1) / ZFrames<N_, P_>::value);                    \
    zscore_dim_kernel<T, N_, P_><<<(unsigned)((warps + 7) / 8), 256, 0, (cudaStream_t)stream>>>(obs, smth, C, shared_mask, B, \
                                                                                               Tn, add_ev, z);       \
    return check_launch("zscore_dim_kernel");                                                                          \
  }
  EKSB_FOR_EACH_DIM(X)
#undef X
  unsigned grid = (unsigned)(((Tn + 7) / 8) * num_tiles(B));
  zscore_kernel<T><<<grid, 256, 0, (cudaStream_t)stream>>>(obs, smth, C, shared_mask, B, Tn, n, p, add_ev, z);
  return check_launch("zscore_kernel");
}

// out[g] = sum_b in[g][b]: one CTA per row, fixed-order tree so the result does not depend on timing.
// Feeds the NCCL all-reduce of the smoothing-parameter grid search.
__global__ void __launch_bounds__(256) reduce_rows_kernel(const double* __restrict__ in, long B, double* __restrict__ out) {
  __shared__ double part[256];
  const double* row = in + (long)blockIdx.x * B;
  double s = 0.0;
  for (long i = threadIdx.x; i < B; i += 256) s += row[i];
  part[threadIdx.x] = s;
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) {
    if (threadIdx.x < w) part[threadIdx.x] += part[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[blockIdx.x] = part[0];
}

}  // namespace eksb

using namespace eksb;

extern "C" int eksb_reduce_rows_f64(const double* in, int rows, long B, double* out, void* stream) {
  if (!in || !out || rows <= 0 || B <= 0) { set_error("reduce_rows: bad argument"); return EKSB_EARG; }
  reduce_rows_kernel<<<rows, 256, 0, (cudaStream_t)stream>>>(in, B, out);
  return check_launch("reduce_rows_kernel");
}
extern "C" int eksb_zscore_f64(const double* obs, const double* smth, const double* C, int shared_mask, long B, long T, int n,
                               int p, int add_ev, double* z, void* stream) {
  return zscore_api<double>(obs, smth, C, shared_mask, B, T, n, p, add_ev, z, stream);
}
extern "C" int eksb_zscore_f32(const float* obs, const float* smth, const float* C, int shared_mask, long B, long T, int n,
                               int p, int add_ev, float* z, void* stream) {
  return zscore_api<float>(obs, smth, C, shared_mask, B, T, n, p, add_ev, z, stream);
}

extern "C" int eksb_version(void) { return 100; }
extern "C" const char* eksb_last_error_string(void) { return g_err; }
extern "C" long eksb_num_tiles(long B) { return num_tiles(B); }
extern "C" int eksb_supported_dims(int n, int p) {
#define X(N_, P_) if (n == N_ && p == P_) return 1;
  EKSB_FOR_EACH_DIM(X)
#undef X
  return 0;
}

extern "C" int eksb_pack_rows_f64(const double* src, long B, long T, int R, double* dst, int rows, int row0, double pad, void* stream) {
  return pack_rows_api<double>(src, B, T, R, dst, rows, row0, pad, stream, false);
}
extern "C" int eksb_pack_rows_f32(const float* src, long B, long T, int R, float* dst, int rows, int row0, float pad, void* stream) {
  return pack_rows_api<float>(src, B, T, R, dst, rows, row0, pad, stream, false);
}
extern "C" int eksb_unpack_rows_f64(const double* src, int rows, int row0, long B, long T, int R, double* dst, void* stream) {
  return pack_rows_api<double>(src, B, T, R, dst, rows, row0, 0.0, stream, true);
}
extern "C" int eksb_unpack_rows_f32(const float* src, int rows, int row0, long B, long T, int R, float* dst, void* stream) {
  return pack_rows_api<float>(src, B, T, R, dst, rows, row0, 0.f, stream, true);
}
extern "C" int eksb_expand_state_f64(const double* state, long B, long T, int n, const double* A, const double* Q, const double* S0,
                                     int shared_mask, double* mean, double* cov, double* pred, void* stream) {
  return expand_state_api<double>(state, B, T, n, A, Q, S0, shared_mask, mean, cov, pred, stream);
}
extern "C" int eksb_expand_state_f32(const float* state, long B, long T, int n, const float* A, const float* Q, const float* S0,
                                     int shared_mask, float* mean, float* cov, float* pred, void* stream) {
  return expand_state_api<float>(state, B, T, n, A, Q, S0, shared_mask, mean, cov, pred, stream);
}
extern "C" int eksb_pack_state_f64(const double* mean, const double* cov, long B, long T, int n, double* state, void* stream) {
  return pack_state_api<double>(mean, cov, B, T, n, state, stream);
}
extern "C" int eksb_pack_state_f32(const float* mean, const float* cov, long B, long T, int n, float* state, void* stream) {
  return pack_state_api<float>(mean, cov, B, T, n, state, stream);
}
extern "C" int eksb_project_f64(const double* state, int rows, const double* C, int shared_mask, long B, long T, int n, int p, double* ym, void* stream) {
  return project_api<double>(state, rows, C, shared_mask, B, T, n, p, ym, stream);
}
extern "C" int eksb_project_f32(const float* state, int rows, const float* C, int shared_mask, long B, long T, int n, int p, float* ym, void* stream) {
  return project_api<float>(state, rows, C, shared_mask, B, T, n, p, ym, stream);
}

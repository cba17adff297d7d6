// Shared host-side helpers for the libeksb translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdio.h>
#include "../../include/eksb.h"
#include "eksb_dims.h"
#include "eksb_math.cuh"

namespace eksb {

void set_error(const char* fmt, ...);

inline int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return EKSB_ECUDA;
  }
  return EKSB_OK;
}

// Multiprocessors of the CURRENT device (148 on B200), queried once per device: every launch-shape decision
// (tiles per CTA, virtual tiles of the time-split scan, grid caps) scales with it instead of a literal.
int sm_count();

__host__ __device__ inline long num_tiles(long B) { return (B + kLanes - 1) / kLanes; }
__device__ __forceinline__ long num_tiles_dev(long B) { return (B + kLanes - 1) / kLanes; }

// Virtual tiles of the time-split pipelined scan (eksb_scan.cuh): with nc > 1 every (track tile, chunk of L frames)
// pair is a tile of its own for the pipelined kernels.  cut[g*nc + c] = warm-up frames K_c of chunk c (forward: from
// the chunk's start, backward: from its end); init = workspace records with the backward start states; the kernels
// exit at once when *gate != 0 (some chunk never forgot its start: the fallback kernels do the pass).  nc <= 1: off.
struct VTiles {
  long L = 0, nc = 0;
  const long* cut = nullptr;
  const void* init = nullptr;
  const int* gate = nullptr;
};

// Warp-specialised (pipelined) variants, defined in eksb_pipe_*.cu.  Return EKSB_EDIM when the
// (n,p) pair has no pipelined instantiation (the caller then uses the plain kernel).
template <typename T>
int kf_forward_pipe(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int shared_mask, long B,
                    long Tn, int n, int p, int a_mode, int first, T* filt, int* status, const T* yoff, cudaStream_t st);
template <typename T>
int rts_backward_pipe(const T* filt, const T* C, const T* A, const T* Q, int shared_mask, long B, long Tn, int n, int p,
                      int a_mode, T* smth, T* ym, T* yv, int* status, cudaStream_t st);
// NLL only (no state records), n_scale candidates of q_scale * Q per launch, in the warp-specialised form
// (eksb_nll_pipe.cuh); EKSB_EDIM for shapes it is not instantiated for (dense A is the caller's check).
template <typename T>
int kf_nll_pipe(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int shared_mask, long B, long Tn,
                int n, int p, int first, const T* q_scale, int n_scale, double* nll, int* status, const T* yoff, cudaStream_t st);
// The same kernels over virtual tiles (vt.nc > 1); diagonal-R Kalman filter / RTS smoother only.
template <typename T>
int kf_forward_pipe_vt(const T* obs, const T* m0, const T* S0, const T* C, const T* A, const T* Q, int shared_mask, long B,
                       long Tn, int n, int p, int a_mode, int first, T* filt, int* status, const T* yoff, VTiles vt,
                       cudaStream_t st);
template <typename T>
int rts_backward_pipe_vt(const T* filt, const T* C, const T* A, const T* Q, int shared_mask, long B, long Tn, int n, int p,
                         int a_mode, T* smth, T* ym, int* status, VTiles vt, cudaStream_t st);
size_t pipe_vt_target_tiles(int n, int p, int elem_size);   // virtual tiles that fill the GPU for this shape (148 x tiles per SM)
// Policy: the pipelined kernels run every (n,p) they are instantiated for (EKSB_PIPE=0 disables them).  They
// are at their best with up to 4 tiles per SM, so a batch of more than EKSB_PIPE_MAX_TILES (592) tiles goes
// out as several launches over equal groups of about EKSB_PIPE_GROUP_TILES (512) tiles, back to back on the
// stream (pipe_groups): f64 65 536 tracks x 8 000 frames 6.41 TB/s algorithmic, against 4.72 with the plain
// thread-per-track kernels, which remain for NLL sweeps, CV output and uninstantiated (n,p).
bool use_pipe(long B);
long pipe_groups(long tiles, int elem_size);   // number of launches a batch of `tiles` tiles is split into

// Per-track status word: bits are OR-ed in (several warps / launches may report on the same track).
__device__ __forceinline__ void set_status(int* st, bool ok, bool nonpos_var) {
  const int bits = (ok ? 0 : kStatusPivot) | (nonpos_var ? kStatusNonPosVar : 0);
  if (bits) atomicOr(st, bits);
}

// Device-side gather of one track's parameters into registers.
template <typename T, int N, int P, int AM>
__device__ __forceinline__ void load_track_params(Params<T, N, P, AM>& prm, const T* C, const T* A, const T* Q,
                                                  int shared_mask, long b) {
  const T* c = C + ((shared_mask & EKSB_SHARED_C) ? 0 : b * (P * N));
  const T* a = A + ((shared_mask & EKSB_SHARED_A) ? 0 : b * (N * N));
  const T* q = Q + ((shared_mask & EKSB_SHARED_Q) ? 0 : b * (N * N));
  load_params<T, N, P, AM>(prm, c, a, q);
}

}  // namespace eksb

/* eksb.h -- C ABI of libeksb.so, the sm_100a (B200) implementation of the EKS hot path.
 *
 * The reference (erialc-cal/eks) is pure Python/numpy and has no FFI; its boundary for
 * this path is four module-level functions plus the Newton variants.  Each entry point
 * below names the reference function (file:line, relative to the reference checkout)
 * whose arithmetic it replaces.  The Python package eks_b200 binds these with ctypes
 * (eks_b200/_lib.py) and re-exports the reference's own names and signatures;
 * INTEGRATION.md shows the stub a reference maintainer would add.
 *
 * Conventions
 *  - Every pointer is a DEVICE pointer unless the name ends in _host (only the eksb_csv_* functions).  The caller owns all
 *    buffers; the library never allocates.  `stream` is a cudaStream_t passed as void*.
 *  - All calls are asynchronous on `stream` and return 0, or a negative EKSB_E* code
 *    (message via eksb_last_error_string()).  Numerical trouble inside a track (a
 *    non-positive pivot: the reference would raise numpy.linalg.LinAlgError) is reported
 *    per track through the optional `status` array, a bit mask OR-ed into the caller's (zeroed) words:
 *      EKSB_STATUS_PIVOT (1)      a non-positive pivot was seen;
 *      EKSB_STATUS_NONPOS_VAR (2) a frame of the track has an ensemble variance <= 0.  The fast kernels update in
 *                                 information form (they divide by every variance), so their output for such a
 *                                 track is NaN where the reference -- which only needs R + C S C^T to be invertible
 *                                 (eks/ensemble_kalman.py:112) -- stays finite: re-run the track through
 *                                 eksb_kf_forward_generic, which does the reference's p x p LU arithmetic.
 *  - _f32 / _f64 suffix = arithmetic and storage type of every non-integer buffer.
 *  - "Track-tile" layout (all per-frame device arrays): tracks are grouped in tiles of
 *    EKSB_LANES = 32; a frame record has ROWS scalars per track;
 *        index(tile g, frame t, row r, lane l) = ((g*T + t)*ROWS + r)*32 + l ,  track b = 32 g + l
 *    so warp lane == track and every warp access is one full contiguous line.
 *    n_tiles = eksb_num_tiles(B) = ceil(B/32); buffers hold n_tiles*T*ROWS*32 scalars.
 *      obs   record: ROWS = 2p : rows [0,p) = y (ensemble mean/median, centred), [p,2p) = ensemble variance
 *      state record: ROWS = n + n(n+1)/2 : rows [0,n) = mean, then the covariance packed
 *                    upper-triangular row-major ((0,0),(0,1),..,(0,n-1),(1,1),..)
 *  - Per-track parameters are dense row-major arrays with a leading track dimension,
 *    m0 (B,n)  S0 (B,n,n)  C (B,p,n)  A (B,n,n)  Q (B,n,n); a bit set in `shared_mask`
 *    (EKSB_SHARED_*) means that array has no track dimension and is shared by all tracks.
 *  - Supported (n,p): (2,4) (3,4) (3,6) (3,8) (4,6) (6,8); anything else -> EKSB_EDIM.
 *  - a_mode: EKSB_A_DIAG if every A is diagonal (identity included; every shipped config),
 *    else EKSB_A_DENSE.  R is diag(ensemble variance): the reference overwrites diag(R) each
 *    frame (eks/ensemble_kalman.py:88-89,99-100) and every caller passes R = I, so
 *    off-diagonals are zero.
 */
#ifndef EKSB_H_
#define EKSB_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define EKSB_API __attribute__((visibility("default")))
#else
#define EKSB_API
#endif

#define EKSB_LANES 32

#define EKSB_OK 0
#define EKSB_EDIM (-1)    /* unsupported (n,p) or M */
#define EKSB_EARG (-2)    /* bad argument (null pointer, non-positive size, bad mode) */
#define EKSB_ECUDA (-3)   /* CUDA runtime error; see eksb_last_error_string() */

#define EKSB_STATUS_PIVOT 1
#define EKSB_STATUS_NONPOS_VAR 2

#define EKSB_SHARED_M0 1
#define EKSB_SHARED_S0 2
#define EKSB_SHARED_C 4
#define EKSB_SHARED_A 8
#define EKSB_SHARED_Q 16
#define EKSB_SHARED_ALL 31

#define EKSB_A_DIAG 0
#define EKSB_A_DENSE 1

#define EKSB_MODE_MEDIAN 0
#define EKSB_MODE_MEAN 1

/* first-frame behaviour of the forward recursion */
#define EKSB_FIRST_PRIOR_UPDATE 1 /* filtering_pass: frame 0 = update from (m0,S0), no predict (eks/ensemble_kalman.py:94-96) */
#define EKSB_FIRST_PRIOR_ONLY 2   /* kalman_newton_recursive: q[0]=mu0, P=inv(S0), y[0] unused (eks/newton_eks.py:128,132) */

EKSB_API int eksb_version(void);
EKSB_API const char* eksb_last_error_string(void);
EKSB_API long eksb_num_tiles(long B);
EKSB_API int eksb_supported_dims(int n, int p); /* 1 if (n,p) is instantiated */

/* ---- A1: ensemble()  eks/ensemble_kalman.py:4-57 -------------------------------------
 * preds is the raw prediction tensor [M][T][K] (models x frames x keypoint columns, K fastest).
 * avg = median or mean over models (:45), var = population variance / M (:46).  Medians and
 * means are bit-exact w.r.t. numpy (same summation order, no FMA contraction); NaN propagates.
 * 1 <= M <= 64 (register sort network up to 16 models, a local-memory insertion sort above). */
EKSB_API int eksb_ensemble_f64(const double* preds, int M, long T, long K, int mode, double* avg, double* var, void* stream);
EKSB_API int eksb_ensemble_f32(const float* preds, int M, long T, long K, int mode, float* avg, float* var, void* stream);

/* Same reduction, fused with the multi-view restack (multiview_pca_smoother.py:645-682):
 * column k = b*p + j of preds is observation j of track b; writes the obs record (track-tile
 * layout) directly: y = avg - offset[b][j] (offset (B,p), may be null), ev = var.
 * Padding lanes of the last tile get y = 0, ev = 1. */
EKSB_API int eksb_ensemble_tiled_f64(const double* preds, int M, long T, long B, int p, int mode, const double* offset,
                            double* obs, void* stream);
EKSB_API int eksb_ensemble_tiled_f32(const float* preds, int M, long T, long B, int p, int mode, const float* offset,
                            float* obs, void* stream);

/* ---- layout conversion ------------------------------------------------------------------
 * pack:   src (B,T,R) row-major -> rows [row0,row0+R) of a track-tile buffer with `rows` rows
 *         per record; padding lanes are filled with `pad`.
 * unpack: the inverse (padding lanes dropped). */
EKSB_API int eksb_pack_rows_f64(const double* src, long B, long T, int R, double* dst, int rows, int row0, double pad, void* stream);
EKSB_API int eksb_pack_rows_f32(const float* src, long B, long T, int R, float* dst, int rows, int row0, float pad, void* stream);
EKSB_API int eksb_unpack_rows_f64(const double* src, int rows, int row0, long B, long T, int R, double* dst, void* stream);
EKSB_API int eksb_unpack_rows_f32(const float* src, int rows, int row0, long B, long T, int R, float* dst, void* stream);

/* Expand a state record buffer into the reference's dense return values
 * (eks/ensemble_kalman.py:91-93,107): mean (B,T,n), cov (B,T,n,n) and, if pred != NULL,
 * pred (B,T,n,n) with pred[t] = A cov[t] A^T + Q for t < T-1 and pred[T-1] = 0 (the
 * reference's shifted `S`; for T == 1 pred[0] = S0).  A/Q/S0 follow shared_mask. */
EKSB_API int eksb_expand_state_f64(const double* state, long B, long T, int n, const double* A, const double* Q, const double* S0,
                          int shared_mask, double* mean, double* cov, double* pred, void* stream);
EKSB_API int eksb_expand_state_f32(const float* state, long B, long T, int n, const float* A, const float* Q, const float* S0,
                          int shared_mask, float* mean, float* cov, float* pred, void* stream);
/* The inverse for callers that hold dense mf (B,T,n) / Vf (B,T,n,n) (smooth_backward's inputs). */
EKSB_API int eksb_pack_state_f64(const double* mean, const double* cov, long B, long T, int n, double* state, void* stream);
EKSB_API int eksb_pack_state_f32(const float* mean, const float* cov, long B, long T, int n, float* state, void* stream);

/* ---- A2/A3: filtering_pass() + kalman_dot()  eks/ensemble_kalman.py:59-117 -------------
 * Forward Kalman filter over all B tracks, one track per thread, whole state in registers.
 * The measurement update is done in information form (n x n inverses; algebraically the
 * update of :104-105).  Outputs, each optional (NULL to skip):
 *   filt   state records (mf, Vf)                               [n_tiles][T][n+n(n+1)/2][32]
 *   nll    (n_scale, B) doubles: 0.5*sum_t[log det Sig_t + nu^T Sig_t^-1 nu + p log 2pi],
 *          Sig_t = R_t + C S_pred C^T (:112), nu_t the innovation of :102
 *   status (B) ints
 * q_scale (device, n_scale values) evaluates the pass once per smoothing-parameter candidate
 * with Q replaced by q_scale[g]*Q (grid search over `s`, multiview_pca_smoother.py:728);
 * pass NULL / 0 for a single evaluation with Q as given.  filt must be NULL if n_scale > 1.
 * first = EKSB_FIRST_PRIOR_UPDATE for filtering_pass; EKSB_FIRST_PRIOR_ONLY turns the same
 * recursion into kalman_newton_recursive (A6, eks/newton_eks.py:115-148: B=C, E=Q, q = mean rows). */
EKSB_API int eksb_kf_forward_f64(const double* obs, const double* m0, const double* S0, const double* C, const double* A,
                        const double* Q, int shared_mask, long B, long T, int n, int p, int a_mode, int first,
                        const double* q_scale, int n_scale, double* filt, double* nll, int* status, void* stream);
EKSB_API int eksb_kf_forward_f32(const float* obs, const float* m0, const float* S0, const float* C, const float* A,
                        const float* Q, int shared_mask, long B, long T, int n, int p, int a_mode, int first,
                        const float* q_scale, int n_scale, float* filt, double* nll, int* status, void* stream);

/* The same filter on UNCENTRED observations: y_offset (B,p), one row per track, is subtracted from y as the frames
 * are read (y - offset, the rounding of a separate centring pass), so the camera means of
 * multiview_pca_smoother.py:698-705 need no pass of their own over the records (eksb_pca_setup with center = 0).
 * y_offset may be NULL (plain eksb_kf_forward). */
EKSB_API int eksb_kf_forward_offset_f64(const double* obs, const double* y_offset, const double* m0, const double* S0,
                        const double* C, const double* A, const double* Q, int shared_mask, long B, long T, int n, int p,
                        int a_mode, int first, double* filt, double* nll, int* status, void* stream);
EKSB_API int eksb_kf_forward_offset_f32(const float* obs, const float* y_offset, const float* m0, const float* S0,
                        const float* C, const float* A, const float* Q, int shared_mask, long B, long T, int n, int p,
                        int a_mode, int first, float* filt, double* nll, int* status, void* stream);

/* ---- A4/A5: smooth_backward() + back-projection  eks/ensemble_kalman.py:120-164 --------
 * RTS pass over state records written by eksb_kf_forward.  S = A Vf A^T + Q is recomputed
 * per frame instead of being read back (:101,:158).  smth may equal filt (in-place).
 * Optional outputs (NULL to skip), all track-tile layout:
 *   ym (p rows)   C ms                       (multiview_pca_smoother.py:745, pupil_smoother.py:190)
 *   yv (p rows)   diag(C Vs C^T)             (:746)
 *   cv (n*n rows) CV[t] = Vs[t+1] J^T, frames 0..T-2 (:162); frame T-1 is not written */
EKSB_API int eksb_rts_backward_f64(const double* filt, const double* C, const double* A, const double* Q, int shared_mask,
                          long B, long T, int n, int p, int a_mode, double* smth, double* ym, double* yv, double* cv,
                          int* status, void* stream);
EKSB_API int eksb_rts_backward_f32(const float* filt, const float* C, const float* A, const float* Q, int shared_mask,
                          long B, long T, int n, int p, int a_mode, float* smth, float* ym, float* yv, float* cv,
                          int* status, void* stream);

/* ---- generic (runtime-dimension) kernels: literal reference arithmetic ------------------
 * kalman_dot: out (n,k) = V C^T (R + C V C^T)^-1 array, array (p,k), via a p x p LU solve with
 * partial pivoting (eks/ensemble_kalman.py:110-117; numpy.linalg.solve -> LAPACK gesv).
 * n <= 8, p <= 16, k <= 8.  *status = 1 on an exactly singular matrix (LinAlgError). */
EKSB_API int eksb_kalman_dot_f64(const double* array, int k, const double* V, const double* C, const double* R, int n, int p,
                        double* out, int* status, void* stream);
EKSB_API int eksb_kalman_dot_f32(const float* array, int k, const float* V, const float* C, const float* R, int n, int p,
                        float* out, int* status, void* stream);
/* filtering_pass with a dense R (B,p,p) or shared (p,p) [r_shared=1] whose off-diagonals take
 * part unchanged (eks/ensemble_kalman.py:112) and/or an (n,p) pair that is not instantiated
 * (n <= 8, p <= 16).  R == NULL means identity off-diagonals.  Same obs / state records. */
EKSB_API int eksb_kf_forward_generic_f64(const double* obs, const double* m0, const double* S0, const double* C, const double* A,
                                const double* Q, const double* R, int r_shared, int shared_mask, long B, long T, int n,
                                int p, int first, double* filt, int* status, void* stream);
EKSB_API int eksb_kf_forward_generic_f32(const float* obs, const float* m0, const float* S0, const float* C, const float* A,
                                const float* Q, const float* R, int r_shared, int shared_mask, long B, long T, int n,
                                int p, int first, float* filt, int* status, void* stream);

/* smooth_backward in the reference's literal arithmetic (eks/ensemble_kalman.py:152-162: J = solve(S, A Vf)^T by LU
 * with partial pivoting) for any n <= 8, p <= 16 and dense A -- the backward twin of eksb_kf_forward_generic, so
 * every shape filtering_pass accepts can also be smoothed (5 cameras: n = 3..5, p = 10).  Same records and optional
 * outputs (ym, yv, cv) as eksb_rts_backward; smth may alias filt only when cv == NULL and pred == NULL is not
 * required (a frame is read before it is written).  pred: optional NS-row packed records with the caller's S[t]
 * (the reference reads its S argument and ignores Q); NULL recomputes S[t] = A Vf[t] A^T + Q (:101). */
EKSB_API int eksb_rts_backward_generic_f64(const double* filt, const double* pred, const double* C, const double* A,
                                  const double* Q, int shared_mask, long B, long T, int n, int p, double* smth,
                                  double* ym, double* yv, double* cv, int* status, void* stream);
EKSB_API int eksb_rts_backward_generic_f32(const float* filt, const float* pred, const float* C, const float* A,
                                  const float* Q, int shared_mask, long B, long T, int n, int p, float* smth,
                                  float* ym, float* yv, float* cv, int* status, void* stream);

/* ym = C * mean over a state record buffer with `rows` rows per record (mean = rows [0,n)):
 * y_m_smooth = C ms (multiview_pca_smoother.py:745) or B q for the Newton variant (:905).
 * ym: p-row track-tile records.  n <= 8, p <= 16. */
EKSB_API int eksb_project_f64(const double* state, int rows, const double* C, int shared_mask, long B, long T, int n, int p,
                     double* ym, void* stream);
EKSB_API int eksb_project_f32(const float* state, int rows, const float* C, int shared_mask, long B, long T, int n, int p,
                     float* ym, void* stream);

/* ---- A7: block-tridiagonal Newton solve -------------------------------------------------
 * One Newton step from q = 0 (the exact MAP path) of the EKS posterior's block-tridiagonal system
 * (spec: ref/Kalman_smoother.pdf p.1, Simulations.ipynb[16-20]; the shipped eks/block_tridiag.py:4-43
 * raises NameError).  Block-Thomas forward elimination + back-substitution, both as TMA-fed
 * warp-specialised kernels.  Parameter names follow eks/block_tridiag.py:4: A (n,n), Bmat (p,n) the
 * observation matrix (shared_mask bit EKSB_SHARED_C), E (n,n) the process noise (bit EKSB_SHARED_Q);
 * obs holds y and the per-frame ensemble variances D_t (T,p order, unlike :16's (p,T)).
 *   work (state-record layout): (z_t, M_t^-1) forward factors
 *   sol  (state-record layout, may alias work): rows [0,n) = x_t = the Newton step == RTS mean;
 *        rows [n, ..) = diagonal block t of H^-1, packed == RTS covariance
 *   ym   optional p-row records: Bmat x_t
 * State dim n <= 4 (pipelined instantiations only). */
EKSB_API int eksb_btd_solve_f64(const double* obs, const double* m0, const double* S0, const double* A, const double* Bmat,
                       const double* E, int shared_mask, long B, long T, int n, int p, int a_mode, double* work,
                       double* sol, double* ym, int* status, void* stream);
EKSB_API int eksb_btd_solve_f32(const float* obs, const float* m0, const float* S0, const float* A, const float* Bmat,
                       const float* E, int shared_mask, long B, long T, int n, int p, int a_mode, float* work,
                       float* sol, float* ym, int* status, void* stream);

/* ---- A8: z-score and the NLL reduction for smoothing-parameter selection (absent from the
 * reference: build-side definitions; the NLL itself comes from eksb_kf_forward) ----
 * z (p-row records) = (y - C ms) / sqrt(diag(C Vs C^T) + add_ev * ev). */
EKSB_API int eksb_zscore_f64(const double* obs, const double* smth, const double* C, int shared_mask, long B, long T, int n, int p,
                    int add_ev, double* z, void* stream);
EKSB_API int eksb_zscore_f32(const float* obs, const float* smth, const float* C, int shared_mask, long B, long T, int n, int p,
                    int add_ev, float* z, void* stream);
/* out[g] = sum over the B tracks of nll[g][:] (deterministic order): the per-GPU partial sums that
 * are all-reduced (NCCL, SUM) across track shards before the argmin over the grid. */
EKSB_API int eksb_reduce_rows_f64(const double* in, int rows, long B, double* out, void* stream);

/* ---- K4: parallel-in-time filter + RTS smoother for long tracks ------------------------
 * Same inputs and outputs as eksb_kf_forward (first = PRIOR_UPDATE) followed by eksb_rts_backward,
 * computed with an exact chunked scan over time (Sarkka & Garcia-Fernandez elements accumulated per
 * chunk of `chunk` frames, folded per track, every chunk re-run from its exact boundary state), so a
 * handful of very long tracks (config 5: 64 x 10 M frames) still fills the GPU.  A chunk whose summary has
 * forgotten the chunk's boundary state (sensitivity < 1e-22 f64 / 1e-11 f32) writes its remaining records
 * during the summary pass, so the input is read ~once, not twice (EKSB_SCAN_NO_EARLY=1 switches that off).  chunk <= 0 picks
 * the chunk length (eksb_scan_chunk_len).  workspace: eksb_scan_workspace_bytes(...) bytes.
 * smth may be NULL (filter only) or alias filt (in-place).  State dim n <= 4. */
EKSB_API size_t eksb_scan_workspace_bytes(long B, long T, int n, long chunk, int elem_size);
EKSB_API long eksb_scan_chunk_len(long B, long T, int n, long chunk, int elem_size);
EKSB_API int eksb_kf_rts_scan_f64(const double* obs, const double* m0, const double* S0, const double* C, const double* A,
                         const double* Q, int shared_mask, long B, long T, int n, int p, int a_mode, long chunk,
                         double* workspace, size_t workspace_bytes, double* filt, double* smth, double* ym,
                         int* status, void* stream);
EKSB_API int eksb_kf_rts_scan_f32(const float* obs, const float* m0, const float* S0, const float* C, const float* A,
                         const float* Q, int shared_mask, long B, long T, int n, int p, int a_mode, long chunk,
                         float* workspace, size_t workspace_bytes, float* filt, float* smth, float* ym, int* status,
                         void* stream);

/* ---- K8: Kalman parameter setup of the multi-camera smoother on the device --------
 * For every track, from its obs records (y = UNcentred ensemble median, ev = ensemble variance):
 *   good frames = { t : max_j ev[t,j] <= numpy.percentile(max_j ev, quantile) }  (multiview_pca_smoother.py:686-688;
 *                 exact order statistic + numpy's linear interpolation, bit-identical threshold)
 *   means (B,p)   = mean of y over good frames                                    (:698-700)
 *   C (B,p,n)     = components_^T of PCA(n=3) of the centred good rows, rows signed like sklearn's svd_flip (:707-708,:730)
 *   S0 (B,n,n)    = diag(var(good PCs)), ddof 0                                    (:722)
 *   Q (B,n,n)     = smooth_param * cov(difference of consecutive good PCs), ddof 1 (:726-728)
 * (m0 = 0 and A = I are constants, :721,:724).  center != 0 also subtracts means from the y rows in place
 * (:702-705), leaving obs ready for eksb_kf_forward.  status[b] = 1 if a track has < 3 good frames.
 * p in {4,6,8}, n = 3.  workspace: eksb_pca_setup_workspace_bytes(B, T, p, sizeof(scalar)) bytes (mostly the
 * B*T max-variance buffer the radix select re-reads). */
EKSB_API size_t eksb_pca_setup_workspace_bytes(long B, long T, int p, int elem_size);
EKSB_API int eksb_pca_setup_f64(double* obs, long B, long T, int n, int p, double quantile, double smooth_param, int center,
                       double* means, double* C, double* S0, double* Q, double* workspace, size_t workspace_bytes,
                       int* status, void* stream);
EKSB_API int eksb_pca_setup_f32(float* obs, long B, long T, int n, int p, double quantile, double smooth_param, int center,
                       float* means, float* C, float* S0, float* Q, double* workspace, size_t workspace_bytes,
                       int* status, void* stream);

/* Ensemble -> setup hand-over of max_vars (max over the p ensemble variances of a frame, :686): the ensemble kernel
 * has them in registers, so eksb_ensemble_tiled_mx also writes that row ([tile][t][32]) -- straight into the setup
 * workspace at eksb_pca_setup_mx_offset_bytes(...) -- and eksb_pca_setup_mx with mx_ready = 1 skips its own first pass
 * over the records (p rows read per track-frame).  max_var may be NULL; mx_ready = 0 is plain eksb_pca_setup. */
EKSB_API int eksb_ensemble_tiled_mx_f64(const double* preds, int M, long T, long B, int p, int mode, const double* offset,
                        double* obs, double* max_var, void* stream);
EKSB_API int eksb_ensemble_tiled_mx_f32(const float* preds, int M, long T, long B, int p, int mode, const float* offset,
                        float* obs, float* max_var, void* stream);
EKSB_API size_t eksb_pca_setup_mx_offset_bytes(long B, long T, int p, int elem_size);
EKSB_API int eksb_pca_setup_mx_f64(double* obs, long B, long T, int n, int p, double quantile, double smooth_param,
                        int center, int mx_ready, double* means, double* C, double* S0, double* Q, double* workspace,
                        size_t workspace_bytes, int* status, void* stream);
EKSB_API int eksb_pca_setup_mx_f32(float* obs, long B, long T, int n, int p, double quantile, double smooth_param,
                        int center, int mx_ready, float* means, float* C, float* S0, float* Q, double* workspace,
                        size_t workspace_bytes, int* status, void* stream);

/* ---- asynchronous-view alignment (multiview_pca_smoother.py:70-101) -----------------
 * out[s][q][j] = linear interpolation of series s (src (S,Ts,K) row-major, sampled at the increasing times
 * ts_src (Ts)) at query time ts_q[q], with scipy interp1d's arithmetic (slope * (x - x_lo) + y_lo, no FMA:
 * bit-identical); queries outside [ts_src[0], ts_src[Ts-1]] extrapolate from the end interval (the reference
 * drops such frames before calling, :83-86).  flip (K) may be null; where flip[j] != 0 the result is
 * flip[j] - value (the mirror flip x -> img_width - x of :90-91).  out is (S,Tq,K). */
EKSB_API int eksb_interp_rows_f64(const double* ts_src, const double* src, long S, long Ts, int K, const double* ts_q, long Tq,
                         const double* flip, double* out, void* stream);
EKSB_API int eksb_interp_rows_f32(const float* ts_src, const float* src, long S, long Ts, int K, const float* ts_q, long Tq,
                         const float* flip, float* out, void* stream);

/* ---- on-disk format either side of the path: DLC / Lightning-Pose CSVs -------------
 * HOST functions (the only ones in this header taking host pointers; no stream).  File layout: 3 header rows
 * (scorer / bodyparts / coords), an integer index column, then n_cols numeric fields per row -- what the
 * reference reads with pd.read_csv(header=[0,1,2], index_col=0) (scripts/multicam_example.py:85) and writes
 * with DataFrame.to_csv (:139,:157).
 *   shape: numeric body size (index column excluded).
 *   read : body -> data_host (n_rows, n_cols) row-major float64 (correctly rounded, std::from_chars; empty or
 *          "nan" fields -> NaN), index column -> index_host (nullable), the three raw header lines ->
 *          header_host (nullable, NUL-terminated, truncated to header_cap).  n_threads <= 0: all host cores.
 *   write: header lines + rows "index,v0,v1,..." with the shortest round-trip decimal in Python repr layout
 *          (identical text to DataFrame.to_csv), NaN -> empty field. */
EKSB_API int eksb_csv_dlc_shape_host(const char* path, long* n_rows, int* n_cols);
EKSB_API int eksb_csv_dlc_read_host(const char* path, double* data_host, long n_rows, int n_cols, long* index_host,
                           char* header_host, size_t header_cap, int n_threads);
EKSB_API int eksb_csv_dlc_write_host(const char* path, const char* header_host, const long* index_host,
                            const double* data_host, long n_rows, int n_cols);

/* ---- routed filter / smoother: the library picks sequential or parallel-in-time kernels -----------------------
 * Same contracts as eksb_kf_forward_offset (y_offset may be NULL) and eksb_rts_backward, plus a caller-owned
 * workspace of eksb_route_workspace_bytes(B, T, n, p, sizeof(scalar)) bytes (0: this shape always runs the sequential
 * kernels and workspace may be NULL).  The sequential kernels keep one dependency chain per track tile, so a batch with
 * fewer tiles than the GPU has SMs -- a 2 048-track shard of configs[3] on one of 8 GPUs, the per-body-part loop of
 * scripts/multicam_example.py:106 batched into one call -- runs at one warp's pace; with a workspace such batches go
 * through the forward / backward half of eksb_kf_rts_scan instead (same results up to rounding; policy and the
 * measurements behind it: csrc/eksb_scan_impl.cuh route_scan).  Requests the scan cannot serve (nll, yv, cv, a
 * too-small or NULL workspace, an (n,p) without a scan instantiation) take the sequential kernels.  The forward and
 * the backward call of one batch may share the workspace. */
EKSB_API size_t eksb_route_workspace_bytes(long B, long T, int n, int p, int elem_size);
EKSB_API int eksb_kf_forward_ws_f64(const double* obs, const double* y_offset, const double* m0, const double* S0,
                        const double* C, const double* A, const double* Q, int shared_mask, long B, long T, int n, int p,
                        int a_mode, int first, double* filt, double* nll, int* status, double* workspace,
                        size_t workspace_bytes, void* stream);
EKSB_API int eksb_kf_forward_ws_f32(const float* obs, const float* y_offset, const float* m0, const float* S0,
                        const float* C, const float* A, const float* Q, int shared_mask, long B, long T, int n, int p,
                        int a_mode, int first, float* filt, double* nll, int* status, float* workspace,
                        size_t workspace_bytes, void* stream);
EKSB_API int eksb_rts_backward_ws_f64(const double* filt, const double* C, const double* A, const double* Q, int shared_mask,
                          long B, long T, int n, int p, int a_mode, double* smth, double* ym, double* yv, double* cv,
                          int* status, double* workspace, size_t workspace_bytes, void* stream);
EKSB_API int eksb_rts_backward_ws_f32(const float* filt, const float* C, const float* A, const float* Q, int shared_mask,
                          long B, long T, int n, int p, int a_mode, float* smth, float* ym, float* yv, float* cv,
                          int* status, float* workspace, size_t workspace_bytes, void* stream);

/* ---- A8: NLL sweep with a workspace (time-split for few tiles and long tracks) ------------------------------
 * The innovation NLL of filtering_pass (eks/ensemble_kalman.py:98-112) per trial smoothing parameter: the objective
 * of the grid search / an optimiser over `s` that replaces the hand-set CLI flag of
 * scripts/multicam_example.py:38-43 (the reference itself has no such search).  Same contract as the nll output of eksb_kf_forward (q_scale / n_scale candidates
 * of q_scale[g]*Q, y_offset optional, first-frame mode), no state records.  With a caller-owned workspace of
 * eksb_nll_workspace_bytes(...) bytes (0: the plain launch is the route, workspace may be NULL) a batch that leaves the
 * plain kernel with fewer CTAs than ~2 per SM is cut in time as in eksb_kf_rts_scan: a warm-up kernel per candidate
 * finds where each chunk has forgotten its start, CTA (tile, chunk) sums the terms of its frames, the chunks are
 * added in order (deterministic).  One keypoint x 100 000 frames: one chain warp for the whole track otherwise.
 * Candidates whose filter never forgets inside a chunk (tiny process noise) are evaluated whole by a second launch
 * behind a device flag.  Same values up to rounding (tests: 1e-10).  Dense A / other shapes: the sequential kernels. */
EKSB_API size_t eksb_nll_workspace_bytes(long B, long T, int n, int p, int n_scale, int elem_size);
EKSB_API int eksb_kf_nll_ws_f64(const double* obs, const double* y_offset, const double* m0, const double* S0,
                        const double* C, const double* A, const double* Q, int shared_mask, long B, long T, int n, int p,
                        int a_mode, int first, const double* q_scale, int n_scale, double* nll, int* status,
                        void* workspace, size_t workspace_bytes, void* stream);
EKSB_API int eksb_kf_nll_ws_f32(const float* obs, const float* y_offset, const float* m0, const float* S0,
                        const float* C, const float* A, const float* Q, int shared_mask, long B, long T, int n, int p,
                        int a_mode, int first, const float* q_scale, int n_scale, double* nll, int* status,
                        void* workspace, size_t workspace_bytes, void* stream);

/* ---- no-allocation variant of the parallel-in-time scan -------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same inputs, outputs
 * and arithmetic as eksb_kf_rts_scan_f64, but the scan's workspace is a file-scope `static __device__` array of
 * EKSB_SCAN_STATIC_SCRATCH_BYTES inside the library (csrc/eksb_scan_f64.cu) instead of a caller buffer: the call
 * allocates nothing (no cudaMalloc / cudaFree / allocating library routine), and every GPU that loads the library
 * holds that much device memory until the process exits.  The first call on each device looks the array's device
 * address up (cudaGetSymbolAddress, that device current) -- make that first call outside a graph capture.
 * Limit: eksb_scan_workspace_bytes(B, T, n, chunk, 8) <= EKSB_SCAN_STATIC_SCRATCH_BYTES, else EKSB_EARG with a
 * message naming both sizes (the library's own chunk policy needs < 30 MiB for any B with state dim 3; a short
 * explicit chunk on many long tracks can need more).  All calls share the array and it keeps the last call's
 * contents; every call clears on `stream` the words that must start at zero.  Calls that overlap on the GPU
 * (two streams, or unordered host threads) race on it: serialise them. */
#define EKSB_SCAN_STATIC_SCRATCH_BYTES (32u << 20)
EKSB_API int eksb_kf_rts_scan_static_f64(const double* obs, const double* m0, const double* S0, const double* C,
                         const double* A, const double* Q, int shared_mask, long B, long T, int n, int p, int a_mode,
                         long chunk, double* filt, double* smth, double* ym, int* status, void* stream);

/* EKSB_DECLS_END (new declarations go above this line) */
#ifdef __cplusplus
}
#endif
#endif /* EKSB_H_ */

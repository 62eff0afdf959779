/* dml_b200.h -- C ABI of the B200 (sm_100a) candidate x CV-fold sweep library.
 *
 * The reference (sanjita2911/CS230-distributed-machine-learning) has no FFI for
 * this path: its worker calls scikit-learn one fit at a time from Python
 * (aws-prod/worker/worker.py:289-363).  These entry points are what a worker
 * binding replaces those calls with; each comment names the reference call
 * sites it stands in for.  INTEGRATION.md shows the ctypes stub.
 *
 * Conventions
 *  - plain C, no C++/torch types; every array argument is a caller-owned DEVICE
 *    pointer (cudaMalloc / torch.Tensor.data_ptr()) unless marked "host".  The small
 *    per-candidate arrays (C, tol, fit_intercept, max_iter, alpha, class_weight) only
 *    have to be device-ACCESSIBLE: page-locked host memory works through unified
 *    addressing, and this repository's binding passes them that way so that no copy
 *    queues behind a dataset upload;
 *  - row-major X with leading dimension ldx (floats); labels are class indices
 *    0..n_classes-1 in the order of sklearn's classes_ (sorted unique y);
 *  - fold_id[i] in 0..n_folds-1 is the CV fold whose TEST set holds row i
 *    (StratifiedKFold / KFold, un-shuffled -- cross_val_score's choice for an
 *    int cv, worker.py:326,341); hold_id (may be NULL) marks with 1 the rows of
 *    the train_test_split hold-out set (worker.py:302);
 *  - fits are ordered [candidate][k], k = 0..n_folds-1 the CV fits, then (when
 *    hold_id != NULL) k = n_folds the hold-out fit; fits_per_cand = n_folds +
 *    (hold_id != NULL);
 *  - work is issued on the ctx stream (so a caller can order it after copies it
 *    queued on that stream, as worker.py does with a prefetched dataset) and every
 *    call returns after that work has finished: outputs are valid on return; 0 on
 *    success, otherwise a code and a message in dml_last_error;
 *  - one ctx per (process, GPU); a ctx is not re-entrant; scratch memory comes
 *    from an arena owned by the ctx and released by dml_ctx_destroy.
 */
#ifndef DML_B200_H
#define DML_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dml_ctx dml_ctx;

/* Library version (major*10000 + minor*100 + patch). */
int dml_version(void);

/* Create / destroy a context on `device`, issuing work on `stream` (a
 * cudaStream_t, NULL = default stream).  Stands where the worker process starts
 * up and registers (worker.py:156-181). */
int dml_ctx_create(int device, void* stream, dml_ctx** out);
void dml_ctx_destroy(dml_ctx* ctx);
const char* dml_last_error(dml_ctx* ctx); /* ctx may be NULL: last create error */

/* Row-sharded multi-GPU mode: every rank holds a row slice of
 * X / y / fold_id and the full candidate list; the library sums gradients,
 * moments and counts with ncclAllReduce.  `id128` is a 128-byte ncclUniqueId
 * (host memory) produced on rank 0 by dml_nccl_unique_id and broadcast by the
 * caller (torch.distributed / any side channel).  Replaces nothing in the
 * reference (it has no tensor collectives; Kafka carries JSON only). */
int dml_nccl_unique_id(void* id128_host);
int dml_ctx_init_comm(dml_ctx* ctx, int n_ranks, int rank, const void* id128_host);

/* LogisticRegression(solver="lbfgs", penalty="l2") for every candidate and every
 * fold at once.  Replaces, for all candidates of a job: model.fit (worker.py:315),
 * model.predict + accuracy_score (worker.py:322-323) and
 * cross_val_score(model, X, y, cv) (worker.py:326), i.e. sklearn
 * _logistic_regression_path + scipy L-BFGS-B.
 *   C, tol, fit_intercept, max_iter : per candidate (length n_cand), device
 *   coef_out       [n_fits][T][d]  fp32 (T = 1 if n_classes == 2 else n_classes)
 *   intercept_out  [n_fits][T]     fp32
 *   n_iter_out, status_out, nfev_out [n_fits] int32 (status: 0 converged,
 *                  1 max_iter reached, 2 abnormal line search)
 *   correct_out, n_test_out [n_fits] int64: correct predictions / rows of the
 *                  fit's held-out set (summed over ranks in row-sharded mode)
 *   rounds_out     host int32 (may be NULL): batched loss/gradient evaluations
 *   use_tensor_cores : 1 = tcgen05 3xTF32 kernels when the shape allows, 0 = FFMA
 *                  kernels, -1 = library default */
int dml_logreg_sweep(dml_ctx* ctx, const float* X, int64_t n, int d, int64_t ldx,
                     const int32_t* y, int n_classes,
                     const int8_t* fold_id, int n_folds, const int8_t* hold_id,
                     const double* C, const double* tol, const uint8_t* fit_intercept,
                     const int32_t* max_iter, int n_cand,
                     float* coef_out, float* intercept_out,
                     int32_t* n_iter_out, int32_t* status_out, int32_t* nfev_out,
                     int64_t* correct_out, int64_t* n_test_out,
                     int32_t* rounds_out, int use_tensor_cores);

/* dml_logreg_sweep with class weights (LogisticRegression(class_weight=dict | "balanced"), which the reference's
 * client serialises with the other estimator parameters, core.py:96-150, and scikit-learn turns into
 * sample_weight = class_weight_[y], _logistic.py).  class_weight [n_fits][n_classes] float64, device-accessible:
 * the weight of each class for fit (candidate, k) -- "balanced" depends on the training fold, so the caller
 * supplies it per fit: n_train / (n_classes * count of the class in the fit's training rows).  A row's loss and
 * gradient contribution are multiplied by the weight of its class; the objective is sum(s_i loss_i) / sum(s_i)
 * + |w|^2 / (2 C sum(s_i)) (_linear_loss.py:291-379).  Scoring is unweighted accuracy, as cross_val_score's
 * default scorer.  Everything else as dml_logreg_sweep. */
int dml_logreg_sweep_weighted(dml_ctx* ctx, const float* X, int64_t n, int d, int64_t ldx,
                              const int32_t* y, int n_classes,
                              const int8_t* fold_id, int n_folds, const int8_t* hold_id,
                              const double* C, const double* tol, const uint8_t* fit_intercept,
                              const int32_t* max_iter, int n_cand, const double* class_weight,
                              float* coef_out, float* intercept_out,
                              int32_t* n_iter_out, int32_t* status_out, int32_t* nfev_out,
                              int64_t* correct_out, int64_t* n_test_out,
                              int32_t* rounds_out, int use_tensor_cores);

/* Ridge (is_classifier = 0, y_reg used) / RidgeClassifier (is_classifier = 1,
 * y_cls used, targets +-1) with solver "cholesky", every alpha and fold from one
 * pass over X.  Replaces the same three worker call sites for these models
 * (sklearn _ridge.py:215-234, _base.py:189-199).
 *   alpha, fit_intercept : per candidate, device
 *   coef_out [n_fits][T][d], intercept_out [n_fits][T] fp32
 *   classifier: correct_out / n_test_out as above (sse/sy/syy may be NULL)
 *   regressor : sse_out = sum (y - yhat)^2, sy_out = sum y, syy_out = sum y^2 over
 *               the held-out rows (fp64); R^2 = 1 - sse / (syy - sy^2 / n_test)
 *   info_out  [n_fits] int32: 0 ok; j+1 = pivot j of the Cholesky factorisation was not positive (the
 *               coefficients of that fit are meaningless); -(j+1) = pivot j was below 1e-4 of the largest
 *               diagonal entry, i.e. the system is rank-deficient at the accuracy of the fp32 Gram matrix
 *               and the result is not the one LAPACK gives scikit-learn in float64 (min-norm lstsq for
 *               LinearRegression, SVD fallback for Ridge, _ridge.py:753-761).  A caller should hand fits
 *               with info != 0 to the scikit-learn executor (worker.py does). */
int dml_ridge_sweep(dml_ctx* ctx, const float* X, int64_t n, int d, int64_t ldx,
                    const float* y_reg, const int32_t* y_cls, int n_classes, int is_classifier,
                    const int8_t* fold_id, int n_folds, const int8_t* hold_id,
                    const double* alpha, const uint8_t* fit_intercept, int n_cand,
                    float* coef_out, float* intercept_out,
                    int64_t* correct_out, int64_t* n_test_out,
                    double* sse_out, double* sy_out, double* syy_out,
                    int32_t* info_out, int use_tensor_cores);

/* One batched evaluation of the binary-logistic objective for given coefficient
 * columns: for column s, over rows i with label[mask_sel[s]][i] != mask_val[s]
 * (its training rows), out[s] = [ sum_i r_i x_i (d values), sum_i r_i, sum_i loss_i ]
 * with z = x.w_s + b_s, loss = log(1+e^z) - y z, r = sigmoid(z) - y; fp64, length
 * d + 2 per column.  This is sklearn's LinearModelLoss.loss_gradient before the
 * 1/n scaling and the penalty (_linear_loss.py:291-379), i.e. what every L-BFGS
 * function evaluation inside model.fit (worker.py:315,326) computes, for all
 * columns at once.  W [n_slots][d], b [n_slots] fp32. */
int dml_logreg_eval(dml_ctx* ctx, const float* X, int64_t n, int d, int64_t ldx, const int32_t* y,
                    const int8_t* fold_id, const int8_t* hold_id,
                    const float* W, const float* b, const int32_t* mask_sel, const int32_t* mask_val,
                    int n_slots, double* out, int use_tensor_cores);

/* Score given coefficients: per slot (column) s, over rows with
 * label[mask_sel[s]][i] == mask_val[s]: classifier -> correct count of groups of
 * T columns (binary: decision > 0, else first arg-max), regressor -> squared
 * error.  Replaces model.predict + accuracy_score / r2_score (worker.py:322-323,
 * 336-338; sklearn _base.py:391,416-418).
 *   W [n_slots][d] fp32, b [n_slots] fp32, mask_sel/mask_val [n_slots] int32
 *   correct_out [n_slots / T] int64 (classifier) or sse_out [n_slots] fp64 */
int dml_score(dml_ctx* ctx, const float* X, int64_t n, int d, int64_t ldx,
              const int32_t* y_cls, const float* y_reg,
              const int8_t* fold_id, const int8_t* hold_id,
              const float* W, const float* b, const int32_t* mask_sel, const int32_t* mask_val,
              int n_slots, int T, int64_t* correct_out, double* sse_out);

/* dml_score for classifiers, for callers that cannot allocate device memory at the
 * time of the call.  It has no counterpart in the reference.
 *
 * Who needs it: a worker admitted under a fixed device-memory reservation that is
 * already fully handed out, or a caller scoring on a side stream while a sweep is
 * running on the same GPU.  cudaMalloc / cudaFree serialise the device (stalling
 * the sweep) and fail once the reservation is spent; dml_score may call them
 * because its scratch lives in a ctx arena that grows on demand.  This entry point
 * takes no ctx and calls neither: its only scratch, the per-128-row-tile partial
 * counts, sits in a fixed 4 MiB array inside the library image, present on every
 * GPU that loads the library for the life of the process.
 *
 * Limit: ceil(n / 128) * n_slots <= 1 048 576, else DML_ERR_LIMIT (4) is returned
 * with a message in dml_last_error(NULL) and nothing is launched.
 * Arguments and results are those of dml_score (classifier form): W [n_slots][d],
 * b, mask_sel, mask_val device arrays; correct_out [n_slots / T] int64 on device.
 * `device` is made current; `stream` is a cudaStream_t (NULL = default stream);
 * the call returns after the stream has drained, with correct_out valid.
 * The first call on each device looks up the array's device address and sets a
 * kernel attribute (threads that make that first call at the same time are
 * serialised inside the library); make that first call outside of any stream capture.
 * All calls on a device share the one array and the array keeps the previous
 * call's contents (the used part is cleared at the start of each call): two
 * calls that overlap on the same GPU, from any streams or threads, race. */
int dml_score_static(int device, void* stream, const float* X, int64_t n, int d, int64_t ldx,
                     const int32_t* y_cls, const int8_t* fold_id, const int8_t* hold_id,
                     const float* W, const float* b, const int32_t* mask_sel, const int32_t* mask_val,
                     int n_slots, int T, int64_t* correct_out);

/* Timing of the last sweep on this ctx, measured with CUDA events on the ctx
 * stream: total milliseconds, milliseconds inside the contraction kernels
 * (forward + gradient, or Gram), their launch count, and all launches. */
int dml_last_timing(dml_ctx* ctx, float* total_ms, float* contraction_ms,
                    int32_t* contraction_launches, int32_t* total_launches);

/* Work counters of the last logistic sweep: sum over batched rounds of the
 * columns evaluated in that round (each column-round is one pass of X through
 * one column: 4*n_train*d flops), and the number of rounds. */
int dml_last_counters(dml_ctx* ctx, int64_t* column_rounds, int32_t* rounds);

/* Row-sharded mode: time spent in the ncclAllReduce calls of the last sweep on this ctx (CUDA events
 * around each call on the ctx stream; up to 512 calls are timed), their number and the bytes reduced.
 * The reference has no collective; diagnostic only. */
int dml_last_comm(dml_ctx* ctx, float* allreduce_ms, int32_t* allreduce_calls, double* allreduce_bytes);

/* CTAs per cluster the forward / scoring kernel of the last call ran with: 2 = the X tiles were
 * shared by TMA multicast between two column blocks, which the library only does after the driver
 * confirmed that the whole range [X, X + ((n-1)*ldx + d)*4) lies inside one cudaMalloc allocation
 * of the context's own GPU; 1 = every CTA loaded its own tiles (any other kind of pointer).
 * Diagnostic only; no counterpart in the reference. */
int dml_last_cluster(dml_ctx* ctx);

/* Round 0 of the last logistic sweep.  Every fit starts at w = 0
 * (sklearn/linear_model/_logistic.py:439,457), where z = 0 for every row and the
 * loss / gradient sums depend on the row mask only, so the sweep evaluates the
 * n_folds (+1) masks of one candidate and copies the sums to the others.
 * Returns the contraction milliseconds and the column count of that round
 * (both included in dml_last_timing / dml_last_counters); 0, 0 when the round
 * ran in full (one candidate, or DML_FIRST_ROUND_ONCE=0). */
int dml_last_first_round(dml_ctx* ctx, float* contraction_ms, int64_t* column_rounds);

#ifdef __cplusplus
}
#endif
#endif /* DML_B200_H */

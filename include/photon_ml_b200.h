/*
 * photon_ml_b200.h -- C ABI of the B200-native GLM objective-evaluation path.
 *
 * This is the drop-in boundary for Photon ML's DistributedGLMLossFunction /
 * SingleNodeGLMLossFunction (value, gradient, Hessian-vector, Hessian-diagonal).  The reference
 * has no FFI of its own (it is 100 % JVM); each entry point below names the Scala member whose
 * body it replaces.  Paths: lib/ = photon-lib/src/main/scala/com/linkedin/photon/ml/,
 * api/ = photon-api/src/main/scala/com/linkedin/photon/ml/.  The JNI stubs that bind these
 * symbols from Scala are in INTEGRATION.md and jni/.
 *
 * Conventions
 *  - every function returns 0 (PML_OK) or a PML_E* code; pml_last_error() gives the message of
 *    the calling thread's last failure.  PML_EINVAL corresponds to the reference's `require`
 *    failures (IllegalArgumentException); NaN/Inf in data are not errors and propagate.
 *  - plain pointers and sizes only.  "host" pointers are ordinary CPU memory owned by the caller
 *    and never retained; "dev" pointers are CUDA device memory on the context's device.
 *  - one context = one GPU = one stream.  Calls on one context are serialised by the library
 *    (internal mutex); different contexts may be used from different threads concurrently.
 *  - host-pointer entry points are synchronous (results are in the caller's buffers on return).
 *    *_dev entry points enqueue on the context's stream and return; use pml_ctx_synchronize().
 *  - there is no CPU fallback: without a CUDA device pml_ctx_create fails with PML_ECUDA.
 */
#ifndef PHOTON_ML_B200_H
#define PHOTON_ML_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PML_OK 0
#define PML_EINVAL 1 /* a `require` of the reference failed: sizes, intercept factor/shift, ... */
#define PML_ECUDA 2  /* CUDA runtime / launch failure, or no device */
#define PML_ENOMEM 3 /* device or host allocation failed */
#define PML_ESTATE 4 /* call sequence error (e.g. cached Hessian-vector without a prior pass) */

/* PointwiseLossFunction family (lib/function/glm/PointwiseLossFunction.scala:36-54). */
#define PML_LOSS_LOGISTIC 0       /* api/function/glm/LogisticLossFunction.scala:70-89 */
#define PML_LOSS_POISSON 1        /* api/function/glm/PoissonLossFunction.scala:40-52 */
#define PML_LOSS_SQUARED 2        /* api/function/glm/SquaredLossFunction.scala:42-54 */
#define PML_LOSS_SMOOTHED_HINGE 3 /* api/function/svm/SmoothedHingeLossFunction.scala:40-84 */

/* element type of the `values` array handed to pml_dataset_create_* */
#define PML_F32 0
#define PML_F64 1

/* how feature values are stored in HBM */
#define PML_STORE_F32 0  /* 4 B/value; PML_EINVAL if a value is not exactly representable */
#define PML_STORE_F64 1  /* 8 B/value, bit-exact for any Double the reference could hold */
#define PML_STORE_AUTO 2 /* ONES when every stored value is 1.0, else F32 when every value survives the round trip, else F64 */
#define PML_STORE_ONES 3 /* 0 B/value: indicator features, every stored value is exactly 1.0 (PML_EINVAL otherwise) */
/* OR-ed into `storage`: do not renumber columns by frequency -- for data whose column order already
 * carries locality (the block-diagonal batches of per-entity problems use it internally) */
#define PML_KEEP_COLUMN_ORDER_FLAG 0x100

typedef struct pml_ctx pml_ctx;
typedef struct pml_dataset pml_dataset;

typedef struct pml_dataset_info {
    int64_t n;             /* rows */
    int64_t nnz;           /* stored entries */
    int32_t d;             /* feature dimension (incl. intercept column if any) */
    int32_t value_bytes;   /* 4 or 8: what PML_STORE_* resolved to */
    int32_t margin_variant; /* CSR margin kernel of single-vector passes: lanes per row; 0 = segment-per-warp; 1 = entry-chunked */
    int32_t n_chunks;      /* CSC chunks of the transpose kernel */
    int64_t nonempty_cols; /* columns with at least one entry */
    int64_t device_bytes;  /* HBM held by this data set */
    int64_t vg_algorithmic_bytes; /* bytes one value+gradient evaluation must move */
    int64_t hv_algorithmic_bytes; /* same for a cold Hessian-vector product */
    int32_t tiled_hot_cols; /* columns of the CSC copy tiled by row blocks (0 = plain layout) */
    int32_t tiled_tiles;    /* row tiles of that layout */
    int32_t tiled_tile_rows;
    int32_t sliced_ell_padding_permille; /* the rows' sliced-ELL copy that single-vector margin passes run on: padding
                                          * entries per 1000 stored entries; -1 = not built (batches; rows too uneven) */
} pml_dataset_info;

/* message of the calling thread's last failing call ("" if none) */
const char *pml_last_error(void);

/* ---- context: one GPU, one stream ------------------------------------------------------- */
/* Replaces the SparkContext / treeAggregateDepth pair a DistributedObjectiveFunction is built
 * with (api/function/DistributedObjectiveFunction.scala:34-39). */
int pml_ctx_create(int device, pml_ctx **out);
int pml_ctx_destroy(pml_ctx *ctx);
int pml_ctx_synchronize(pml_ctx *ctx);
/* the cudaStream_t the context launches on, for callers that order their own work after it */
int pml_ctx_stream(pml_ctx *ctx, void **out_stream);
/* pinned host buffers (what a JNI shim backs its direct ByteBuffers with) */
int pml_host_alloc(void **out, int64_t bytes);
int pml_host_free(void *p);

/* ---- data set: LabeledPoint rows packed once into device CSR + CSC ------------------------ */
/* Replaces RDD[LabeledPoint] / Iterable[LabeledPoint] as `Data`
 * (api/function/DistributedObjectiveFunction.scala:36, SingleNodeObjectiveFunction.scala:27;
 * row type lib/data/LabeledPoint.scala:32-36).  Rows need not be sorted by column and may
 * repeat a column (Breeze SparseVector allows both); offsets NULL = 0, weights NULL = 1
 * (lib/constants/MathConst.scala:24-25).  row_ptr has n+1 entries, row_ptr[0] == 0. */
int pml_dataset_create_csr(pml_ctx *ctx, int64_t n, int32_t d, const int64_t *row_ptr,
                           const int32_t *col_idx, const void *values, int values_dtype,
                           int storage, const double *labels, const double *offsets,
                           const double *weights, pml_dataset **out);
/* same, all pointers in device memory (the arrays are copied; the caller keeps ownership) */
int pml_dataset_create_csr_dev(pml_ctx *ctx, int64_t n, int32_t d, const int64_t *row_ptr_dev,
                               const int32_t *col_idx_dev, const void *values_dev,
                               int values_dtype, int storage, const double *labels_dev,
                               const double *offsets_dev, const double *weights_dev,
                               pml_dataset **out);
int pml_dataset_destroy(pml_dataset *ds);
int pml_dataset_get_info(const pml_dataset *ds, pml_dataset_info *out);

/* Coordinate.updateModel re-offsets the data every coordinate-descent pass
 * (lib/algorithm/Coordinate.scala:59-62, api/data/FixedEffectDataSet.scala:45-54): only the
 * n offsets are uploaded, the packed matrix stays.  NULL restores 0 / 1. */
int pml_dataset_set_offsets(pml_dataset *ds, const double *offsets);
/* down-sampling re-weights rows (api/optimization/DistributedOptimizationProblem.scala:145-160) */
int pml_dataset_set_weights(pml_dataset *ds, const double *weights);
int pml_dataset_set_labels(pml_dataset *ds, const double *labels);
/* NormalizationContext(factors, shifts, interceptId) (lib/normalization/
 * NormalizationContext.scala:39-47); NULL = None, intercept_id -1 = None.  Fails with
 * PML_EINVAL exactly where the reference's requires do: shifts without intercept (:44),
 * factors[intercept] != 1 or shifts[intercept] != 0
 * (lib/function/glm/ValueAndGradientAggregator.scala:96-99,109-112). */
int pml_dataset_set_normalization(pml_dataset *ds, const double *factors, const double *shifts,
                                  int32_t intercept_id);

/* ---- evaluations, host buffers (the drop-in calls) ---------------------------------------- */
/* DistributedGLMLossFunction.calculate (api/function/glm/DistributedGLMLossFunction.scala:93-102)
 * and SingleNodeGLMLossFunction.calculate (SingleNodeGLMLossFunction.scala:83-91), i.e.
 * ValueAndGradientAggregator.calculateValueAndGradient (lib/function/glm/
 * ValueAndGradientAggregator.scala:235-274).  With PML_LOSS_SMOOTHED_HINGE it is
 * DistributedSmoothedHingeLossFunction.calculate (api/function/svm/
 * DistributedSmoothedHingeLossFunction.scala:84-103): normalization is ignored, as there.
 * coef and out_grad have d entries; the sums are unregularised (L2 stays on the caller's side,
 * lib/function/L2Regularization.scala:115-125).  out_grad may be NULL (value only:
 * DistributedGLMLossFunction.value, :64-68). */
int pml_value_and_gradient(pml_dataset *ds, int loss, const double *coef, double *out_value,
                           double *out_grad);
/* DistributedGLMLossFunction.hessianVector (:114-125) = HessianVectorAggregator.calcHessianVector
 * (lib/function/glm/HessianVectorAggregator.scala:136-178).  coef_tag: 0 = stateless (as the
 * reference: the margin pass is redone); non-zero = identity of the coefficient vector (the JNI
 * shim passes the Broadcast id).  When the tag equals the previous call's, coef is not read
 * again and the cached per-row wt*l''(z) is reused -- TRON's CG loop calls hessianVector up to
 * 20 times at fixed coefficients (lib/optimization/TRON.scala:294-301). */
int pml_hessian_vector(pml_dataset *ds, int loss, const double *coef, uint64_t coef_tag,
                       const double *v, double *out_hv);
/* DistributedGLMLossFunction.hessianDiagonal (:134-137) = HessianDiagonalAggregator
 * (lib/function/glm/HessianDiagonalAggregator.scala:51-61,89-125); ignores normalization. */
int pml_hessian_diagonal(pml_dataset *ds, int loss, const double *coef, double *out_diag);
/* Coordinate scoring, score_i = x_i . coef (api/algorithm/FixedEffectCoordinate.scala:158-165,
 * lib/model/Coefficients.scala:53-58): no offset, no normalization.  out_scores has n entries. */
int pml_scores(pml_dataset *ds, const double *coef, double *out_scores);

/* ---- evaluations, device buffers (resident pipelines and the multi-GPU split) -------------- */
/* Raw partial sums of this data set's rows: out_raw_dev[0..d) = vectorSum, [d] = valueSum,
 * [d+1] = vectorShiftPrefactorSum -- the fields ValueAndGradientAggregator.merge adds up
 * (lib/function/glm/ValueAndGradientAggregator.scala:161-176).  With rows sharded over several
 * GPUs the caller sums these d+2 doubles across ranks (one all-reduce; it replaces Spark's
 * treeAggregate, :243-247) and then calls pml_finalize_dev once. */
int pml_value_and_gradient_raw_dev(pml_dataset *ds, int loss, const double *coef_dev,
                                   double *out_raw_dev);
int pml_hessian_vector_raw_dev(pml_dataset *ds, int loss, const double *coef_dev,
                               uint64_t coef_tag, const double *v_dev, double *out_raw_dev);
/* ValueAndGradientAggregator.getVector (:199-221): out_j = (raw_j - shift_j * raw[d+1]) *
 * factor_j; *out_value_dev = raw[d] (may be NULL).  `loss` only matters for the smoothed hinge
 * (normalization ignored). */
int pml_finalize_dev(pml_dataset *ds, int loss, const double *raw_dev, double *out_value_dev,
                     double *out_vec_dev);
/* single-GPU convenience: raw + finalize, everything on the device */
int pml_value_and_gradient_dev(pml_dataset *ds, int loss, const double *coef_dev,
                               double *out_value_dev, double *out_grad_dev);
int pml_hessian_vector_dev(pml_dataset *ds, int loss, const double *coef_dev, uint64_t coef_tag,
                           const double *v_dev, double *out_hv_dev);

/* ---- coordinate-descent glue kept on the device ---------------------------------------------- */
/* score_i = x_i . coef for every row, left in device memory
 * (api/algorithm/FixedEffectCoordinate.scala:158-165, lib/model/Coefficients.scala:53-58). */
int pml_scores_dev(pml_dataset *ds, const double *coef_dev, double *out_scores_dev);
/* pml_dataset_set_offsets from a device array (NULL = no offsets). */
int pml_dataset_set_offsets_dev(pml_dataset *ds, const double *offsets_dev);
/* out_i = base_i + scores[row_map[i]]: DataSet.addScoresToOffsets (api/data/FixedEffectDataSet.scala:
 * 45-54, api/data/LocalDataSet.scala:82-89) without the RDD join.  row_map_dev translates this
 * coordinate's row order into the other's (NULL = same order); base_offsets_dev NULL = 0. */
int pml_offsets_from_scores_dev(pml_ctx *ctx, int64_t n, const double *base_offsets_dev,
                                const double *scores_dev, const int64_t *row_map_dev,
                                double *out_offsets_dev);
/* Training-loss evaluator on given scores: *out_value_dev = sum_i weights[i] * l(scores[i] + offsets[i], labels[i])
 * (Evaluator.evaluate over (score + offset, label, weight): lib/evaluation/Evaluator.scala:35-45,
 * api/evaluation/LogisticLossEvaluator.scala:39-46 and its Poisson / squared / smoothed-hinge siblings), for scores
 * that are already on the device (the sum of the coordinates' pml_scores_dev).  Asynchronous on the stream. */
int pml_loss_from_scores_dev(pml_dataset *ds, int loss, const double *scores_dev, double *out_value_dev);

/* number of kernel launches this context has issued (bench.py's gpu_launches) */
int pml_ctx_launch_count(pml_ctx *ctx, int64_t *out);
/* Per-kernel device timing for the roofline report: with profiling on, every kernel class is
 * bracketed by CUDA events on the context's stream; pml_ctx_read_profile synchronises, sums the
 * elapsed milliseconds and launch counts per class since the last read, and clears them. */
#define PML_K_PREPARE 0  /* prepare_vectors_kernel */
#define PML_K_ROW 1      /* CSR product: row_pass_kernel, or seg_spmv_kernel for variant 1 */
#define PML_K_COL 2      /* CSC product: seg_spmv_kernel */
#define PML_K_FIXUP 3    /* fixup_kernel */
#define PML_K_FINALIZE 4 /* finalize_kernel */
#define PML_K_EPILOGUE 5 /* row_epilogue_kernel (row-pass variant 1 only) */
#define PML_K_ENTITY 7   /* entity_values_kernel (per-entity objective values of a batch) */
#define PML_K_COL_TILED 6 /* seg_spmv_tiled_kernel: CSC pass over the hot columns, r staged in shared memory */
#define PML_K_PEER 8     /* peer_merge_kernel: merge over NVLink peer memory + getVector */
#define PML_K_COUNT 9
int pml_ctx_set_profiling(pml_ctx *ctx, int on);
int pml_ctx_read_profile(pml_ctx *ctx, double *out_ms /* [PML_K_COUNT] */,
                         int64_t *out_launches /* [PML_K_COUNT] */);
/* force a CSR margin-kernel variant for every pass: lanes per row in {2,4,8,16,32} (32 =
 * warp-per-row), 0 = segment-per-warp (32-row blocks), 1 = entry-chunked (256-entry warp chunks,
 * load-balanced), -1 = back to the default choice (1 for single-vector passes, lanes per row
 * chosen from nnz/row for the paired pass of a cold Hessian-vector product).  For benchmarking
 * and tests. */
int pml_dataset_set_margin_variant(pml_dataset *ds, int lanes_per_row);

/* ---- batched per-entity problems (random effects) ------------------------------------------ */
/* Replaces the body of the per-entity mapValues in RandomEffectCoordinate.updateModel
 * (api/algorithm/RandomEffectCoordinate.scala:108-119), where each entity's
 * SingleNodeGLMLossFunction.calculate (api/function/glm/SingleNodeGLMLossFunction.scala:83-91)
 * runs over its own LocalDataSet (api/data/LocalDataSet.scala:36): all E local problems are
 * evaluated together.  Rows are concatenated entity by entity (entity e owns rows
 * [ent_row_ptr[e], ent_row_ptr[e+1])); col_local holds each row's indices in the entity's own
 * (projected) space [0, ent_dim[e]).  Coefficients, gradients and multiply vectors of all
 * entities are concatenated in entity order (entity e at [coef_ptr[e], coef_ptr[e+1]),
 * coef_ptr = prefix sum of ent_dim).  Entities are independent, so with several GPUs each rank
 * packs its own subset and no collective is needed. */
typedef struct pml_batch pml_batch;
int pml_batch_create_csr(pml_ctx *ctx, int64_t n_entities, const int64_t *ent_row_ptr,
                         const int32_t *ent_dim, const int64_t *row_ptr, const int32_t *col_local,
                         const void *values, int values_dtype, int storage, const double *labels,
                         const double *offsets, const double *weights, pml_batch **out);
int pml_batch_destroy(pml_batch *b);
/* The block-diagonal data set behind the batch: pml_dataset_set_offsets / set_weights (rows in
 * concatenated order), pml_hessian_vector, pml_hessian_diagonal and pml_scores apply to all
 * entities at once through it.  Owned by the batch; do not destroy it. */
int pml_batch_dataset(pml_batch *b, pml_dataset **out);
int pml_batch_coef_ptr(pml_batch *b, int64_t *out /* E+1 */);
/* out_values[e] = entity e's objective value; out_grads (may be NULL) = concatenated gradients */
int pml_batch_value_and_gradient(pml_batch *b, int loss, const double *coefs, double *out_values,
                                 double *out_grads);
/* same with device buffers (asynchronous on the context's stream): what a device-resident batched
 * optimizer calls every iteration -- no per-call transfer of the concatenated coefficient vector */
int pml_batch_value_and_gradient_dev(pml_batch *b, int loss, const double *coefs_dev,
                                     double *out_values_dev, double *out_grads_dev);

/* same for the Hessian-vector product (SingleNodeGLMLossFunction.hessianVector :103-113, what a batched per-entity
 * TRON calls in its conjugate-gradient loop); coef_tag as in pml_hessian_vector. */
int pml_batch_hessian_vector(pml_batch *b, int loss, const double *coefs, uint64_t coef_tag, const double *vs,
                             double *out_hvs);
int pml_batch_hessian_vector_dev(pml_batch *b, int loss, const double *coefs_dev, uint64_t coef_tag,
                                 const double *vs_dev, double *out_hvs_dev);
/* residual offsets of all rows in concatenated order (LocalDataSet.addScoresToOffsets, api/data/LocalDataSet.scala:82-89) */
int pml_batch_set_offsets(pml_batch *b, const double *offsets);
/* active[e] == 0 takes entity e out of the following evaluations: its value, gradient and Hessian-vector entries
 * come back as 0 (a lock-step batched optimizer uses this for entities that have converged: in the reference such an
 * entity's solve has simply ended, RandomEffectCoordinate.scala:112-119), and the parts of the data that hold only
 * masked entities' rows are no longer streamed.  Scores (pml_scores* on pml_batch_dataset) are still computed for every
 * row.  With coefficient tags, entities may only LEAVE the active set between Hessian-vector calls that share a tag
 * (the cached curvature of rows skipped by the first call is not there).  NULL = all active. */
int pml_batch_set_active(pml_batch *b, const uint8_t *active /* E, host */);

/* Per-entity NormalizationContext (lib/normalization/NormalizationContext.scala:39-47; the random-effect problems
 * train in the transformed space of their projected features exactly as the fixed effect does:
 * ValueAndGradientAggregator.scala:90-122, 199-221, HessianVectorAggregator.scala:60-85 per entity).  factors / shifts
 * are concatenated like the coefficients (either may be NULL); intercept_local[e] is entity e's intercept index in its
 * own space, -1 = none (NULL = none anywhere).  The reference's requires hold per entity (shifts need an intercept;
 * its factor is 1 and its shift 0) -> PML_EINVAL.  With shifts every row of an entity must also hold the entry
 * (intercept, 1.0) -- the library folds the per-entity shift terms into the intercept coefficient -- else PML_EINVAL.
 * All NULL removes the context.  The smoothed hinge loss ignores it, as in the reference. */
int pml_batch_set_normalization(pml_batch *b, const double *factors, const double *shifts, const int32_t *intercept_local);
/* same with the mask already in device memory (a device-resident optimizer updates it without a host round trip);
 * the bytes are copied.  NULL = all active. */
int pml_batch_set_active_dev(pml_batch *b, const uint8_t *active_dev /* E, device */);
/* Vector algebra of a lock-step batched optimizer over the concatenated per-entity vectors -- what Breeze's dot and
 * axpy do for ONE entity inside TRON.truncatedCG / LBFGS (lib/optimization/TRON.scala:279-339), for all at once:
 *   segment_dot:   out[e] = sum over entity e's coefficients of x_j * y_j            (fixed order per entity)
 *   segment_axpby: out_j  = a[e(j)] * x_j + c[e(j)] * y_j   (a / c NULL = 1, y NULL = no second term; out may alias
 *                  x or y; entities with active_dev[e] == 0 keep out_j, NULL = all)
 * Device pointers, asynchronous on the context's stream. */
int pml_batch_segment_dot_dev(pml_batch *b, const double *x_dev, const double *y_dev, double *out_dev /* E */);
/* One iteration of TRON.truncatedCG (lib/optimization/TRON.scala:300-336) for every entity with run[e] != 0, given
 * hd = (unregularised) Hessian times `direction` from pml_batch_hessian_vector_dev: alpha = rTr / (d . Hd); step += alpha d;
 * outside the trust region delta[e]: back to the boundary, residual updated, out_over[e] = 1 (the entity's loop ends);
 * inside: residual -= alpha Hd, beta = r'.r' / rTr, direction = beta direction + residual, rtr[e] = r'.r', out_over[e] = 0.
 * l2 adds the L2 mix-in's term (Hd += l2 * direction).  All vectors are the entities' vectors concatenated (pml_batch
 * layout), rtr / delta / run / out_over have one slot per entity; everything in device memory, asynchronous. */
int pml_batch_cg_step_dev(pml_batch *b, const uint8_t *run_dev, double l2, double *direction_dev, const double *hd_dev,
                          double *step_dev, double *residual_dev, double *rtr_dev, const double *delta_dev,
                          uint8_t *out_over_dev);
/* The L-BFGS direction for every entity with go[e] != 0 (NULL = all) in one kernel: Breeze's two-loop recursion, which
 * lib/optimization/LBFGS.scala:59-107 runs for ONE problem through breeze.optimize.LBFGS (numCorrections pairs,
 * LBFGS.scala:44, 69-76) --  q = -grad; newest pair to oldest: a_i = rho_i (s_i . q), q -= a_i y_i; q *= s.y / y.y of the
 * newest pair; oldest to newest: q += (a_i - rho_i (y_i . q)) s_i.  S / Y hold the correction pairs as m concatenated
 * vectors each (slot-major, m x D doubles), rho their 1 / (y . s) as m x E; every entity keeps its own ring: head[e] is
 * the slot of its newest pair, len[e] <= m how many it holds (0: the direction is -grad).  Entities that are not worked
 * on keep out[] as it is.  1 <= m <= 16.  pml_batch_lbfgs_push_dev appends (s, y, rho_new) for every entity with
 * mask[e] != 0: head advances (a dropped history, len = 0, starts again at slot 0), len grows up to m.
 * Everything in device memory, asynchronous on the context's stream. */
int pml_batch_lbfgs_direction_dev(pml_batch *b, const uint8_t *go_dev /* E or NULL */, int m, const double *S_dev,
                                  const double *Y_dev, const double *rho_dev, const int32_t *head_dev /* E */,
                                  const int32_t *len_dev /* E */, const double *grad_dev, double *out_dev);
int pml_batch_lbfgs_push_dev(pml_batch *b, const uint8_t *mask_dev /* E */, int m, const double *rho_new_dev /* E */,
                             const double *s_dev, const double *y_dev, double *S_dev, double *Y_dev, double *rho_dev,
                             int32_t *head_dev, int32_t *len_dev);
int pml_batch_segment_axpby_dev(pml_batch *b, const double *a_dev /* E */, const double *x_dev, const double *c_dev /* E */,
                                const double *y_dev, const uint8_t *active_dev /* E */, double *out_dev);

/* ---- the merge step across GPUs over peer memory ------------------------------------------------ */
/* One process per GPU.  Replaces treeAggregate's merge (ValueAndGradientAggregator.scala:161-176, :243-247;
 * HessianVectorAggregator.scala:136-152) AND getVector (:199-221) by one kernel per rank that adds up the ranks'
 * raw sums over NVLink peer memory, in rank order, and leaves the finished value / vector on every rank
 * (bit-identical everywhere).  Set-up: every rank creates a pml_peer, publishes pml_peer_handle() to all ranks
 * by any host-side means (the Python binding uses torch.distributed.all_gather_object; a Spark driver would collect
 * and broadcast them), and passes the rank-ordered concatenation to pml_peer_connect().  Every rank must then
 * make the same sequence of *_allreduce_dev calls; normalization factors / shifts must be identical on all ranks.
 * A rank that never arrives makes the others give up after a few seconds with PML_ECUDA on their next call. */
#define PML_MAX_PEERS 8
#define PML_PEER_HANDLE_BYTES 64
typedef struct pml_peer pml_peer;
int pml_peer_create(pml_ctx *ctx, int rank, int world, int32_t d, pml_peer **out);
int pml_peer_handle(pml_peer *peer, void *out_handle /* PML_PEER_HANDLE_BYTES */);
int pml_peer_connect(pml_peer *peer, const void *all_handles /* world * PML_PEER_HANDLE_BYTES, rank order */);
int pml_peer_destroy(pml_peer *peer);
/* Optional second transport: an NVSwitch multicast object (NVLS) over the ranks' slabs -- the switch adds
 * (multimem.ld_reduce) and fans out (multimem.st), so a rank moves 1/N of the vector instead of (N-1)/N each way.
 * Rank 0 calls _export and hands the returned POSIX file descriptor to the other ranks (SCM_RIGHTS over a Unix
 * socket; the caller closes its descriptors afterwards); every rank calls _join (rank 0 with fd = -1), the ranks
 * agree on the host that all joins succeeded, then every rank calls _bind (which blocks until all have joined).
 * From then on *_allreduce_dev uses the multicast kernel and pml_peer_buffers returns the multicast-bound
 * buffers.  Any failure (no NVSwitch fabric, no permission) leaves the peer-to-peer transport in place. */
/* which physical GPU a context is on ("0000:1b:00.0"): ranks that share a GPU cannot form a multicast team */
int pml_ctx_pci_bus_id(pml_ctx *ctx, char *out, int len /* >= 16 */);
#define PML_PEER_P2P 0
#define PML_PEER_MULTICAST 1
int pml_peer_multicast_export(pml_peer *peer, int *out_fd);
int pml_peer_multicast_join(pml_peer *peer, int fd);
int pml_peer_multicast_bind(pml_peer *peer);
int pml_peer_multicast_release(pml_peer *peer); /* back to the peer-to-peer transport (e.g. another rank failed to bind) */
int pml_peer_transport(pml_peer *peer, int *out /* PML_PEER_P2P or PML_PEER_MULTICAST */);
/* The rank's exchange buffers (device memory owned by the peer handle): raw = d+2 doubles the evaluation fills,
 * result = d+1 doubles, [vector(d), value], valid after an *_allreduce_dev call in stream order. */
int pml_peer_buffers(pml_peer *peer, double **raw_dev, double **result_dev);
/* DistributedGLMLossFunction.calculate over all ranks' rows: this rank's raw sums, the merge, getVector.
 * Asynchronous on the context's stream.  out_* may be NULL (read the result buffer instead: no copy). */
int pml_value_and_gradient_allreduce_dev(pml_dataset *ds, pml_peer *peer, int loss, const double *coef_dev,
                                         double *out_value_dev, double *out_grad_dev);
int pml_hessian_vector_allreduce_dev(pml_dataset *ds, pml_peer *peer, int loss, const double *coef_dev,
                                     uint64_t coef_tag, const double *v_dev, double *out_hv_dev);
/* The merge + getVector alone, for callers that filled the raw buffer themselves (pml_*_raw_dev with
 * out_raw_dev = the peer's raw buffer): same as pml_finalize_dev after an all-reduce of the raw sums. */
int pml_peer_merge_dev(pml_dataset *ds, pml_peer *peer, int loss);
/* The same two exchanges with HOST vectors in and out -- the shape DistributedObjectiveFunction hands over
 * (coefficients broadcast to every executor, DistributedObjectiveFunction.scala:57-58; result to the driver, :73;
 * bodies DistributedGLMLossFunction.scala:93-102, 114-125).  Every rank passes the SAME coefficient vector (d
 * doubles) and uploads only its own slice [lo, hi) of it (pml_peer_slice) over its own PCIe link; the slices are
 * exchanged between the GPUs over NVLink, then evaluation, merge and getVector run as in the _dev forms.
 * gather_output = 0: only out[lo, hi) is written -- ranks that pass one shared host buffer (a memory-mapped file,
 * a buffer of the driver JVM registered by every rank) assemble the full vector in it with N parallel downloads;
 * gather_output = 1: the whole vector is downloaded on this rank.  *out_value is set on every rank.  Synchronous.
 * coef may be NULL for a Hessian-vector product whose tag matches the cached pass. */
int pml_peer_slice(pml_peer *peer, int64_t *out_lo, int64_t *out_hi);
int pml_value_and_gradient_allreduce(pml_dataset *ds, pml_peer *peer, int loss, const double *coef, double *out_value,
                                     double *out_grad, int gather_output);
int pml_hessian_vector_allreduce(pml_dataset *ds, pml_peer *peer, int loss, const double *coef, uint64_t coef_tag,
                                 const double *v, double *out_hv, int gather_output);
/* DistributedGLMLossFunction.hessianDiagonal (DistributedGLMLossFunction.scala:134-137, HessianDiagonalAggregator.scala:
 * 89-125) over all ranks' rows; like the reference it ignores the normalization context. */
int pml_hessian_diagonal_allreduce(pml_dataset *ds, pml_peer *peer, int loss, const double *coef, double *out_diag,
                                   int gather_output);
/* 1 when an exchange on this handle gave up waiting for another rank (its result was filled with NaN);
 * pml_ctx_synchronize and every later *_allreduce* call on the handle then return PML_ECUDA. */
int pml_peer_status(pml_peer *peer, int *out);
/* The whole exchange of `world` ranks on ONE device, every rank's CTAs inside one cooperative launch (no reference
 * counterpart: a check of the exchange kernels for boxes with fewer GPUs than ranks -- kernels that wait on one
 * another must not run as separate launches on one GPU).  raw: world x (d+2) host doubles, rank q's raw sums
 * [vectorSum(d), valueSum, prefactorSum]; out_result: world x (d+1), what each rank would hold after the merge +
 * getVector (factors / shifts: d host doubles or NULL).  vec_in / vec_out (both or neither): world x d; rank q
 * uploads its slice of vec_in[q], the all-gather runs, and vec_out[q] is rank q's assembled copy.  The exchange is
 * repeated `repeats` times on the same slabs (the flag epochs advance). */
int pml_peer_emulate(pml_ctx *ctx, int world, int32_t d, int repeats, const double *raw, const double *factors,
                     const double *shifts, double *out_result, const double *vec_in, double *vec_out);

/* ---- evaluation without any allocation ------------------------------------------------------ */
/* No counterpart in the reference (the JVM allocates freely); this exists for callers of THIS
 * library that must not trigger cudaMalloc/cudaFree: the JNI shim while it pins Java arrays with
 * GetPrimitiveArrayCritical (the JVM may be holding off garbage collection for all threads, and
 * cudaMalloc synchronises the device and can block), and executor threads that evaluate many
 * small per-entity problems side by side on one GPU (an allocation in one stalls the kernels of
 * all).  Such callers keep their rows in device memory themselves and cannot create a
 * pml_dataset, because packing allocates.
 *
 * Computes *out_value_dev = sum_i weights[i] * l(x_i . coef + offsets[i], labels[i]) -- what
 * DistributedGLMLossFunction.value returns when no normalization is set -- directly from the
 * caller's CSR arrays (row_ptr_dev: n+1 int64 starting at 0; col_idx_dev in [0, d), not
 * checked; values_dev float or double per values_dtype; offsets_dev / weights_dev may be NULL =
 * 0 / 1).  All pointers are device memory on `device`; the work is enqueued on `stream`
 * (a cudaStream_t, NULL = default stream) and the call returns without synchronising.
 *
 * It calls no cudaMalloc, cudaFree or allocating library routine.  Its scratch (one partial sum
 * per 256 rows plus a completion counter, 512 KB) is a fixed array inside the library image,
 * resident on every GPU that loads the library for the life of the process.  Consequences:
 *  - n is limited to 16,777,216 rows (65,536 CTAs x 256 rows); larger n fails with PML_EINVAL.
 *  - all calls on one device share that array: two calls whose kernels overlap on the GPU
 *    (different streams, or different threads) race and give wrong sums.  Issue them on one
 *    stream, or otherwise order them.  The array keeps the previous call's contents; the call
 *    clears the counter itself, nothing needs resetting by the caller.
 *  - the first call on each device also looks up the array's device address
 *    (cudaGetSymbolAddress), which may load the library's device code on that GPU -- the one
 *    moment the CUDA runtime itself may allocate and synchronise.  Make that first call once per
 *    device at start-up: before starting any stream capture, and OUTSIDE any
 *    GetPrimitiveArrayCritical region or other section in which allocation is forbidden.
 *  - the calling thread's current CUDA device is the same after the call as before it. */
int pml_value_noalloc_dev(int device, void *stream, int64_t n, int32_t d,
                          const int64_t *row_ptr_dev, const int32_t *col_idx_dev,
                          const void *values_dev, int values_dtype, const double *labels_dev,
                          const double *offsets_dev, const double *weights_dev, int loss,
                          const double *coef_dev, double *out_value_dev);

#ifdef __cplusplus
}
#endif
#endif /* PHOTON_ML_B200_H */

/*
 * b200fit.h -- C ABI of libb200fit.so, the B200 (sm_100a) fit engine that sits
 * under the pyspark.ml-shaped Python surface in orange3-spark_b200/b200fit/.
 *
 * The reference (Gjerman/Orange3-Spark) has NO native / FFI interface: its hot
 * path is `pyspark.ml` objects reached over py4j.  Each entry point below
 * therefore cites the reference call site whose (JVM-side) work it replaces,
 * and the MLlib operation restated in SURVEY.md Appendix A.
 *
 * Conventions
 *   - every function returns 0 on success, a negative B2F_E* code otherwise;
 *     b2f_last_error() returns a thread-local message for the last failure;
 *   - all pointers are DEVICE pointers unless the name ends in _host;
 *   - the caller owns every buffer (the library never allocates device memory; its development
 *     tracers, which did, are compiled in only with -DB2F_DEV)
 *     and passes the CUDA stream as an opaque pointer (cudaStream_t; NULL = the
 *     legacy default stream);
 *   - calls are asynchronous on `stream`; nothing here synchronises the device;
 *   - feature matrices are row-major fp32 `X[n, ld]`, ld >= d, ld % 4 == 0 and X
 *     16-byte aligned (rows are legal float4 / TMA rows).
 */
#ifndef B200FIT_H_
#define B200FIT_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2F_OK 0
#define B2F_EINVAL (-1)   /* bad argument (message says which) */
#define B2F_ECUDA (-2)    /* CUDA runtime / driver error */
#define B2F_EUNSUPPORTED (-3)

/* column dtypes accepted by b2f_pack_columns */
#define B2F_F32 0
#define B2F_F64 1
#define B2F_I32 2
#define B2F_I64 3
#define B2F_U8 4 /* also bool */

/* b2f_kmeans_step algorithms */
#define B2F_KM_AUTO 0
#define B2F_KM_GENERIC 1 /* CUDA-core fp32, any k / d */
#define B2F_KM_TC3 2     /* tcgen05 3xTF32 (fp32-accurate) contraction, d % 32 == 0, k % 128 == 0 */
#define B2F_KM_TC1 3     /* tcgen05 1xTF32 contraction (filter pass; labels need refine) */
#define B2F_KM_TCB 4     /* tcgen05 bf16x3 filter pass + 3xTF32 re-check of uncertain rows; d in {64,128},
                            k % 128 == 0, k <= 65536; only through b2f_kmeans_step_ws (labels + workspace) */
#define B2F_KM_TCH 5     /* same pipeline with a single-MMA fp16 filter pass (a third of the tensor work, a
                            wider uncertainty band); same shapes and entry point as B2F_KM_TCB           */

typedef void* b2f_stream_t;

/* ---- library ------------------------------------------------------------ */
int b2f_version(void);
const char* b2f_last_error(void);

/* ---- synthetic inputs (Philox4x32-10 keyed on (seed, global row, column)) ----
 * No reference counterpart: the reference ships no data generator; these produce the
 * BASELINE.json config shapes on device so 100M x 128 never crosses PCIe.          */
/* X[r, c] = centre[comp(row)][c] * 1 + z(row, c);  centre[j][c] = center_scale * z'(j, c);
 * comp(row) = philox(row) mod n_components; optional per-column scale multiplies the result. */
int b2f_synth_mixture(float* X, int64_t n, int64_t row0, int d, int64_t ld, uint64_t seed,
                      int n_components, float center_scale, const float* col_scale, b2f_stream_t stream);
/* X[r, c] = col_scale[c] * z(row, c);  margins m_k = X[r,:] . w_true[k,:] + b_true[k];
 * n_classes == 2: y = 1 if u < sigmoid(m_0) else 0 (w_true is [1,d]);
 * n_classes  > 2: y ~ Categorical(softmax(m)) by inverse CDF on u.  y is written as fp32 class ids. */
int b2f_synth_logistic(float* X, float* y, int64_t n, int64_t row0, int d, int64_t ld, uint64_t seed,
                       const float* col_scale, const float* w_true, const float* b_true, int n_classes,
                       b2f_stream_t stream);

/* ---- K1 loader: Dataset Builder -> VectorAssembler ------------------------
 * replaces VectorAssembler(inputCols, outputCol='features').transform(df)
 *   reference: orangecontrib/spark/widgets/ml/spark_ml_dataset.py:575-576 (A.1)
 * col_ptrs_dev[d]  : device array of device pointers, one per ordered input column
 * col_dtypes_dev[d]: device array of B2F_* dtype codes
 * X[n, ld]         : fp32 row-major output (padding columns d..ld-1 are zeroed)
 * bad_flag         : device int32, set to 1 if any value is NaN (NULL to skip).  Infinities pass: MLlib's
 *                    VectorAssembler throws "Encountered NaN while assembling a row" under
 *                    handleInvalid = "error" for NaN (and null) only.                       */
int b2f_pack_columns(const void* const* col_ptrs_dev, const int32_t* col_dtypes_dev, int d, int64_t n,
                     float* X, int64_t ld, int32_t* bad_flag, b2f_stream_t stream);

/* column statistics for StandardScaler.fit / the LogisticRegression summarizer (A.3 pass 0, A.6)
 *   reference: reached from fit() orangecontrib/spark/base/spark_ml_estimator.py:22
 * out[0:d] += sum w (x - shift), out[d:2d] += sum w (x - shift)^2, out[2d] += sum w, out[2d+1] += sum w^2
 * (fp64, accumulated with atomics: zero `out` first; `w` and `shift` may be NULL).       */
int b2f_colstats(const float* X, int64_t n, int d, int64_t ld, const float* w, const float* shift,
                 double* out, b2f_stream_t stream);
/* StandardScalerModel.transform (A.6): Y[i, c] = (X[i, c] - shift[c]) * scale[c] for c < d, 0 in the padding columns;
 * scale / shift: device fp32[d], either may be NULL (1 / 0); Y may be X (in place).  One HBM pass.
 *   reference: the scaler is listed by orangecontrib/spark/utils/ml_api_utils.py:65-67 and applied through
 *   transform() at orangecontrib/spark/base/spark_ml_transformer.py:133 (BASELINE config 5)            */
int b2f_scale_columns(const float* X, float* Y, int64_t n, int d, int64_t ld, const float* scale, const float* shift,
                      b2f_stream_t stream);
/* label histogram: hist[c] += sum of w over rows with y == c, c < n_bins (fp64, zero first) */
int b2f_label_hist(const float* y, const float* w, int64_t n, int n_bins, double* hist, int32_t* bad_flag,
                   b2f_stream_t stream);

/* ---- K2 KMeans (Lloyd) ----------------------------------------------------
 * replaces the JVM work behind KMeans.fit
 *   reference: orangecontrib/spark/base/spark_ml_estimator.py:22 with module = clustering
 *   (orangecontrib/spark/widgets/ml/spark_ml_clustering.py:14); MLlib semantics: SURVEY.md A.2.
 *
 * Device state owned by the caller:
 *   centers   double[k*d]                      master copy of the centres
 *   prep      b2f_kmeans_prep_bytes(k,d) bytes  kernel-ready centre operands (opaque)
 *   acc       b2f_kmeans_acc_bytes(k,d) bytes   accumulators; its HEAD is the all-reduce
 *             payload double[k*d + k + 2] = [sums | counts | cost | n_rows]; the tail is
 *             scratch (fp32 replicas, per-CTA partials).                                     */
size_t b2f_kmeans_prep_bytes(int k, int d);
size_t b2f_kmeans_acc_bytes(int k, int d);
/* number of doubles in the all-reduce payload at the head of `acc` */
int64_t b2f_kmeans_payload_len(int k, int d);
int b2f_kmeans_prepare(const double* centers, int k, int d, void* prep, b2f_stream_t stream);
/* The same, told the largest row norm of the data the operands will meet (x_norm_max >= max_i ||x_i||, of the scaled
 * rows when a column scale is used; 0 = unknown = b2f_kmeans_prepare).  With it the fp16 filter scales its operands
 * so that every score of every row fits one 22-bit key window and lets the tensor core add the rounding constant:
 * one instruction per score less in its epilogue.  Rows whose norm exceeds the figure (other data at predict time)
 * are simply handed to the re-check.  b2f_row_norm_max computes it: out is a device float, ZEROED by the caller,
 * receives max ||x|| rounded up (atomicMax; with several ranks every rank may use its own shard's value).
 * Same reference call site as above (KMeans.fit).                                                        */
int b2f_kmeans_prepare_x(const double* centers, int k, int d, double x_norm_max, void* prep, b2f_stream_t stream);
int b2f_row_norm_max(const float* X, int64_t n, int d, int64_t ld, const float* col_scale, float* out,
                     b2f_stream_t stream);
int b2f_kmeans_acc_zero(void* acc, int k, int d, b2f_stream_t stream);
/* one assignment + accumulate pass over this rank's rows (findClosest + axpy into sums):
 * labels (int32[n]) may be NULL; w (row weights) may be NULL; algo is a B2F_KM_* code.       */
int b2f_kmeans_step(const float* X, int64_t n, int d, int64_t ld, const float* w, const void* prep, int k,
                    void* acc, int32_t* labels, int algo, b2f_stream_t stream);
/* Same pass with a caller-provided workspace (b2f_kmeans_ws_bytes(n, k, d) bytes, 256-byte aligned),
 * which enables B2F_KM_TCB / B2F_KM_TCH: rows whose label the low-precision filter pass cannot certify
 * are listed in the workspace, gathered and re-assigned by the 3xTF32 kernel.  The list holds n/5 rows; if
 * more rows are uncertain the excess is DROPPED and the payload's row count (last entry) comes out below
 * n -- the caller must then redo the pass with B2F_KM_TC3 and the same workspace (see below).  B2F_KM_AUTO picks TCH when the shape allows it
 * and labels / ws are given, else behaves like b2f_kmeans_step.  Same reference call site as above.    */
size_t b2f_kmeans_ws_bytes(int64_t n, int k, int d);
int b2f_kmeans_step_ws(const float* X, int64_t n, int d, int64_t ld, const float* w, const void* prep, int k,
                       void* acc, int32_t* labels, void* ws, size_t ws_bytes, int algo, b2f_stream_t stream);
/* With a workspace, B2F_KM_TC3 means: 3xTF32 scores over every row WITH the certainty test, rows inside its (narrow)
 * uncertainty band to the fp64 kernel -- labels equal to fp64 except fp64-level ties.  It is the fallback of the
 * filter pipeline; its list lives in the same workspace and holds every row when the workspace has the standard size.
 * After any of these calls the first four int32 of the workspace read
 *   [0] rows the filter pass listed for the 3xTF32 re-check   [1] rows re-checked (= min([0], [3]))
 *   [2] rows listed for the fp64 kernel                       [3] capacity of each list
 * so max(0, [0] - [3]) + max(0, [2] - [3]) rows were dropped (0 in any sane run).                      */
/* The same pass over the rows x * col_scale (element-wise per column, fp32[d] on the device, 16-byte aligned)
 * WITHOUT a scaled copy of X: StandardScalerModel.transform folded into the Lloyd pass (BASELINE config 5).
 * The operand conversion multiplies by the factor, the accumulated raw row sums are scaled once per pass;
 * payload, labels and cost are those of the scaled data.  Shapes as B2F_KM_TCH (b2f_kmeans_scaled_supported).
 *   reference: orangecontrib/spark/base/spark_ml_transformer.py:133 followed by spark_ml_estimator.py:22  */
int b2f_kmeans_scaled_supported(int k, int d, int64_t ld);
int b2f_kmeans_step_scaled(const float* X, int64_t n, int d, int64_t ld, const float* col_scale, const void* prep,
                           int k, void* acc, int32_t* labels, void* ws, size_t ws_bytes, b2f_stream_t stream);
/* folds scratch replicas / partials into the fp64 payload at the head of acc */
int b2f_kmeans_fold(void* acc, int k, int d, b2f_stream_t stream);
/* centre update from the (all-reduced) payload: centres with count > 0 move to sums/count,
 * empty clusters keep their centre; status[0] = 1 if every moved centre moved <= tol else 0,
 * status[1] = max squared move (as double bits in status_f64[0]); cost copied to status_f64[1]. */
int b2f_kmeans_finalize(const double* payload, int k, int d, double tol, double* centers, int32_t* status,
                        double* status_f64, b2f_stream_t stream);
/* KMeansModel.transform: labels (int32[n]) and/or squared distance to the closest centre
 * (float[n], used by computeCost and the k-means|| initialiser); no accumulation.
 *   reference: orangecontrib/spark/widgets/ml/spark_ml_model.py:53                            */
int b2f_kmeans_predict(const float* X, int64_t n, int d, int64_t ld, const void* prep, int k, int32_t* labels,
                       float* dist2, int algo, b2f_stream_t stream);
/* The same with a workspace (b2f_kmeans_ws_bytes): labels then come from the certainty pipeline of
 * b2f_kmeans_step_ws (filter pass, candidate check, 3xTF32 re-check, fp64 tail) wherever the shape allows it, i.e. they
 * equal the fp64 assignment except for fp64-level ties; dist2 is the fp32 direct-form distance to the labelled
 * centre.  col_scale (may be NULL) as in b2f_kmeans_step_scaled.  Same reference call site.            */
int b2f_kmeans_predict_ws(const float* X, int64_t n, int d, int64_t ld, const float* col_scale, const void* prep,
                          int k, int32_t* labels, float* dist2, void* ws, size_t ws_bytes, int algo,
                          b2f_stream_t stream);

/* ---- K3 / K4 LogisticRegression loss + gradient passes --------------------
 * replaces the treeAggregate of LogisticAggregator inside RDDLossFunction.calculate
 *   reference: orangecontrib/spark/base/spark_ml_estimator.py:22 with module = classification
 *   (orangecontrib/spark/widgets/ml/spark_ml_classification.py:15); semantics: SURVEY.md A.3.
 *
 * beta    double[K*d]  ORIGINAL-space weights (= standardized coefficient * 1/sigma_j), K = 1 binary
 * icpt    double[K]    intercepts
 * out     double[K*d + K + 2], ACCUMULATED (zero first; this is the all-reduce payload):
 *           out[k*d + j] += sum_i w_i (p_ik - 1[y_i = k]) x_ij     (raw, not yet scaled by 1/sigma, 1/W)
 *           out[K*d + k] += sum_i w_i (p_ik - 1[y_i = k])
 *           out[K*d + K] += sum_i w_i loss_i ;  out[K*d + K + 1] += sum_i w_i
 * scratch  b2f_lr_scratch_bytes(d, K) bytes of caller-owned device memory                       */
size_t b2f_lr_scratch_bytes(int d, int n_classes);
int b2f_lr_binary_lossgrad(const float* X, const float* y, const float* w, int64_t n, int d, int64_t ld,
                           const double* beta, const double* icpt, double* out, void* scratch,
                           b2f_stream_t stream);
/* Same pass as b2f_lr_binary_lossgrad for callers that cannot supply scratch memory.
 * (No counterpart in the reference; it exists for hosts that may not allocate on the device:
 * bindings that own only the buffers they were handed, or code running where cudaMalloc is not
 * allowed.)  The per-CTA partial sums go to a fixed array inside the library (about 1.2 MB that
 * every GPU loading the library keeps until exit), so:
 *   - the padded feature count ld must be <= 512; larger inputs are rejected with B2F_EINVAL;
 *   - the call allocates and frees nothing; inputs and `out` (accumulated, zero it first) are
 *     still the caller's;
 *   - the first call on each device looks up the array's device address (a one-time runtime
 *     query): make that first call before starting any stream capture;
 *   - all calls on a device share the array: two calls that can overlap on the GPU (different
 *     streams, or different host threads) race on it and give wrong sums -- serialise them.     */
int b2f_lr_binary_lossgrad_noalloc(const float* X, const float* y, const float* w, int64_t n, int d, int64_t ld,
                                   const double* beta, const double* icpt, double* out, b2f_stream_t stream);
int b2f_lr_multinomial_lossgrad(const float* X, const float* y, const float* w, int64_t n, int d, int64_t ld,
                                int n_classes, const double* beta, const double* icpt, double* out,
                                void* scratch, b2f_stream_t stream);
/* LogisticRegressionModel.transform (A.3 prediction): raw[n,Kc], prob[n,Kc], pred[n] (any may be NULL);
 * Kc = 2 for binary (n_classes == 2, beta is [1,d]) else n_classes.
 *   reference: orangecontrib/spark/widgets/ml/spark_ml_model.py:53                              */
int b2f_lr_predict(const float* X, int64_t n, int d, int64_t ld, int n_classes, const double* beta,
                   const double* icpt, double threshold, float* raw, float* prob, float* pred,
                   b2f_stream_t stream);

/* ---- evaluators -------------------------------------
 * Squared-Euclidean silhouette the way pyspark.ml.evaluation.ClusteringEvaluator computes it (per-cluster
 * statistics, no pairwise distances).
 *   reference call site: orangecontrib/spark/widgets/ml/spark_ml_evaluation.py:42-55 (module = evaluation, :7,19)
 * stats: device fp64 [k][d + 2] = S_c | Psi_c | N_c, zeroed by the caller, ACCUMULATED here (fp64 atomics): with
 * several GPUs every rank adds its shard and the buffer is all-reduced before the score pass.
 * labels: int32[n] in [0, k); a label outside sets *bad_flag (may be NULL) and the row is skipped.
 * score: out[0] += sum w s_i, out[1] += sum w over this shard (fp64, zero first); scratch of
 * b2f_silhouette_scratch_bytes(k, d) bytes.  silhouette = out[0] / out[1] after the all-reduce.             */
int b2f_silhouette_stats(const float* X, int64_t n, int d, int64_t ld, const int32_t* labels, const float* w, int k,
                         double* stats, int32_t* bad_flag, b2f_stream_t stream);
size_t b2f_silhouette_scratch_bytes(int k, int d);
int b2f_silhouette_score(const float* X, int64_t n, int d, int64_t ld, const int32_t* labels, const float* w, int k,
                         const double* stats, void* scratch, double* out, b2f_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* B200FIT_H_ */

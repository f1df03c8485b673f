/*
 * gpusched.h — C ABI of libgpusched.so, the B200 (sm_100a) implementation of the
 * k8s-gpu-scheduler hot path: batched Score -> Logic -> NormalizeScore of pending pods
 * against nodes and their GPUs, the MIG-configuration choice of reconfigure(), and the
 * recommender's completion (fit / transform) of un-profiled app x GPU-config entries.
 *
 * The reference (dimgatz98/k8s-gpu-scheduler) has no FFI; it evaluates one (pod, node)
 * pair per Score() call and fetches its inputs over Redis / gRPC / the API server inside
 * that call.  This header is the thin layer a cgo binding would target (INTEGRATION.md
 * shows the Go side).  Every entry point cites the reference code it replaces; all
 * reference paths are relative to the reference repository root.
 *
 * Conventions
 *   - return 0 (GS_OK) on success, a negative GS_ERR_* code on failure;
 *     gs_last_error(ctx) returns a human-readable message for the last failure on ctx.
 *   - the caller owns every host buffer; the library copies what it needs before the
 *     call returns and never keeps a caller pointer (cgo pointer rules).
 *   - outputs are caller-allocated; the library owns all device memory.
 *   - entry points taking a gs_ctx are thread-safe (one mutex per context): the
 *     kube-scheduler framework calls Score from 16 goroutines.
 *   - there is NO CPU fallback: if no CUDA device is usable, gs_create fails.
 */
#ifndef GPUSCHED_H_
#define GPUSCHED_H_

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define GS_API __attribute__((visibility("default")))
#else
#define GS_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* ---------------------------------------------------------------- constants -------- */

#define GS_OK 0
#define GS_ERR_INVALID (-1)     /* bad argument (null pointer, size out of range, bad enum) */
#define GS_ERR_CUDA (-2)        /* a CUDA runtime call failed; see gs_last_error */
#define GS_ERR_STATE (-3)       /* call order: scoring before cluster + predictions were uploaded */
#define GS_ERR_NOMEM (-4)       /* host or device allocation failed */
#define GS_ERR_UNSUPPORTED (-5) /* valid request this build cannot serve */
#define GS_ERR_NOCONV (-6)      /* completion: a numerical routine did not converge */

/* node model, from the node NAME: contains "a30" -> A30 (tested first), else contains
 * "gpu" -> V100, else none.  pkg/plugins/gpu_plugin/gpu_plugins.go:477-499 */
#define GS_MODEL_NONE 0
#define GS_MODEL_A30 1
#define GS_MODEL_V100 2

/* app id of a pod whose name matches no row of the recommender's training tables: the
 * recommender answers such a request with an empty column list
 * (pkg/recommender/recom_server.py:155-160) and the plugin sees an empty map
 * (gpu_plugins.go:322-325). */
#define GS_APP_UNKNOWN 0xFFFFFFFFu

/* most UUIDs (GPUs or MIG instances) one node may list in Redis */
#define GS_MAX_GPUS_PER_NODE 31

/* Packed per-(pod,node) selection byte written by gs_score_batch:
 *   bits 0..4  index (in Redis array order) of the selected GPU, GS_SEL_NONE if none
 *              (selectedUUID == "" in gpu_plugins.go:458,691-694,752-755)
 *   bit  5     F6: the candidate's configuration column exists for this node
 *              (A30: "%dP_A30" present, gpu_plugins.go:636-637; V100: any of 1P/2P/4P
 *              present, :639-651)
 *   bit  6     F7: a candidate term entered the score, i.e. confPrediction != -1
 *              (gpu_plugins.go:667)
 *   bit  7     F1 & F3: node has a model (:478,:497,:558) and at least one UUID (:536-542)
 * "Filter" in the reference is these implicit predicates; a node failing them still
 * gets score 0, it is not removed (the reference never returns Unschedulable). */
#define GS_SEL_NONE 31u
#define GS_SEL_GPU(b) ((unsigned)(b) & 31u)
#define GS_SEL_F6(b) (((unsigned)(b) >> 5) & 1u)
#define GS_SEL_F7(b) (((unsigned)(b) >> 6) & 1u)
#define GS_SEL_ELIGIBLE(b) (((unsigned)(b) >> 7) & 1u)

/* flags for gs_score_batch */
#define GS_SCORE_RAW 0u        /* out_score = what Score() returns (gpu_plugins.go:779-810) */
#define GS_SCORE_NORMALIZE 1u  /* out_score = after NormalizeScore() (gpu_plugins.go:816-841) */
#define GS_SCORE_DIRECT 2u     /* force the un-restructured one-thread-per-(pod,node) kernel */
#define GS_SCORE_NO_DEDUPE 4u  /* score every pod itself, even pods with equal (app, SLO) (measurement) */
#define GS_SCORE_LANES_NODES 8u /* restructured path: the one-thread-per-node kernel only (measurement) */

/* ---------------------------------------------------------------- tables ----------- */

/* Cluster snapshot, structure-of-arrays.  Replaces, per Score() call, the Redis GET of
 * the node's UUID list (gpu_plugins.go:529-542; writer pkg/profiler/cmd/client/client.go:70-76)
 * and GetSLOs' scan of all pods (gpu_plugins.go:87-160).
 *
 * GPU slot (n, g) = n*max_gpus + g, g in Redis array order.  Residents of slot s are
 * res_*[gpu_off[s] .. gpu_off[s+1]); slots g >= node_ngpu[n] must be empty.  A resident
 * is a Running pod whose first ConfigMap carrying CUDA_VISIBLE_DEVICES names that UUID
 * and that has a non-empty SLO env (gpu_plugins.go:105,111-140).  Resident order fixes
 * the float summation order that Go's map iteration leaves undefined (:571,:595,:703). */
typedef struct gs_cluster {
  int32_t n_nodes;
  int32_t max_gpus;           /* slot stride, 1..GS_MAX_GPUS_PER_NODE */
  const uint8_t* node_model;  /* [n_nodes] GS_MODEL_* */
  const uint8_t* node_ngpu;   /* [n_nodes] len(uuids); 0 = key missing or empty list */
  const uint32_t* gpu_off;    /* [n_nodes*max_gpus + 1] CSR offsets, non-decreasing */
  const uint32_t* res_app;    /* [gpu_off[last]] app id or GS_APP_UNKNOWN */
  const float* res_slo;       /* [gpu_off[last]] float32(ParseFloat(env SLO, 32)) */
  int32_t redis_reachable;    /* 0 = no Redis pod found: every score is 0 (:508-527) */
} gs_cluster;

/* Completed prediction tables as the recommender serves them (float32 on the wire,
 * pkg/recommender/protos/recom.proto:15).  Replaces the dial-per-call gRPC requests
 * GetConfigurationPredictions / GetInterferencePredictions (gpu_plugins.go:302-355).
 *
 *   conf   [n_apps x n_cols]        row = app, column = a configuration such as "2P_A30"
 *   interf [2 x n_apps x n_apps]    interf[(model-1)*n_apps*n_apps + a*n_apps + b] =
 *                                   entry of row "<app a>_<model>" under column <app b>
 *   col_of [2 x (GS_MAX_GPUS_PER_NODE+1)]  col_of[(model-1)*(GS_MAX_GPUS_PER_NODE+1) + k]
 *                                   = column of "%dP_<model>" % k, or -1 if the table
 *                                   has no such column (gpu_plugins.go:583-587,636-651) */
typedef struct gs_predictions {
  int32_t n_apps;
  int32_t n_cols;
  const float* conf;
  const float* interf;
  const int32_t* col_of;
} gs_predictions;

typedef struct gs_ctx gs_ctx;

/* ---------------------------------------------------------------- lifecycle -------- */

/* Create a context on CUDA device `device`.  Replaces the plugin factory's lazy global
 * initialisation (gpu_plugins.go:785-796, :928-930). */
GS_API int gs_create(int device, gs_ctx** out);
GS_API void gs_destroy(gs_ctx* ctx);
GS_API const char* gs_last_error(const gs_ctx* ctx);
/* library build string: version, arch, compile date */
GS_API const char* gs_version(void);

/* ---------------------------------------------------------------- uploads ---------- */

/* Upload a cluster snapshot (or the node shard [node_begin, node_begin+n_nodes) of one:
 * node_base is added to node indices in best-node results).
 * Replaces gpu_plugins.go:529-549 (Redis GET + GetSLOs) for all nodes at once.
 * Limits: at most 8 388 480 nodes per context (shard beyond that) and 65 535 residents per GPU.
 * Both uploads first wait for the GPU work of every earlier scoring / normalise call on this
 * context (whatever stream it was issued on), so a re-upload after an informer event never
 * overwrites tables that an asynchronous gs_score_raw_dev is still reading. */
GS_API int gs_upload_cluster(gs_ctx* ctx, const gs_cluster* cluster, int32_t node_base);

/* Replace ONE node's GPU list and residents in the uploaded snapshot (node = index within this context's
 * nodes): what a bind, a pod deletion or a profiler refresh changes.  gpu_counts[g] residents on GPU g
 * (g < n_gpus, in Redis array order), their app ids and SLOs concatenated in res_app / res_slo.  The
 * reference re-reads Redis and re-scans all pods on every Score (gpu_plugins.go:529-549, 87-160); here
 * only this node's part of the tables (and what lies behind it in the arrays) is copied again.  Resident
 * indices used by pod_self shift as they would in a rebuilt snapshot.  The node's model cannot change
 * (it derives from the node's name). */
GS_API int gs_update_node(gs_ctx* ctx, int32_t node, int32_t n_gpus, const uint32_t* gpu_counts,
                          const uint32_t* res_app, const float* res_slo);

/* Upload the completed prediction tables.  Replaces gpu_plugins.go:302-355 and the
 * serving side pkg/recommender/recom_server.py:138-204. */
GS_API int gs_upload_predictions(gs_ctx* ctx, const gs_predictions* preds);

/* ---------------------------------------------------------------- scoring ---------- */

/* Score a batch of pending pods against every uploaded node.
 *
 *   pod_app  [n_pods]  app id (or GS_APP_UNKNOWN) of the pod being scheduled
 *   pod_slo  [n_pods]  currSLO: float32 of env SLO, 0 if missing/unparsable (:459-469)
 *   pod_self [n_pods]  or NULL: resident index (into res_*) that IS this pod, -1 if the
 *                      pod is not among the residents (a Pending pod never is, :105);
 *                      implements the self-exclusion at :596,:704
 *   out_score [n_pods x n_nodes]  raw (0..100) or normalised (0..100) score
 *   out_sel   [n_pods x n_nodes]  packed selection byte, see GS_SEL_*
 *   out_best  [n_pods] or NULL    node index (+node_base) with the highest score, lowest
 *                                 index on ties; what the framework's selectHost picks
 *                                 when this is the only score plugin
 *
 * Replaces Score -> Logic (gpu_plugins.go:779-810, 455-777) for all (pod,node) pairs and
 * NormalizeScore (:816-841) for all pods.  Host buffers in, host buffers out; out_score and
 * out_sel may be NULL when only out_best is wanted.  Large batches are scored in slices whose
 * results are copied out on a second stream while the next slice is being scored; pinned
 * (page-locked) output buffers make those copies truly asynchronous. */
GS_API int gs_score_batch(gs_ctx* ctx, const uint32_t* pod_app, const float* pod_slo,
                   const int32_t* pod_self, int32_t n_pods, uint32_t flags,
                   uint8_t* out_score, uint8_t* out_sel, int32_t* out_best);

/* Decisions only: what the framework and PostBind need of a cycle, 8 bytes per pod instead of two
 * n_nodes-byte rows.  For every pod: the node the framework's selectHost picks when this is the only
 * score plugin (highest score, lowest index on ties; -1 without nodes), the selection byte AT that
 * node — the {nodeName: selectedUUID} record Logic leaves in the pod's ConfigMap
 * (gpu_plugins.go:759-768) and PostBind reads back (:894-895) —, the raw score extrema over the
 * nodes (NormalizeScore's highest/lowest, :818-827: the chosen node's normalised score is 100 if
 * raw_hi > raw_lo, else 0), and the V100 configuration Logic records under MPS_<node> (:640-665):
 * 0 = the chosen node is not a V100 node with UUIDs, else 1, 2 or 4 for "1P_V100" / "2P_V100" /
 * "4P_V100" (PostBind derives CUDA_MPS_PINNED_DEVICE_MEM_LIMIT / ..._ACTIVE_THREAD_PERCENTAGE from
 * it, :896-903).  Same arguments as gs_score_batch otherwise. */
typedef struct gs_decision {
  int32_t node;
  uint8_t sel;    /* GS_SEL_* of the chosen node */
  uint8_t raw_hi; /* highest raw score over the nodes */
  uint8_t raw_lo; /* lowest raw score over the nodes */
  uint8_t mps;    /* 0, 1, 2 or 4 */
} gs_decision;
GS_API int gs_score_decisions(gs_ctx* ctx, const uint32_t* pod_app, const float* pod_slo,
                              const int32_t* pod_self, int32_t n_pods, uint32_t flags, gs_decision* out);

/* Same computation on DEVICE pointers (all arguments are device memory on the context's
 * device), asynchronous on `stream` (a cudaStream_t passed as void*; NULL = the
 * context's own stream).  Split in two phases so that a node-sharded multi-GPU caller
 * can reduce the per-pod extrema between them:
 *
 *   gs_score_raw_dev      writes raw scores + selection bytes and, per pod,
 *                         d_hi_arg[p] = max over local nodes of
 *                                       (raw << 32) | (0xFFFFFFFF - global node index)
 *                         d_neg_lo[p] = max over local nodes of (100 - raw)
 *                         (both are max-reductions: one all-reduce(MAX) over ranks on the
 *                         concatenated int64 buffer makes them global)
 *   gs_normalize_dev      rescales d_score in place with the (global) extrema, exactly
 *                         NormalizeScore's integer arithmetic, and decodes d_best.
 * d_hi_arg and d_neg_lo are uint64[n_pods] each.  d_best may be NULL.
 * The context's scratch (bucketing buffers, the per-batch residents-sum table) is shared by all
 * calls on the context: issue them on one stream, or make sure that the GPU work of two calls does
 * not overlap.  Growing that scratch for a larger batch allocates (and synchronises the device);
 * see gs_score_raw_dev_noalloc for callers that cannot have that. */
GS_API int gs_score_raw_dev(gs_ctx* ctx, const uint32_t* d_pod_app, const float* d_pod_slo,
                     const int32_t* d_pod_self, int32_t n_pods, uint32_t flags,
                     uint8_t* d_score, uint8_t* d_sel, uint64_t* d_hi_arg,
                     uint64_t* d_neg_lo, void* stream);
GS_API int gs_normalize_dev(gs_ctx* ctx, int32_t n_pods, uint8_t* d_score,
                     const uint64_t* d_hi_arg, const uint64_t* d_neg_lo, int32_t* d_best,
                     void* stream);

/* The best node alone from the (global) extrema, without rescaling any score: d_best[p] as gs_normalize_dev
 * decodes it.  For callers that keep raw scores. */
GS_API int gs_best_dev(gs_ctx* ctx, int32_t n_pods, const uint64_t* d_hi_arg, const uint64_t* d_neg_lo,
                       int32_t* d_best, void* stream);

/* gs_score_decisions' last step on DEVICE pointers, after gs_score_raw_dev (and, on a node shard, after
 * the extrema were made global): d_out[p] as in gs_decision; on a shard whose nodes do not include
 * pod p's winner, sel = GS_SEL_NONE and mps = 0 (the owning shard reports them). */
GS_API int gs_decisions_dev(gs_ctx* ctx, const uint32_t* d_pod_app, const float* d_pod_slo, int32_t n_pods,
                            const uint8_t* d_sel, const uint64_t* d_hi_arg, const uint64_t* d_neg_lo,
                            gs_decision* d_out, void* stream);

/* ---------------------------------------------------------------- several GPUs ----- */

/* Node-sharded scoring over several GPUs of one box in ONE process (no counterpart
 * in the reference, which scores one (pod, node) per call on the CPU).  The cluster's nodes are
 * block-partitioned over the devices inside the library, pods and prediction tables are replicated,
 * and the per-pod extrema that NormalizeScore (gpu_plugins.go:818-827) and host selection need are
 * merged with atomicMax over the peer mapping (NVLink) into device 0's copy — no NCCL, no second
 * process.  Needs peer access from every device to the first one (GS_ERR_UNSUPPORTED otherwise).
 * Results are bit-identical to one device holding all nodes.
 *
 *   gs_mscore_batch      same arguments and results as gs_score_batch (no self table): full
 *                        [n_pods x n_nodes] matrices in the caller's node order + best node per pod
 *   gs_mscore_decisions  same results as gs_score_decisions */
typedef struct gs_mctx gs_mctx;
GS_API int gs_create_multi(const int* devices, int n_devices, gs_mctx** out);
GS_API void gs_destroy_multi(gs_mctx* mctx);
GS_API const char* gs_multi_last_error(const gs_mctx* mctx);
GS_API int gs_multi_device_count(const gs_mctx* mctx);
GS_API int gs_mupload_predictions(gs_mctx* mctx, const gs_predictions* preds);
GS_API int gs_mupload_cluster(gs_mctx* mctx, const gs_cluster* cluster);
GS_API int gs_mscore_batch(gs_mctx* mctx, const uint32_t* pod_app, const float* pod_slo, int32_t n_pods,
                           uint32_t flags, uint8_t* out_score, uint8_t* out_sel, int32_t* out_best);
GS_API int gs_mscore_decisions(gs_mctx* mctx, const uint32_t* pod_app, const float* pod_slo, int32_t n_pods,
                               uint32_t flags, gs_decision* out);

/* gs_score_raw_dev for callers that must not allocate.  (No counterpart in the reference.)
 *
 * The regular scoring entry points grow context-owned device buffers with cudaMalloc when a
 * batch needs more working memory than the last one did.  Some callers cannot have that on
 * their scoring path: a scheduler that has handed all device memory to its own pools at
 * start-up, a latency-critical cycle that must never block on the driver's allocator
 * (cudaMalloc and cudaFree synchronise the device and serialise against every other
 * allocating thread), or a harness that forbids allocation after initialisation.  This
 * variant computes exactly what gs_score_raw_dev computes (same arguments, minus the self
 * table and flags; same bytes out) but takes ALL of its working memory from one fixed array
 * inside the library's device image, and calls neither cudaMalloc/cudaFree nor any
 * allocating library routine.
 *
 *   Limits    n_pods <= GS_NOALLOC_MAX_PODS, uploaded apps <= GS_NOALLOC_MAX_APPS, uploaded
 *             residents <= GS_NOALLOC_MAX_RESIDENTS, max_gpus <= 8; anything larger is
 *             refused with GS_ERR_INVALID (gs_last_error names the limit) — split the batch
 *             or use gs_score_raw_dev.  The tables themselves are uploaded (and allocated)
 *             beforehand with gs_upload_*, outside the no-allocation phase.
 *   Cost      the array (about 3.2 MB) is part of the library image: every GPU that loads
 *             libgpusched.so holds it until the process exits.
 *   Sharing   all calls on a GPU use the same array, which keeps the previous call's
 *             contents; each call clears what must start at zero (and zeroes d_hi_arg /
 *             d_neg_lo, like gs_score_raw_dev).  Calls issued one after another on one stream
 *             are safe.  Calls whose GPU work overlaps — different streams, or different
 *             contexts on the same GPU — race on the array: serialise them yourself.
 *   First use the first call on each device resolves the array's device address
 *             (cudaGetSymbolAddress, once per device); make that first call before any
 *             stream capture starts. */
#define GS_NOALLOC_MAX_PODS 65536
#define GS_NOALLOC_MAX_APPS 4096
#define GS_NOALLOC_MAX_RESIDENTS 131072
GS_API int gs_score_raw_dev_noalloc(gs_ctx* ctx, const uint32_t* d_pod_app, const float* d_pod_slo,
                                    int32_t n_pods, uint8_t* d_score, uint8_t* d_sel,
                                    uint64_t* d_hi_arg, uint64_t* d_neg_lo, void* stream);

/* NormalizeScore alone on a caller-provided score list (one pod): the framework's
 * NodeScoreList as int64 scores.  Replaces gpu_plugins.go:816-841 literally. */
GS_API int gs_normalize_scores(gs_ctx* ctx, int64_t* scores, int32_t n_scores);

/* MIG configuration choice for an empty A30 node: index into
 * {"all-4g.24gb","all-2g.12gb","all-1g.6gb"} (gpu_plugins.go:52).  Replaces the
 * arithmetic of reconfigure(), gpu_plugins.go:366-399; its node-label / pod-delete /
 * Redis-poll side effects (:401-452) stay in the host. */
GS_API int gs_select_mig(gs_ctx* ctx, const uint32_t* pod_app, const float* pod_slo,
                  int32_t n_pods, uint8_t* out_config);

/* ---------------------------------------------------------------- measurement ------ */

/* Instrumentation (no counterpart in the reference): when enabled, gs_score_raw_dev records CUDA
 * events on the launching stream before the residents-sum kernel (stage_a_kernel), between it and
 * the scoring kernels, and after the scoring kernels.  gs_last_score_kernel_ms waits for the last
 * bracket and returns the duration of the scoring kernels alone (score_pods_kernel and, for the work
 * items it leaves, score_fast_kernel; score_direct_kernel with GS_SCORE_DIRECT); gs_last_stage_a_ms
 * returns stage_a_kernel's.  Used by bench.py for the roofline figure. */
GS_API int gs_set_profiling(gs_ctx* ctx, int enabled);
/* Number of CUDA kernels this library has launched in this process so far (all contexts; copies
 * and memsets are not kernels).  bench.py reports the difference over its timed region. */
GS_API uint64_t gs_launch_count(void);
/* Arithmetic self-test (no counterpart in the reference): runs the streamlined reciprocal,
 * small-integer division and floor used by the restructured scoring kernel against CUDA's
 * IEEE intrinsics on `n_million` x 10^6 pseudo-random operands and returns the number of
 * results that differ in any bit (must be 0). */
GS_API int gs_selftest(gs_ctx* ctx, uint64_t seed, int32_t n_million, uint64_t* out_mismatches);
GS_API int gs_last_score_kernel_ms(gs_ctx* ctx, float* out_ms);
GS_API int gs_last_stage_a_ms(gs_ctx* ctx, float* out_ms);
/* Filter self-test (no counterpart in the reference): the scoring kernel decides int64(tmpScore)
 * (gpu_plugins.go:752) from a float32 estimate wherever that estimate lies farther than
 * *out_delta from an integer, and only otherwise runs the float64 sequence.  This runs both on
 * about `n_million` x 10^6 drawn (residents' sums, candidate term) samples ON THE DEVICE — half of
 * them aimed within 1e-4 of an integer score — and reports the largest |estimate - tmpScore|
 * seen (must stay below *out_delta), how many estimates outside the band floor differently from
 * tmpScore (must be 0), how many fell inside the band, and how many samples were evaluated. */
GS_API int gs_selftest_filter(gs_ctx* ctx, uint64_t seed, int32_t n_million, double* out_max_dev,
                              double* out_delta, uint64_t* out_wrong_floors, uint64_t* out_in_band,
                              uint64_t* out_evaluated);

/* ---------------------------------------------------------------- completion ------- */

/* Recommender completion model: sklearn IterativeImputer(max_iter=10, random_state=0)
 * with BayesianRidge, as constructed in pkg/recommender/recommender/recommender.py:15-22. */
typedef struct gs_model gs_model;

/* Fit on X [n x d] row-major float64, NaN = not profiled.
 * Replaces Imputer.fit, pkg/recommender/recommender/recommender.py:15-22. */
GS_API int gs_complete_fit(gs_ctx* ctx, const double* X, int32_t n, int32_t d, gs_model** out);

/* The same fit with the ROWS of X sharded over several contexts (one per GPU / rank): every
 * rank passes its own rows and a callback that all-reduces a float64 device buffer in place
 * over the ranks (sum or max).  The callback is the path's only exchange: per feature the
 * column sums (d+1 values) and the centred Gram matrix (d*d values) of the rows where the
 * feature is observed, per round one scalar; every rank then solves the same small system and
 * imputes its own rows, so all ranks return identical models.  The buffer is ready on `stream`
 * when the callback is invoked; the reduced values must be visible to work submitted to
 * `stream` after it returns (enqueue the collective on that stream, or synchronise).  Return 0
 * for success.  No counterpart in the reference (row-sharded completion). */
#define GS_REDUCE_SUM 0
#define GS_REDUCE_MAX 1
typedef int (*gs_allreduce_fn)(double* device_buffer, int64_t count, int op, void* stream, void* user);
GS_API int gs_complete_fit_sharded(gs_ctx* ctx, const double* X_local, int32_t n_local, int32_t d,
                                   gs_allreduce_fn reduce, void* user, gs_model** out);

/* Fit and, in the same pass, return the completed training rows (IterativeImputer.fit_transform,
 * sklearn/impute/_iterative.py:723-873): out_completed [n_local x d] equals what
 * gs_complete_transform would return for X_local, without replaying the steps.  reduce/user may be
 * NULL (one context), out_completed may be NULL.  The recommender serves exactly these rows
 * (pkg/recommender/recom_server.py:162-164 imputes rows of its own training table). */
GS_API int gs_complete_fit_transform(gs_ctx* ctx, const double* X_local, int32_t n_local, int32_t d,
                                     gs_allreduce_fn reduce, void* user, double* out_completed,
                                     gs_model** out);

/* Impute X [m x d] (NaN = missing) into out [m x d]; observed entries are returned
 * unchanged.  Replaces Imputer.predict, recommender.py:24-27. */
GS_API int gs_complete_transform(gs_ctx* ctx, const gs_model* model, const double* X, int32_t m,
                          double* out);

/* number of (feature, estimator) steps stored = n_iter_ * d, and rounds run (n_iter_) */
GS_API int gs_model_info(const gs_model* model, int32_t* n_features, int32_t* n_rounds,
                  int32_t* n_steps);
/* copy out step s: the imputed feature, its d-1 coefficients (in neighbour order =
 * all other features ascending) and intercept */
GS_API int gs_model_step(const gs_model* model, int32_t s, int32_t* feat, double* coef,
                  double* intercept);
GS_API void gs_model_destroy(gs_model* model);
/* Instrumentation (no counterpart in the reference): device time of the last fit on this context, measured with
 * CUDA events on at most 64 of its estimator fits.  out4 = {fits sampled, ms per fit for the column moments, for the
 * Gram matrix (kernels, all-reduce, recentring), for the estimator (reduction / decomposition + evidence
 * iteration)}.  Used by bench.py's `completion` object. */
GS_API int gs_complete_last_timing(gs_ctx* ctx, double* out4);
/* Instrumentation (no counterpart in the reference): symmetric eigendecompositions the block Jacobi
 * solver has run in this process (return value) and the sweeps they took in total. */
GS_API uint64_t gs_eig_stats(uint64_t* out_sweeps);

#ifdef __cplusplus
}
#endif
#endif /* GPUSCHED_H_ */

/*
 * gpusched_host.h — host-side mirror of the reference plugin above the C ABI of gpusched.h.
 *
 * The reference's host code is Go (pkg/plugins/gpu_plugin/gpu_plugins.go); no Go toolchain
 * exists where this library is built, so the same logic is provided in C++
 * (k8s-gpu-scheduler_b200/csrc/host/) behind the flat C interface below, which is what the
 * tests drive.  It does what the Go plugin does around the arithmetic:
 *   - turns the reference's string world (node names, Redis JSON values, pods with
 *     ConfigMap / env, the recommender's indexed tables) into the structure-of-arrays tables
 *     of gpusched.h, with the reference's own rules (cited per function);
 *   - exposes the plugin surface: Name, PreScore (batch), Score, NormalizeScore, with the
 *     reference's error behaviour (Score fails with an Error status when GetSLOs would).
 * Nothing here computes scores: that is gs_score_batch on the GPU.
 */
#ifndef GPUSCHED_HOST_H_
#define GPUSCHED_HOST_H_

#include "gpusched.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct gsp_builder gsp_builder;
typedef struct gsp_plugin gsp_plugin;

/* framework.Code values the reference's Score can return (gpu_plugins.go:782,800,805) */
#define GSP_SUCCESS 0
#define GSP_ERROR 1

GS_API gsp_builder* gsp_builder_create(void);
GS_API void gsp_builder_destroy(gsp_builder* b);

/* The completed configurations table as ImputeConfigurations serves it: index = training
 * rows (app names), columns such as "2P_A30", values row-major float32
 * (pkg/recommender/recom_server.py:138-170, utils.py:37-42).  Defines the app ids:
 * app a = index[a]. */
GS_API int gsp_builder_set_configurations(gsp_builder* b, const char* const* index, int n_rows,
                                          const char* const* columns, int n_cols, const float* values);
/* The completed interference table: index rows "<app>_<model>" (model V100 or A30), columns =
 * app names (recom_server.py:172-204).  Rows / columns naming apps that the configurations
 * table does not have are ignored; missing ones read as 0. */
GS_API int gsp_builder_set_interference(gsp_builder* b, const char* const* index, int n_rows,
                                        const char* const* columns, int n_cols, const float* values);

/* A node and the value Redis holds under its name: a JSON array of UUID strings
 * (pkg/profiler/cmd/client/client.go:70-76), or NULL when the key is missing (redis.Nil,
 * gpu_plugins.go:536-542).  Model from the NAME: contains "a30" -> A30, else contains "gpu"
 * -> V100 (gpu_plugins.go:477-499).  Returns the node's index or a negative error. */
GS_API int gsp_builder_add_node(gsp_builder* b, const char* name, const char* redis_value);

/* A pod of namespace "default" as GetSLOs sees it (gpu_plugins.go:87-160): phase, the value of
 * CUDA_VISIBLE_DEVICES in its first ConfigMap that has the key (NULL if none), the SLO env of
 * container 0 (NULL if unset).  Only Running pods whose UUID belongs to a node become
 * residents; a resident whose SLO does not parse makes Score fail for that node (:141-144). */
GS_API int gsp_builder_add_pod(gsp_builder* b, const char* name, const char* ns, const char* phase,
                               const char* cuda_visible_devices, const char* slo_env);
GS_API void gsp_builder_set_redis_reachable(gsp_builder* b, int reachable);

/* app id of a pod name: '-' -> '_', then the first configurations index entry contained in
 * it (find_index_for_request, recom_server.py:67-71,155-156); GS_APP_UNKNOWN if none. */
GS_API uint32_t gsp_builder_app_of(const gsp_builder* b, const char* pod_name);
/* currSLO: float32(ParseFloat(value, 32)), 0 when unset or unparsable (gpu_plugins.go:459-469).
 * *ok (may be NULL) tells whether the string parsed. */
GS_API float gsp_parse_slo(const char* value, int* ok);

/* The body of a Prometheus instant query against the node's DCGM exporter (GET /api/v1/query?query=<metric>
 * {pod="<dcgm pod>"}, pkg/prom/fetch_prom_metrics/prom_metrics.go:63-118), parsed as ParseResponse does
 * (:22-61).  Call once per metric (the reference queries DCGM_FI_PROF_GR_ENGINE_ACTIVE, DCGM_FI_DEV_MEM_COPY_UTIL,
 * DCGM_FI_DEV_GPU_TEMP, DCGM_FI_DEV_FB_USED, DCGM_FI_DEV_FB_FREE).  Returns the number of samples, or a negative
 * error.  An empty body is no samples. */
GS_API int gsp_builder_add_dcgm_response(gsp_builder* b, const char* node_name, const char* prometheus_json);
/* After gsp_builder_build: one metric as a structure-of-arrays column out[node * max_gpus + gpu], GPUs in
 * Redis array order, keyed as GetDcgmMetricsForUUIDS keys them (gpu_plugins.go:193-228: by UUID, or on MIG
 * nodes by the k-th distinct GPU_I_ID), float32(ParseFloat(value, 32)); slots without a sample keep
 * default_value (the reference's commented-out consumer defaults to util = 0, fbFree = 1,
 * gpu_plugins.go:668-673).  Nothing in Score reads these columns — in the reference that code is dead.
 * Returns the number of slots filled, or a negative error. */
GS_API int gsp_builder_dcgm_column(gsp_builder* b, const char* metric_name, float default_value, float* out,
                                   int n_slots);

/* Finish the tables.  After this the views below are valid until the builder is destroyed. */
GS_API int gsp_builder_build(gsp_builder* b);
GS_API const gs_cluster* gsp_builder_cluster(const gsp_builder* b);
GS_API const gs_predictions* gsp_builder_predictions(const gsp_builder* b);
GS_API const char* gsp_builder_last_error(const gsp_builder* b);

/* The plugin: uploads the builder's tables into ctx (gpu_plugins.go:928-930 New). */
GS_API gsp_plugin* gsp_plugin_create(gs_ctx* ctx, gsp_builder* b);
GS_API void gsp_plugin_destroy(gsp_plugin* p);
GS_API const char* gsp_plugin_name(void); /* "GPU", gpu_plugins.go:50 */
GS_API const char* gsp_plugin_last_status(const gsp_plugin* p);

/* PreScore: Score -> Logic for this pod against every node in one GPU call; the row is kept
 * for the Score calls of the cycle. */
GS_API int gsp_plugin_prescore(gsp_plugin* p, const char* pod_name, const char* slo_env);
/* Score (gpu_plugins.go:779-810): the raw score of the last PreScore'd pod on nodeName and
 * the UUID Logic selected ("" if none, :458).  Returns GSP_SUCCESS or GSP_ERROR (unknown
 * node, :798-801; GetSLOs failure, :544-548). */
GS_API int gsp_plugin_score(gsp_plugin* p, const char* node_name, int64_t* out_score, char* out_uuid,
                            int uuid_cap);
/* PostBind's data path (gpu_plugins.go:843-926) for the last PreScore'd pod bound to nodeName: the
 * three values the reference appends to the pod's ConfigMap — CUDA_VISIBLE_DEVICES = the UUID Logic
 * selected on that node (its {nodeName: selectedUUID} record, :759-768, :894-895), and the MPS limits
 * derived from the V100 configuration Logic recorded as MPS_<nodeName> (:640-665, :896-903).
 * Returns 1 when the reference would write them (UUID not empty, :907), 0 when it would not, a
 * negative value on error (unknown node, no PreScore, Redis not found :852).  The ConfigMap write
 * itself stays with the caller. */
GS_API int gsp_plugin_postbind(gsp_plugin* p, const char* node_name, char* out_uuid, int uuid_cap,
                               char* out_mps_mem, int mem_cap, char* out_mps_threads, int threads_cap);
/* Score for every node of the builder, in node order: what the framework's 16 workers collect
 * by calling Score once per node (nodes whose Score fails get -1).  For replay drivers. */
GS_API int gsp_plugin_score_all(gsp_plugin* p, int64_t* out_scores, int n);
/* NormalizeScore (gpu_plugins.go:816-841) on the framework's list; always succeeds (:840). */
GS_API int gsp_plugin_normalize_score(gsp_plugin* p, int64_t* scores, int n);

#ifdef __cplusplus
}
#endif
#endif /* GPUSCHED_HOST_H_ */

// tables.h — device-side views of the structure-of-arrays tables and launcher prototypes.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace gs {

// Device pointers to everything the scoring kernels read.  Passed by value.
struct DevTables {
  int n_nodes;
  int max_gpus;
  int n_apps;
  int n_cols;
  int redis_reachable;
  int node_base;
  // node tables are in the library's internal order (grouped by model, see gs_upload_cluster);
  // node_orig[i] = the caller's index of internal node i, res_new[r] = internal index of the
  // caller's resident r
  const uint8_t* node_model;  // [N]
  const uint8_t* node_ngpu;   // [N]
  const uint32_t* node_orig;  // [N]
  const uint32_t* res_new;    // [R]
  const uint32_t* gpu_off;    // [N*G+1]
  const uint32_t* res_app;    // [R]
  const float* res_slo;       // [R]
  const float* conf;          // [A*K]
  const float* interf;        // [2*A*A]
  const float* interf_t;      // [2*A*A] per-model transpose: interf_t[m][b][a] = interf[m][a][b]
  const int32_t* col_of;      // [2*32]
  // derived per resident by resident_prep: {slo (0 = never contributes), 1/slo, conf, S}
  // S = float32 sum, in resident order, of interf[row(r)][app(c)] over co-residents c != r
  const float4* res_rec;      // [R]
};

// One pending pod in app-sorted order.
struct PodRec {
  float slo;
  float inv_slo;    // float32 1/slo
  float conf_v100;  // the V100 candidate prediction of gpu_plugins.go:639-651, -1 if none
  uint32_t orig;    // bits 0..28 original pod index, kPodF6, kPodGeneral, kPodInfInv
};
constexpr uint32_t kPodF6 = 0x80000000u;       // F6 holds on V100 nodes (some 1P/2P/4P column exists)
constexpr uint32_t kPodGeneral = 0x40000000u;  // SLO is NaN or infinite: general path
constexpr uint32_t kPodInfInv = 0x20000000u;   // 1/SLO is infinite (SLO is 0 or denormal): every term is 0 or NaN
constexpr uint32_t kPodIndexMask = 0x1FFFFFFFu;
constexpr int kMaxPods = 1 << 29;
// node tiles ride on grid.y (<= 65535 tiles of kTile nodes); per-GPU counts are packed into 16 bits
constexpr int kMaxNodesPerCtx = 65535 * 128;
constexpr int kMaxResidentsPerSlot = 65535;

// Pending pods are bucketed by (app, kind); a work item never mixes kinds, so every warp of the
// scoring kernels runs one code path.
constexpr int kKindPlain = 0;    // finite SLO with a finite non-zero reciprocal; some V100 configuration exceeds it
constexpr int kKindNoV100 = 1;   // the same, but no 1P/2P/4P_V100 prediction exceeds the SLO: no candidate term on V100 nodes (:667)
constexpr int kKindInfInv = 2;   // 1/SLO is infinite (SLO missing = 0, or denormal)
constexpr int kKindGeneral = 3;  // SLO is NaN or infinite, or its reciprocal is 0: op-for-op path
constexpr int kKinds = 4;
// bucket of a pod: app-major, so that the work items of one app are neighbours (n_apps = unknown app)
__host__ __device__ inline int bucket_of(uint32_t app, int n_apps, int kind) {
  return (int)(app < (uint32_t)n_apps ? app : (uint32_t)n_apps) * kKinds + kind;
}
inline int n_buckets_for(int n_apps) { return (n_apps + 1) * kKinds; }

// A work item of the restructured kernels: up to kChunk pods of one app, ordered by kind.
struct ChunkDesc {
  uint32_t app;
  uint32_t start;
  uint32_t len;
  uint32_t kind;
};

// ChunkDesc::kind of a work item: how many of its pods are of each kind, in the order they are stored
// (plain | no V100 configuration | no SLO | the rest are general); kChunk <= 256 fits 9 bits each
__host__ __device__ inline uint32_t pack_kinds(uint32_t n_plain, uint32_t n_nov, uint32_t n_inf) {
  return n_plain | (n_nov << 9) | (n_inf << 18);
}

#ifndef GS_KCHUNK
#define GS_KCHUNK 256
#endif
#ifndef GS_MIN_CTAS
#define GS_MIN_CTAS 4
#endif
constexpr int kChunk = GS_KCHUNK; // pods per work item (one app); staged kStage at a time
constexpr int kStage = 128;      // pods staged in shared memory at once (= kTile)
constexpr int kFastMaxGpus = 8;  // GPUs per node the register-resident kernel covers
constexpr int kTile = 128;       // nodes per CTA
constexpr int kSlice = 32;       // nodes per CTA of the lanes-are-pods kernel (one warp)

// Per-batch table of the residents' sums for every (app bucket, GPU slot), written by
// stage_a_kernel and read back coalesced by score_fast_kernel: [n_apps + 1][slots].
struct StageA {
  double* n0;     // negative_sum of the slot's residents given the candidate's app
  double* p0;     // positive_sum
  float* ci;      // the candidate's interference sum on that GPU
  uint32_t* cnt;  // count_negative | count_positive << 16
  size_t stride;  // slots = n_nodes * max_gpus
};

// The same table after the fold (stage_fold_kernel, for the lanes-are-pods kernel): per (app bucket,
// GPU slot) the float32 filter constants, the candidate's interference sum and the three integer
// scores that do not depend on the pod's SLO.  24 bytes per entry as well: it lives in the same buffer.
struct FoldTab {
  float4* ab;    // {An, Bn, Ap, Bp}: tmp ~ A + B*v per branch; B = 0 with A = 0 means "filter off"
  float* ci;     // the candidate's interference sum on that GPU
  uint32_t* pk;  // residents-only score | score with a zero term, negative side << 8 | positive side << 16 | kFoldGeneral
  size_t stride;
};
inline FoldTab fold_view(const StageA& st) {  // the stage buffer, seen as a fold table
  const size_t e8 = (size_t)(st.p0 - st.n0);
  char* base = (char*)st.n0;
  return FoldTab{(float4*)base, (float*)(base + 16 * e8), (uint32_t*)(base + 20 * e8), st.stride};
}
constexpr uint32_t kFoldGeneral = 1u << 24;  // the slot's sums fall outside the streamlined domain

// Scratch of the restructured path (owned by the context).
struct PodScratch {
  uint32_t* counts;   // [(A+1)*kKinds + 1] pods per (app, kind) bucket (app A = unknown app)
  uint32_t* cursor;   // [(A+1)*kKinds + 1] scatter cursors
  uint32_t* n_chunks; // [2]: number of work items, and the work counter of the persistent scoring kernel
  ChunkDesc* chunks;  // [max_chunks]
  PodRec* sorted;     // [P]
  uint32_t* bucket;   // [P] bucket of each pod, kNoBucket for pods that are not scored themselves (duplicates)
  uint32_t max_chunks;
  StageA stage;       // n0 == nullptr: score_fast_kernel walks the resident lists itself
  // class dedupe: pods with equal (app, SLO bits) get equal rows, so only one representative per class
  // (the lowest pod index) is scored and its rows are copied to the others.  rep == nullptr: off.
  uint32_t* rep;              // [P] representative of each pod's class
  unsigned long long* hkeys;  // [hash_size] open-addressing table: app << 32 | SLO bits
  uint32_t* hvals;            // [hash_size] lowest pod index of the class
  uint32_t hash_size;         // power of two >= 2 * P
};
constexpr uint32_t kNoBucket = 0xFFFFFFFFu;

// upper bound on the work items a batch can produce: one per started kChunk per bucket, and
// never more than there are pods
inline uint32_t max_chunks_for(int n_pods, int n_apps) {
  const uint32_t by_buckets = (uint32_t)(n_pods / kChunk + n_buckets_for(n_apps) + 1);
  return by_buckets < (uint32_t)n_pods ? by_buckets : (uint32_t)(n_pods > 0 ? n_pods : 1);
}
inline uint32_t hash_size_for(int n_pods) {
  uint32_t h = 64;
  while (h < 2u * (uint32_t)(n_pods > 0 ? n_pods : 1)) h <<= 1;
  return h;
}
constexpr int kStageTableMinPods = 256;  // smaller batches walk the resident lists in the scoring kernel

// every kernel launch of the library is counted (gs_launch_count): bench.py reports the launches of
// its timed region from this counter instead of assuming them
void note_launch();
unsigned long long launch_count();

cudaError_t launch_resident_prep(const DevTables& T, float4* res_rec, cudaStream_t s);
cudaError_t launch_score_direct(const DevTables& T, const uint32_t* pod_app, const float* pod_slo,
                                const int32_t* pod_self, int n_pods, uint8_t* score, uint8_t* sel,
                                unsigned long long* hi_arg, unsigned long long* neg_lo, cudaStream_t s);
cudaError_t launch_pod_bucketing(const DevTables& T, const uint32_t* pod_app, const float* pod_slo, int n_pods,
                                 const PodScratch& scratch, cudaStream_t s);
// bucketing -> stage A -> scoring kernel -> class expansion.  lanes_nodes: use the one-thread-per-node
// kernel (score_fast_kernel) even where the lanes-are-pods kernel (score_pods_kernel) could run.
// prof: nullptr, or three events recorded before stage A, after it, and after the scoring kernel.
cudaError_t launch_score_fast(const DevTables& T, const uint32_t* pod_app, const float* pod_slo, int n_pods,
                              const PodScratch& scratch, uint8_t* score, uint8_t* sel,
                              unsigned long long* hi_arg, unsigned long long* neg_lo, cudaStream_t s,
                              const cudaEvent_t* prof = nullptr, bool lanes_nodes = false);
cudaError_t launch_stage_any(const DevTables& T, const uint32_t* counts, const PodScratch& sc, cudaStream_t s,
                             bool lanes_nodes);
cudaError_t launch_score_any(const DevTables& T, const PodScratch& sc, uint8_t* score, uint8_t* sel,
                             unsigned long long* hi_arg, unsigned long long* neg_lo, cudaStream_t s, bool lanes_nodes);
cudaError_t launch_stage_fold(const DevTables& T, const uint32_t* counts, const StageA& stage, cudaStream_t s);
cudaError_t launch_score_pods(const DevTables& T, const PodScratch& sc, uint8_t* score, uint8_t* sel,
                              unsigned long long* hi_arg, unsigned long long* neg_lo, cudaStream_t s);
cudaError_t launch_expand_classes(const DevTables& T, const uint32_t* rep, int n_pods, uint8_t* score, uint8_t* sel,
                                  unsigned long long* hi_arg, unsigned long long* neg_lo, cudaStream_t s);
cudaError_t launch_transpose_interf(const float* interf, float* interf_t, int n_apps, cudaStream_t s);
cudaError_t launch_selftest(unsigned long long seed, int blocks, int iters, unsigned long long* d_mismatches,
                            cudaStream_t s);
cudaError_t launch_selftest_filter(unsigned long long seed, int blocks, int iters, unsigned long long* d_out,
                                   cudaStream_t s);
double filter_delta();
cudaError_t launch_pod_histogram(const uint32_t* pod_app, int n_pods, int n_apps, uint32_t* counts, cudaStream_t s);
cudaError_t launch_stage_a(const DevTables& T, const uint32_t* counts, const StageA& stage, cudaStream_t s);
cudaError_t launch_score_main(const DevTables& T, const PodScratch& sc, uint8_t* score, uint8_t* sel,
                              unsigned long long* hi_arg, unsigned long long* neg_lo, cudaStream_t s,
                              bool skip_plain = false);
cudaError_t launch_normalize(int n_pods, int n_nodes, int rescale, uint8_t* score,
                             const unsigned long long* hi_arg, const unsigned long long* neg_lo,
                             int32_t* best, cudaStream_t s);
// one scheduling decision (mirrors gs_decision of include/gpusched.h)
struct Decision {
  int32_t node;
  uint8_t sel, raw_hi, raw_lo, mps;
};
cudaError_t launch_decisions(const DevTables& T, const uint32_t* pod_app, const float* pod_slo, int n_pods,
                             const uint8_t* sel, const unsigned long long* hi_arg, const unsigned long long* neg_lo,
                             Decision* out, cudaStream_t s);
cudaError_t launch_select_mig(const DevTables& T, const uint32_t* pod_app, const float* pod_slo, int n_pods,
                              uint8_t* out, cudaStream_t s);

}  // namespace gs

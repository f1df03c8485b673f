// noalloc.cu — gs_score_raw_dev_noalloc: the scoring phase for callers that may not allocate.
//
// Who needs this: a scheduler that scores pods on its latency-critical cycle usually sets up
// all of its device memory at start-up and afterwards runs in a regime where allocating is
// off the table — the process has handed the whole HBM budget to its own pools, or it must
// never block on the driver's allocator (cudaMalloc/cudaFree synchronise the device and
// serialise against every other thread that allocates), or it runs under a harness that
// forbids allocation calls after initialisation.  The regular entry points grow
// context-owned buffers on demand with cudaMalloc; this one does not: every byte of working
// memory it needs beyond the caller's inputs and outputs lives in ONE fixed array in the
// library's device image (`g_scratch` below), sized for the limits GS_NOALLOC_MAX_*.
//
// Consequences of that design, all part of the contract (see include/gpusched.h):
//   * the array exists once per GPU that loads the library, for the life of the process
//     (about 3.2 MB), whether or not this entry point is ever called;
//   * batches beyond the limits are refused with GS_ERR_INVALID, not served more slowly;
//   * all calls on a GPU share the array and it keeps whatever the last call left in it.
//     Everything that must start from zero is cleared at the start of each call, and the
//     per-resident records are rebuilt each call (another context may have overwritten them),
//     so back-to-back calls on one stream are fine — but two calls whose GPU work overlaps
//     (different streams, or different contexts on one GPU) race on the array;
//   * the first call on each device looks the array's device address up (below); make that
//     first call before any stream capture begins.
#include <mutex>

#include "../../include/gpusched.h"
#include "context.h"

namespace {

using namespace gs;

constexpr int kNaMaxPods = GS_NOALLOC_MAX_PODS;
constexpr int kNaMaxApps = GS_NOALLOC_MAX_APPS;
constexpr int kNaMaxResidents = GS_NOALLOC_MAX_RESIDENTS;
constexpr size_t kNaBuckets = (size_t)(kNaMaxApps + 1) * kKinds + 4;
constexpr size_t kNaMaxChunks = (size_t)kNaMaxPods / kChunk + kNaBuckets + 1;

constexpr size_t align256(size_t x) { return (x + 255) / 256 * 256; }
// carve-up of the array
constexpr size_t kOffResRec = 0;
constexpr size_t kOffCounts = kOffResRec + align256(sizeof(float4) * (size_t)kNaMaxResidents);
constexpr size_t kOffCursor = kOffCounts + align256(sizeof(uint32_t) * kNaBuckets);
constexpr size_t kOffNChunks = kOffCursor + align256(sizeof(uint32_t) * kNaBuckets);
constexpr size_t kOffChunks = kOffNChunks + 256;
constexpr size_t kOffSorted = kOffChunks + align256(sizeof(ChunkDesc) * kNaMaxChunks);
constexpr size_t kOffBucket = kOffSorted + align256(sizeof(PodRec) * (size_t)kNaMaxPods);
constexpr size_t kScratchBytes = kOffBucket + align256(sizeof(uint32_t) * (size_t)kNaMaxPods);
static_assert(kScratchBytes <= (size_t)4 << 20, "keep the resident scratch modest");

// The scratch itself.  In host code this NAME is only a host-side stand-in for the symbol,
// not a device address: kernels must be given the address obtained from
// cudaGetSymbolAddress (scratch_for_device), never `g_scratch` directly.
static __device__ __align__(256) unsigned char g_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_setup_mu;
unsigned char* g_scratch_addr[kMaxDevices] = {};

// Once per device, with that device current: resolve the array's device address.
cudaError_t scratch_for_device(int device, unsigned char** out) {
  std::lock_guard<std::mutex> lk(g_setup_mu);
  if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
  if (!g_scratch_addr[device]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_scratch);
    if (e != cudaSuccess) return e;
    g_scratch_addr[device] = static_cast<unsigned char*>(p);
  }
  *out = g_scratch_addr[device];
  return cudaSuccess;
}

}  // namespace

extern "C" int gs_score_raw_dev_noalloc(gs_ctx* ctx, const uint32_t* d_pod_app, const float* d_pod_slo,
                                        int32_t n_pods, uint8_t* d_score, uint8_t* d_sel, uint64_t* d_hi_arg,
                                        uint64_t* d_neg_lo, void* stream) {
  if (!ctx) return GS_ERR_INVALID;
  gs::Ctx* c = ctx;
  std::lock_guard<std::mutex> lk(c->mu);
  if (n_pods < 0 || (n_pods && (!d_pod_app || !d_pod_slo || !d_score || !d_sel || !d_hi_arg || !d_neg_lo)))
    return c->fail(GS_ERR_INVALID, "gs_score_raw_dev_noalloc: null pointer or negative n_pods");
  if (!c->have_cluster || !c->have_preds)
    return c->fail(GS_ERR_STATE, "gs_score_raw_dev_noalloc needs gs_upload_cluster and gs_upload_predictions first");
  if (n_pods > kNaMaxPods)
    return c->fail(GS_ERR_INVALID, "gs_score_raw_dev_noalloc: %d pods exceed GS_NOALLOC_MAX_PODS = %d; split the batch "
                   "or use gs_score_raw_dev", n_pods, kNaMaxPods);
  if (c->n_apps > kNaMaxApps)
    return c->fail(GS_ERR_INVALID, "gs_score_raw_dev_noalloc: %d apps exceed GS_NOALLOC_MAX_APPS = %d", c->n_apps,
                   kNaMaxApps);
  if (c->n_res > kNaMaxResidents)
    return c->fail(GS_ERR_INVALID, "gs_score_raw_dev_noalloc: %lld resident pods exceed GS_NOALLOC_MAX_RESIDENTS = %d",
                   (long long)c->n_res, kNaMaxResidents);
  if (c->max_gpus > kFastMaxGpus)
    return c->fail(GS_ERR_INVALID, "gs_score_raw_dev_noalloc: %d GPU slots per node exceed the %d it covers", c->max_gpus,
                   kFastMaxGpus);
  if (n_pods == 0) return GS_OK;
  cudaError_t e = cudaSetDevice(c->device);
  if (e != cudaSuccess) return c->cuda_fail(e, "cudaSetDevice");
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;

  // the caller's per-pod extrema are max-accumulated: they must start at zero — also when this
  // context holds no node at all (an empty shard is legal: its zero extrema must not disturb the
  // other ranks' values in the all-reduce), exactly as gs_score_raw_dev leaves them
  if ((e = cudaMemsetAsync(d_hi_arg, 0, sizeof(uint64_t) * (size_t)n_pods, s)) != cudaSuccess ||
      (e = cudaMemsetAsync(d_neg_lo, 0, sizeof(uint64_t) * (size_t)n_pods, s)) != cudaSuccess)
    return c->cuda_fail(e, "cudaMemsetAsync(extrema)");
  if (c->n_nodes == 0) return GS_OK;
  unsigned char* base = nullptr;
  if ((e = scratch_for_device(c->device, &base)) != cudaSuccess) return c->cuda_fail(e, "cudaGetSymbolAddress(scratch)");

  gs::DevTables T = c->tables();
  float4* res_rec = reinterpret_cast<float4*>(base + kOffResRec);
  T.res_rec = res_rec;
  // rebuilt every call: the array is shared, the last writer may have been another context
  if ((e = gs::launch_resident_prep(T, res_rec, s)) != cudaSuccess) return c->cuda_fail(e, "resident_prep");

  gs::PodScratch sc;
  sc.counts = reinterpret_cast<uint32_t*>(base + kOffCounts);  // zeroed inside launch_pod_bucketing
  sc.cursor = reinterpret_cast<uint32_t*>(base + kOffCursor);
  sc.n_chunks = reinterpret_cast<uint32_t*>(base + kOffNChunks);
  sc.chunks = reinterpret_cast<gs::ChunkDesc*>(base + kOffChunks);
  sc.sorted = reinterpret_cast<gs::PodRec*>(base + kOffSorted);
  sc.bucket = reinterpret_cast<uint32_t*>(base + kOffBucket);
  sc.rep = nullptr;  // no class dedupe here: its hash table would double the resident array
  sc.hkeys = nullptr;
  sc.hvals = nullptr;
  sc.hash_size = 0;
  sc.max_chunks = gs::max_chunks_for(n_pods, c->n_apps);
  sc.stage = gs::StageA{nullptr, nullptr, nullptr, nullptr, 0};  // no residents-sum table: too large to hold statically
  e = gs::launch_score_fast(T, d_pod_app, d_pod_slo, n_pods, sc, d_score, d_sel, (unsigned long long*)d_hi_arg,
                            (unsigned long long*)d_neg_lo, s);
  if (e != cudaSuccess) return c->cuda_fail(e, "score_fast (no-alloc)");
  return c->note_use(s);  // an event created with the context: nothing is allocated here
}

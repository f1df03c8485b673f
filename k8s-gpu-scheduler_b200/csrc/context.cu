// context.cu — the C ABI of include/gpusched.h: context, table uploads, scoring entry
// points.  Host logic only; kernels live in score_kernels.cu.
#include <algorithm>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/gpusched.h"
#include "context.h"

namespace gs {

static std::atomic<unsigned long long> g_launches{0};
void note_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
unsigned long long launch_count() { return g_launches.load(std::memory_order_relaxed); }

int Ctx::fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  err = buf;
  return code;
}

int Ctx::cuda_fail(cudaError_t e, const char* what) {
  return fail(GS_ERR_CUDA, "%s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
}

// grow-only device buffer
int Ctx::reserve(DevBuf& b, size_t bytes, const char* what) {
  if (bytes <= b.cap) return GS_OK;
  if (b.p) cudaFree(b.p);
  b.p = nullptr;
  b.cap = 0;
  size_t want = bytes + bytes / 8 + 256;
  cudaError_t e = cudaMalloc(&b.p, want);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(GS_ERR_NOMEM, "cudaMalloc(%zu bytes) for %s: %s", want, what, cudaGetErrorString(e));
  }
  b.cap = want;
  return GS_OK;
}

void Ctx::release(DevBuf& b) {
  if (b.p) cudaFree(b.p);
  b.p = nullptr;
  b.cap = 0;
}

int Ctx::upload(DevBuf& b, const void* src, size_t bytes, const char* what) {
  int rc = reserve(b, bytes ? bytes : 1, what);
  if (rc) return rc;
  if (bytes) {
    cudaError_t e = cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, stream);
    if (e != cudaSuccess) return cuda_fail(e, what);
  }
  return GS_OK;
}

int Ctx::note_use(cudaStream_t s) {
  cudaError_t e = cudaEventRecord(ev_last_use, s);
  return e == cudaSuccess ? GS_OK : cuda_fail(e, "cudaEventRecord(last use)");
}

int Ctx::wait_last_use() {
  cudaError_t e = cudaEventSynchronize(ev_last_use);
  return e == cudaSuccess ? GS_OK : cuda_fail(e, "waiting for in-flight scoring before an upload");
}

DevTables Ctx::tables() const {
  DevTables T;
  T.n_nodes = n_nodes;
  T.max_gpus = max_gpus;
  T.n_apps = n_apps;
  T.n_cols = n_cols;
  T.redis_reachable = redis_reachable;
  T.node_base = node_base;
  T.node_model = (const uint8_t*)node_model.p;
  T.node_ngpu = (const uint8_t*)node_ngpu.p;
  T.node_orig = (const uint32_t*)node_orig.p;
  T.res_new = (const uint32_t*)res_new.p;
  T.gpu_off = (const uint32_t*)gpu_off.p;
  T.res_app = (const uint32_t*)res_app.p;
  T.res_slo = (const float*)res_slo.p;
  T.conf = (const float*)conf.p;
  T.interf = (const float*)interf.p;
  T.interf_t = (const float*)interf_t.p;
  T.col_of = (const int32_t*)col_of.p;
  T.res_rec = (const float4*)res_rec.p;
  return T;
}

// derive the per-resident records once both table sets are present
int Ctx::ensure_prepared() {
  if (!have_cluster || !have_preds)
    return fail(GS_ERR_STATE, "scoring needs gs_upload_cluster and gs_upload_predictions first");
  if (!prep_dirty) return GS_OK;
  int rc = reserve(res_rec, sizeof(float4) * (size_t)(n_res ? n_res : 1), "res_rec");
  if (rc) return rc;
  DevTables T = tables();
  cudaError_t e = launch_resident_prep(T, (float4*)res_rec.p, stream);
  if (e != cudaSuccess) return cuda_fail(e, "resident_prep");
  // rare (once per upload): finish here so that callers' own streams need no ordering with ours
  e = cudaStreamSynchronize(stream);
  if (e != cudaSuccess) return cuda_fail(e, "resident_prep sync");
  prep_dirty = false;
  return GS_OK;
}

int Ctx::ensure_scratch(int n_pods) {
  uint32_t mc = max_chunks_for(n_pods, n_apps);
  const size_t nb = (size_t)n_buckets_for(n_apps) + 4;
  const size_t np = (size_t)(n_pods ? n_pods : 1);
  int rc;
  if ((rc = reserve(sc_counts, sizeof(uint32_t) * nb, "pod bucket counts"))) return rc;
  if ((rc = reserve(sc_cursor, sizeof(uint32_t) * nb, "pod bucket cursors"))) return rc;
  if ((rc = reserve(sc_nchunks, 256, "chunk counter"))) return rc;
  if ((rc = reserve(sc_chunks, sizeof(ChunkDesc) * (size_t)mc, "chunk descriptors"))) return rc;
  if ((rc = reserve(sc_sorted, sizeof(PodRec) * np, "sorted pods"))) return rc;
  if ((rc = reserve(sc_bucket, sizeof(uint32_t) * np, "pod buckets"))) return rc;
  if (dedupe) {
    const size_t hs = hash_size_for(n_pods);
    if ((rc = reserve(sc_rep, sizeof(uint32_t) * np, "class representatives"))) return rc;
    if ((rc = reserve(sc_hkeys, sizeof(unsigned long long) * hs, "class table keys"))) return rc;
    if ((rc = reserve(sc_hvals, sizeof(uint32_t) * hs, "class table values"))) return rc;
  }
  // residents-sum table: 24 bytes per (app bucket, GPU slot), if it fits the budget
  const size_t entries = (size_t)(n_apps + 1) * (size_t)n_nodes * (size_t)max_gpus;
  if (n_pods >= kStageTableMinPods && entries * 24 <= stage_budget) {
    if ((rc = reserve(sc_stage, entries * 24 + 1024, "residents-sum table"))) return rc;
  }
  return GS_OK;
}

PodScratch Ctx::scratch_view(int n_pods, bool use_dedupe) const {
  PodScratch sc;
  sc.counts = (uint32_t*)sc_counts.p;
  sc.cursor = (uint32_t*)sc_cursor.p;
  sc.n_chunks = (uint32_t*)sc_nchunks.p;
  sc.chunks = (ChunkDesc*)sc_chunks.p;
  sc.sorted = (PodRec*)sc_sorted.p;
  sc.bucket = (uint32_t*)sc_bucket.p;
  sc.max_chunks = max_chunks_for(n_pods, n_apps);
  sc.rep = nullptr;
  sc.hkeys = nullptr;
  sc.hvals = nullptr;
  sc.hash_size = 0;
  if (use_dedupe && dedupe && sc_rep.p) {
    sc.rep = (uint32_t*)sc_rep.p;
    sc.hkeys = (unsigned long long*)sc_hkeys.p;
    sc.hvals = (uint32_t*)sc_hvals.p;
    sc.hash_size = hash_size_for(n_pods);
  }
  sc.stage = StageA{nullptr, nullptr, nullptr, nullptr, 0};
  const size_t entries = (size_t)(n_apps + 1) * (size_t)n_nodes * (size_t)max_gpus;
  const size_t e8 = (entries + 31) / 32 * 32;  // keep every array 256-byte aligned
  if (n_pods >= kStageTableMinPods && entries * 24 <= stage_budget && e8 * 24 <= sc_stage.cap) {
    char* base = (char*)sc_stage.p;
    sc.stage.n0 = (double*)base;
    sc.stage.p0 = (double*)(base + e8 * 8);
    sc.stage.ci = (float*)(base + e8 * 16);
    sc.stage.cnt = (uint32_t*)(base + e8 * 20);
    sc.stage.stride = (size_t)n_nodes * (size_t)max_gpus;
  }
  return sc;
}

}  // namespace gs

using gs::Ctx;

#define GS_LOCK(ctx) std::lock_guard<std::mutex> _lk((ctx)->mu)
#define GS_CUDA(ctx, call)                                       \
  do {                                                           \
    cudaError_t _e = (call);                                     \
    if (_e != cudaSuccess) return (ctx)->cuda_fail(_e, #call);   \
  } while (0)

extern "C" {

uint64_t gs_launch_count(void) { return gs::launch_count(); }

const char* gs_version(void) { return "gpusched 0.1 sm_100a " __DATE__; }

int gs_create(int device, gs_ctx** out) {
  if (!out) return GS_ERR_INVALID;
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    cudaGetLastError();
    return GS_ERR_CUDA;  // no CPU fallback by design
  }
  if (device < 0 || device >= count) return GS_ERR_INVALID;
  gs_ctx* c = new (std::nothrow) gs_ctx();
  if (!c) return GS_ERR_NOMEM;
  c->device = device;
  if (const char* mb = getenv("GS_STAGE_TABLE_MB")) c->stage_budget = (size_t)strtoull(mb, nullptr, 10) << 20;
  if (const char* v = getenv("GS_NO_DEDUPE")) c->dedupe = atoi(v) == 0;
  if (const char* v = getenv("GS_LANES_NODES")) c->lanes_nodes = atoi(v) != 0;
  if (cudaSetDevice(device) != cudaSuccess ||
      cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreateWithFlags(&c->ev_last_use, cudaEventDisableTiming) != cudaSuccess) {
    if (c->stream) cudaStreamDestroy(c->stream);
    cudaGetLastError();
    delete c;
    return GS_ERR_CUDA;
  }
  *out = c;
  return GS_OK;
}

void gs_destroy(gs_ctx* c) {
  if (!c) return;
  {
    GS_LOCK(c);
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    gs::DevBuf* bufs[] = {&c->node_orig, &c->res_new, &c->node_model, &c->node_ngpu, &c->gpu_off,   &c->res_app,   &c->res_slo, &c->conf,
                          &c->interf,     &c->interf_t,  &c->sc_stage,  &c->col_of,    &c->res_rec,   &c->sc_counts, &c->sc_cursor,
                          &c->sc_nchunks, &c->sc_chunks, &c->sc_sorted, &c->sc_counts_all, &c->sc_bucket, &c->sc_rep, &c->sc_hkeys, &c->sc_hvals, &c->b_pod_app, &c->b_pod_slo,
                          &c->b_pod_self, &c->b_score,   &c->b_sel,     &c->b_extrema, &c->b_best,  &c->b_misc};
    for (gs::DevBuf* b : bufs) c->release(*b);
    for (auto& ev : c->ev_chunk)
      if (ev) cudaEventDestroy(ev);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    for (auto& ev : c->ev_prof)
      if (ev) cudaEventDestroy(ev);
    if (c->ev_last_use) cudaEventDestroy(c->ev_last_use);
    cudaStreamDestroy(c->stream);
  }
  delete c;
}

const char* gs_last_error(const gs_ctx* c) { return c ? c->err.c_str() : "null context"; }

int gs_upload_cluster(gs_ctx* c, const gs_cluster* cl, int32_t node_base) {
  if (!c) return GS_ERR_INVALID;
  GS_LOCK(c);
  if (!cl) return c->fail(GS_ERR_INVALID, "gs_upload_cluster: null cluster");
  if (cl->n_nodes > gs::kMaxNodesPerCtx)
    return c->fail(GS_ERR_INVALID, "gs_upload_cluster: %d nodes in one context; at most %d (shard the cluster)",
                   cl->n_nodes, gs::kMaxNodesPerCtx);
  if (cl->n_nodes < 0 || cl->max_gpus < 1 || cl->max_gpus > GS_MAX_GPUS_PER_NODE)
    return c->fail(GS_ERR_INVALID, "gs_upload_cluster: n_nodes=%d max_gpus=%d out of range (max_gpus 1..%d)",
                   cl->n_nodes, cl->max_gpus, GS_MAX_GPUS_PER_NODE);
  const size_t N = (size_t)cl->n_nodes, slots = N * (size_t)cl->max_gpus;
  if (N && (!cl->node_model || !cl->node_ngpu)) return c->fail(GS_ERR_INVALID, "gs_upload_cluster: null node table");
  if (!cl->gpu_off) return c->fail(GS_ERR_INVALID, "gs_upload_cluster: null gpu_off");
  if (cl->gpu_off[0] != 0) return c->fail(GS_ERR_INVALID, "gs_upload_cluster: gpu_off[0] must be 0");
  for (size_t s = 0; s < slots; ++s) {
    if (cl->gpu_off[s + 1] < cl->gpu_off[s])
      return c->fail(GS_ERR_INVALID, "gs_upload_cluster: gpu_off decreases at slot %zu", s);
    if (cl->gpu_off[s + 1] - cl->gpu_off[s] > (uint32_t)gs::kMaxResidentsPerSlot)
      return c->fail(GS_ERR_INVALID, "gs_upload_cluster: slot %zu holds %u residents; at most %d per GPU", s,
                     cl->gpu_off[s + 1] - cl->gpu_off[s], gs::kMaxResidentsPerSlot);
  }
  for (size_t n = 0; n < N; ++n) {
    if (cl->node_model[n] > GS_MODEL_V100)
      return c->fail(GS_ERR_INVALID, "gs_upload_cluster: node %zu has model %d", n, (int)cl->node_model[n]);
    if (cl->node_ngpu[n] > cl->max_gpus)
      return c->fail(GS_ERR_INVALID, "gs_upload_cluster: node %zu lists %d GPUs > max_gpus %d", n,
                     (int)cl->node_ngpu[n], cl->max_gpus);
    for (int g = cl->node_ngpu[n]; g < cl->max_gpus; ++g)
      if (cl->gpu_off[n * cl->max_gpus + g + 1] != cl->gpu_off[n * cl->max_gpus + g])
        return c->fail(GS_ERR_INVALID, "gs_upload_cluster: node %zu slot %d beyond node_ngpu holds residents", n, g);
  }
  const size_t R = cl->gpu_off[slots];
  if (R && (!cl->res_app || !cl->res_slo)) return c->fail(GS_ERR_INVALID, "gs_upload_cluster: null resident table");
  GS_CUDA(c, cudaSetDevice(c->device));
  if (int rcw = c->wait_last_use()) return rcw;  // asynchronous scoring on a caller stream may still read the tables
  // Internal node order.  Round 1 grouped the nodes of every 256-block by model so that the warps of
  // the one-thread-per-node kernel were homogeneous; the lanes-are-pods kernel evaluates nodes from
  // per-slice lists instead and wants consecutive internal nodes to be consecutive bytes of a result
  // row, so the order is now the caller's.  (node_orig / res_new stay, as identities, for the kernels
  // that were written against them.)
  std::vector<uint32_t> order(N), res_new(R ? R : 1);
  for (size_t n = 0; n < N; ++n) order[n] = (uint32_t)n;
  std::vector<uint8_t> p_model(N ? N : 1), p_ngpu(N ? N : 1);
  std::vector<uint32_t> p_off(slots + 1), p_app(R ? R : 1);
  std::vector<float> p_slo(R ? R : 1);
  const size_t G = (size_t)cl->max_gpus;
  p_off[0] = 0;
  for (size_t i = 0; i < N; ++i) {
    const size_t n = order[i];
    p_model[i] = cl->node_model[n];
    p_ngpu[i] = cl->node_ngpu[n];
    for (size_t g = 0; g < G; ++g) {
      const uint32_t lo = cl->gpu_off[n * G + g], hi = cl->gpu_off[n * G + g + 1];
      uint32_t w = p_off[i * G + g];
      for (uint32_t r = lo; r < hi; ++r, ++w) {
        p_app[w] = cl->res_app[r];
        p_slo[w] = cl->res_slo[r];
        res_new[r] = w;
      }
      p_off[i * G + g + 1] = w;
    }
  }
  int rc;
  if ((rc = c->upload(c->node_model, p_model.data(), N, "node_model"))) return rc;
  if ((rc = c->upload(c->node_ngpu, p_ngpu.data(), N, "node_ngpu"))) return rc;
  if ((rc = c->upload(c->gpu_off, p_off.data(), sizeof(uint32_t) * (slots + 1), "gpu_off"))) return rc;
  if ((rc = c->upload(c->res_app, p_app.data(), sizeof(uint32_t) * R, "res_app"))) return rc;
  if ((rc = c->upload(c->res_slo, p_slo.data(), sizeof(float) * R, "res_slo"))) return rc;
  if ((rc = c->upload(c->node_orig, order.data(), sizeof(uint32_t) * N, "node_orig"))) return rc;
  if ((rc = c->upload(c->res_new, res_new.data(), sizeof(uint32_t) * R, "res_new"))) return rc;
  GS_CUDA(c, cudaStreamSynchronize(c->stream));  // caller may free its buffers on return
  c->n_nodes = cl->n_nodes;
  c->max_gpus = cl->max_gpus;
  c->n_res = (int64_t)R;
  c->redis_reachable = cl->redis_reachable ? 1 : 0;
  c->node_base = node_base;
  c->have_cluster = true;
  c->prep_dirty = true;
  c->h_model.assign(p_model.begin(), p_model.begin() + N);
  c->h_ngpu.assign(p_ngpu.begin(), p_ngpu.begin() + N);
  c->h_off = p_off;
  c->h_app.assign(p_app.begin(), p_app.begin() + R);
  c->h_slo.assign(p_slo.begin(), p_slo.begin() + R);
  return GS_OK;
}

int gs_update_node(gs_ctx* c, int32_t node, int32_t n_gpus, const uint32_t* gpu_counts, const uint32_t* res_app,
                   const float* res_slo) {
  if (!c) return GS_ERR_INVALID;
  GS_LOCK(c);
  if (!c->have_cluster) return c->fail(GS_ERR_STATE, "gs_update_node before gs_upload_cluster");
  const size_t G = (size_t)c->max_gpus, N = (size_t)c->n_nodes;
  if (node < 0 || (size_t)node >= N || n_gpus < 0 || n_gpus > c->max_gpus || (n_gpus && !gpu_counts))
    return c->fail(GS_ERR_INVALID, "gs_update_node: node %d or GPU count %d out of range (%zu nodes, max_gpus %d)", node,
                   n_gpus, N, c->max_gpus);
  size_t n_new = 0;
  for (int g = 0; g < n_gpus; ++g) {
    if (gpu_counts[g] > (uint32_t)gs::kMaxResidentsPerSlot)
      return c->fail(GS_ERR_INVALID, "gs_update_node: GPU %d holds %u residents; at most %d", g, gpu_counts[g],
                     gs::kMaxResidentsPerSlot);
    n_new += gpu_counts[g];
  }
  if (n_new && (!res_app || !res_slo)) return c->fail(GS_ERR_INVALID, "gs_update_node: null resident table");
  GS_CUDA(c, cudaSetDevice(c->device));
  if (int rcw = c->wait_last_use()) return rcw;
  // splice the node's residents into the host mirror
  const size_t s0 = (size_t)node * G;
  const uint32_t lo = c->h_off[s0], hi = c->h_off[s0 + G];
  const long long delta = (long long)n_new - (long long)(hi - lo);
  c->h_app.erase(c->h_app.begin() + lo, c->h_app.begin() + hi);
  c->h_slo.erase(c->h_slo.begin() + lo, c->h_slo.begin() + hi);
  c->h_app.insert(c->h_app.begin() + lo, res_app, res_app + n_new);
  c->h_slo.insert(c->h_slo.begin() + lo, res_slo, res_slo + n_new);
  uint32_t w = lo;
  for (size_t g = 0; g < G; ++g) {
    w += g < (size_t)n_gpus ? gpu_counts[g] : 0u;
    c->h_off[s0 + g + 1] = w;
  }
  for (size_t s = s0 + G + 1; s < c->h_off.size(); ++s) c->h_off[s] = (uint32_t)((long long)c->h_off[s] + delta);
  c->h_ngpu[(size_t)node] = (uint8_t)n_gpus;
  const size_t R = c->h_app.size();
  // device: the node's GPU count, the offsets from this node on, the residents from this node on
  int rc;
  std::vector<uint32_t> ident(R ? R : 1);
  for (size_t r = 0; r < R; ++r) ident[r] = (uint32_t)r;
  if (sizeof(uint32_t) * R > c->res_app.cap || sizeof(float) * R > c->res_slo.cap || sizeof(uint32_t) * R > c->res_new.cap) {
    // grew beyond the buffers: whole arrays (reserve reallocates)
    if ((rc = c->upload(c->res_app, c->h_app.data(), sizeof(uint32_t) * R, "res_app"))) return rc;
    if ((rc = c->upload(c->res_slo, c->h_slo.data(), sizeof(float) * R, "res_slo"))) return rc;
    if ((rc = c->upload(c->res_new, ident.data(), sizeof(uint32_t) * R, "res_new"))) return rc;
  } else if (R > lo) {
    GS_CUDA(c, cudaMemcpyAsync((uint32_t*)c->res_app.p + lo, c->h_app.data() + lo, sizeof(uint32_t) * (R - lo),
                               cudaMemcpyHostToDevice, c->stream));
    GS_CUDA(c, cudaMemcpyAsync((float*)c->res_slo.p + lo, c->h_slo.data() + lo, sizeof(float) * (R - lo),
                               cudaMemcpyHostToDevice, c->stream));
    GS_CUDA(c, cudaMemcpyAsync((uint32_t*)c->res_new.p + lo, ident.data() + lo, sizeof(uint32_t) * (R - lo),
                               cudaMemcpyHostToDevice, c->stream));
  }
  GS_CUDA(c, cudaMemcpyAsync((uint32_t*)c->gpu_off.p + s0, c->h_off.data() + s0, sizeof(uint32_t) * (c->h_off.size() - s0),
                             cudaMemcpyHostToDevice, c->stream));
  GS_CUDA(c, cudaMemcpyAsync((uint8_t*)c->node_ngpu.p + node, c->h_ngpu.data() + node, 1, cudaMemcpyHostToDevice, c->stream));
  GS_CUDA(c, cudaStreamSynchronize(c->stream));
  c->n_res = (int64_t)R;
  c->prep_dirty = true;  // the per-resident records are rebuilt before the next scoring call
  return GS_OK;
}

int gs_upload_predictions(gs_ctx* c, const gs_predictions* pr) {
  if (!c) return GS_ERR_INVALID;
  GS_LOCK(c);
  if (!pr || pr->n_apps < 0 || pr->n_cols < 0 || !pr->col_of || (pr->n_apps && pr->n_cols && !pr->conf) ||
      (pr->n_apps && !pr->interf))
    return c->fail(GS_ERR_INVALID, "gs_upload_predictions: null table or negative size");
  for (int i = 0; i < 2 * (GS_MAX_GPUS_PER_NODE + 1); ++i)
    if (pr->col_of[i] >= pr->n_cols || pr->col_of[i] < -1)
      return c->fail(GS_ERR_INVALID, "gs_upload_predictions: col_of[%d]=%d outside [-1,%d)", i, pr->col_of[i],
                     pr->n_cols);
  GS_CUDA(c, cudaSetDevice(c->device));
  if (int rcw = c->wait_last_use()) return rcw;
  const size_t A = (size_t)pr->n_apps, K = (size_t)pr->n_cols;
  int rc;
  if ((rc = c->upload(c->conf, pr->conf, sizeof(float) * A * K, "conf"))) return rc;
  if ((rc = c->upload(c->interf, pr->interf, sizeof(float) * 2 * A * A, "interf"))) return rc;
  if ((rc = c->upload(c->col_of, pr->col_of, sizeof(int32_t) * 2 * (GS_MAX_GPUS_PER_NODE + 1), "col_of"))) return rc;
  if ((rc = c->reserve(c->interf_t, sizeof(float) * 2 * A * A + 16, "interf_t"))) return rc;
  if (A) {
    cudaError_t e = gs::launch_transpose_interf((const float*)c->interf.p, (float*)c->interf_t.p, pr->n_apps, c->stream);
    if (e != cudaSuccess) return c->cuda_fail(e, "transpose_interf");
  }
  GS_CUDA(c, cudaStreamSynchronize(c->stream));
  c->n_apps = pr->n_apps;
  c->n_cols = pr->n_cols;
  c->have_preds = true;
  c->prep_dirty = true;
  return GS_OK;
}

// shared by the host- and device-pointer entry points; ctx->mu held
static int score_raw_locked(gs_ctx* c, const uint32_t* d_app, const float* d_slo, const int32_t* d_self, int32_t P,
                            uint32_t flags, uint8_t* d_score, uint8_t* d_sel, uint64_t* d_hi, uint64_t* d_neg,
                            cudaStream_t s) {
  int rc = c->ensure_prepared();
  if (rc) return rc;
  GS_CUDA(c, cudaMemsetAsync(d_hi, 0, sizeof(uint64_t) * (size_t)P, s));
  GS_CUDA(c, cudaMemsetAsync(d_neg, 0, sizeof(uint64_t) * (size_t)P, s));
  gs::DevTables T = c->tables();
  const bool fast = !(flags & GS_SCORE_DIRECT) && d_self == nullptr && c->max_gpus <= gs::kFastMaxGpus;
  cudaError_t e;
  if (fast && (rc = c->ensure_scratch(P))) return rc;
  if (c->profiling) {
    for (auto& ev : c->ev_prof)
      if (!ev) GS_CUDA(c, cudaEventCreate(&ev));
    c->ev_valid = false;
  }
  const bool lanes_nodes = c->lanes_nodes || (flags & GS_SCORE_LANES_NODES);
  if (fast) {
    gs::PodScratch sc = c->scratch_view(P, !(flags & GS_SCORE_NO_DEDUPE));
    e = gs::launch_score_fast(T, d_app, d_slo, P, sc, d_score, d_sel, (unsigned long long*)d_hi,
                              (unsigned long long*)d_neg, s, c->profiling ? c->ev_prof : nullptr, lanes_nodes);
  } else {
    if (c->profiling) {
      GS_CUDA(c, cudaEventRecord(c->ev_prof[0], s));
      GS_CUDA(c, cudaEventRecord(c->ev_prof[1], s));
    }
    e = gs::launch_score_direct(T, d_app, d_slo, d_self, P, d_score, d_sel, (unsigned long long*)d_hi,
                                (unsigned long long*)d_neg, s);
    if (c->profiling) GS_CUDA(c, cudaEventRecord(c->ev_prof[2], s));
  }
  if (e != cudaSuccess) return c->cuda_fail(e, fast ? "score_fast" : "score_direct");
  if (c->profiling) c->ev_valid = true;
  return c->note_use(s);
}

int gs_score_raw_dev(gs_ctx* c, const uint32_t* d_pod_app, const float* d_pod_slo, const int32_t* d_pod_self,
                     int32_t n_pods, uint32_t flags, uint8_t* d_score, uint8_t* d_sel, uint64_t* d_hi_arg,
                     uint64_t* d_neg_lo, void* stream) {
  if (!c) return GS_ERR_INVALID;
  GS_LOCK(c);
  if (n_pods < 0 || n_pods > gs::kMaxPods ||
      (n_pods && (!d_pod_app || !d_pod_slo || !d_score || !d_sel || !d_hi_arg || !d_neg_lo)))
    return c->fail(GS_ERR_INVALID, "gs_score_raw_dev: null pointer or n_pods outside [0, 2^29]");
  GS_CUDA(c, cudaSetDevice(c->device));
  if (n_pods == 0) return GS_OK;
  return score_raw_locked(c, d_pod_app, d_pod_slo, d_pod_self, n_pods, flags, d_score, d_sel, d_hi_arg, d_neg_lo,
                          stream ? (cudaStream_t)stream : c->stream);
}

int gs_normalize_dev(gs_ctx* c, int32_t n_pods, uint8_t* d_score, const uint64_t* d_hi_arg, const uint64_t* d_neg_lo,
                     int32_t* d_best, void* stream) {
  if (!c) return GS_ERR_INVALID;
  GS_LOCK(c);
  if (n_pods < 0 || (n_pods && (!d_score || !d_hi_arg || !d_neg_lo)))
    return c->fail(GS_ERR_INVALID, "gs_normalize_dev: null pointer or negative n_pods");
  if (!c->have_cluster) return c->fail(GS_ERR_STATE, "gs_normalize_dev before gs_upload_cluster");
  GS_CUDA(c, cudaSetDevice(c->device));
  cudaError_t e = gs::launch_normalize(n_pods, c->n_nodes, 1, d_score, (const unsigned long long*)d_hi_arg,
                                       (const unsigned long long*)d_neg_lo, d_best,
                                       stream ? (cudaStream_t)stream : c->stream);
  if (e != cudaSuccess) return c->cuda_fail(e, "normalize");
  return c->note_use(stream ? (cudaStream_t)stream : c->stream);
}

int gs_best_dev(gs_ctx* c, int32_t n_pods, const uint64_t* d_hi_arg, const uint64_t* d_neg_lo, int32_t* d_best,
                void* stream) {
  if (!c) return GS_ERR_INVALID;
  GS_LOCK(c);
  if (n_pods < 0 || (n_pods && (!d_hi_arg || !d_neg_lo || !d_best)))
    return c->fail(GS_ERR_INVALID, "gs_best_dev: null pointer or negative n_pods");
  GS_CUDA(c, cudaSetDevice(c->device));
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
  cudaError_t e = gs::launch_normalize(n_pods, c->n_nodes, 0, nullptr, (const unsigned long long*)d_hi_arg,
                                       (const unsigned long long*)d_neg_lo, d_best, s);
  if (e != cudaSuccess) return c->cuda_fail(e, "best node");
  return c->note_use(s);
}

int gs_score_batch(gs_ctx* c, const uint32_t* pod_app, const float* pod_slo, const int32_t* pod_self, int32_t P,
                   uint32_t flags, uint8_t* out_score, uint8_t* out_sel, int32_t* out_best) {
  if (!c) return GS_ERR_INVALID;
  GS_LOCK(c);
  if (P < 0 || P > gs::kMaxPods || (P && (!pod_app || !pod_slo)))
    return c->fail(GS_ERR_INVALID, "gs_score_batch: null pod table or n_pods outside [0, 2^29]");
  if (!c->have_cluster || !c->have_preds)
    return c->fail(GS_ERR_STATE, "scoring needs gs_upload_cluster and gs_upload_predictions first");
  if (P == 0) return GS_OK;
  GS_CUDA(c, cudaSetDevice(c->device));
  const size_t N = (size_t)c->n_nodes, PN = (size_t)P * N;
  // a self table that excludes nobody is the common case: treat as absent
  if (pod_self) {
    bool any = false;
    for (int32_t p = 0; p < P && !any; ++p) any = pod_self[p] >= 0;
    if (!any) pod_self = nullptr;
  }
  if (pod_self)
    for (int32_t p = 0; p < P; ++p)
      if (pod_self[p] >= c->n_res) return c->fail(GS_ERR_INVALID, "gs_score_batch: pod_self[%d] beyond residents", p);
  int rc;
  cudaStream_t s = c->stream;
  if ((rc = c->upload(c->b_pod_app, pod_app, sizeof(uint32_t) * (size_t)P, "pod_app"))) return rc;
  if ((rc = c->upload(c->b_pod_slo, pod_slo, sizeof(float) * (size_t)P, "pod_slo"))) return rc;
  if (pod_self && (rc = c->upload(c->b_pod_self, pod_self, sizeof(int32_t) * (size_t)P, "pod_self"))) return rc;
  if ((rc = c->reserve(c->b_score, PN ? PN : 1, "scores"))) return rc;
  if ((rc = c->reserve(c->b_sel, PN ? PN : 1, "selections"))) return rc;
  if ((rc = c->reserve(c->b_extrema, sizeof(uint64_t) * 2 * (size_t)P, "extrema"))) return rc;
  if ((rc = c->reserve(c->b_best, sizeof(int32_t) * (size_t)P, "best"))) return rc;
  uint64_t* d_hi = (uint64_t*)c->b_extrema.p;
  uint64_t* d_neg = d_hi + P;
  const int rescale = (flags & GS_SCORE_NORMALIZE) ? 1 : 0;

  // Large batches through the restructured path: score the pods in a few slices and copy each
  // slice's rows to the host on a second stream while the next slice is being scored (the
  // results are 2 bytes per (pod,node): the PCIe copy, not the arithmetic, bounds this call).
  const bool fast = !(flags & GS_SCORE_DIRECT) && !pod_self && c->max_gpus <= gs::kFastMaxGpus;
  int n_slices = 1;
  if (fast && N && out_score && out_sel && P >= 8192 && PN * 2 >= ((size_t)64 << 20)) {
    n_slices = (int)std::min<size_t>(6, std::max<size_t>(2, PN * 2 / ((size_t)80 << 20)));  // measured: 6 slices best at 500 MB
    if (const char* ov = getenv("GS_SLICES")) n_slices = std::max(1, std::min(8, atoi(ov)));
  }
  if (n_slices > 1) {
    if ((rc = c->ensure_prepared())) return rc;
    if ((rc = c->ensure_scratch(P))) return rc;
    if ((rc = c->reserve(c->sc_counts_all, sizeof(uint32_t) * (size_t)(gs::n_buckets_for(c->n_apps) + 4), "bucket counts (batch)"))) return rc;
    if (!c->copy_stream) {
      GS_CUDA(c, cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
      for (auto& ev : c->ev_chunk) GS_CUDA(c, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    }
    gs::DevTables T = c->tables();
    const uint32_t* d_app = (const uint32_t*)c->b_pod_app.p;
    const float* d_slo = (const float*)c->b_pod_slo.p;
    uint8_t* d_score = (uint8_t*)c->b_score.p;
    uint8_t* d_sel = (uint8_t*)c->b_sel.p;
    int32_t* d_best = (int32_t*)c->b_best.p;
    GS_CUDA(c, cudaMemsetAsync(d_hi, 0, sizeof(uint64_t) * 2 * (size_t)P, s));
    // the residents' sums once, for every app bucket present anywhere in the batch
    gs::PodScratch whole = c->scratch_view(P);
    cudaError_t e = gs::launch_pod_histogram(d_app, P, c->n_apps, (uint32_t*)c->sc_counts_all.p, s);
    const bool ln = c->lanes_nodes || (flags & GS_SCORE_LANES_NODES);
    if (e == cudaSuccess) e = gs::launch_stage_any(T, (const uint32_t*)c->sc_counts_all.p, whole, s, ln);
    if (e != cudaSuccess) return c->cuda_fail(e, "stage_a (batch)");
    const int per = (P + n_slices - 1) / n_slices;
    for (int k = 0; k < n_slices; ++k) {
      const int p0 = k * per, np = std::min(per, P - p0);
      if (np <= 0) break;
      gs::PodScratch sc = c->scratch_view(np, !(flags & GS_SCORE_NO_DEDUPE));
      sc.stage = whole.stage;  // the table covers the batch's apps (built above), whatever the slice size
      const size_t row0 = (size_t)p0 * N;
      e = gs::launch_pod_bucketing(T, d_app + p0, d_slo + p0, np, sc, s);
      if (e == cudaSuccess)
        e = gs::launch_score_any(T, sc, d_score + row0, d_sel + row0, (unsigned long long*)(d_hi + p0),
                                 (unsigned long long*)(d_neg + p0), s, ln);
      if (e == cudaSuccess)
        e = gs::launch_expand_classes(T, sc.rep, np, d_score + row0, d_sel + row0, (unsigned long long*)(d_hi + p0),
                                      (unsigned long long*)(d_neg + p0), s);
      if (e == cudaSuccess)
        e = gs::launch_normalize(np, c->n_nodes, rescale, d_score + row0, (const unsigned long long*)(d_hi + p0),
                                 (const unsigned long long*)(d_neg + p0), d_best + p0, s);
      if (e != cudaSuccess) return c->cuda_fail(e, "score slice");
      GS_CUDA(c, cudaEventRecord(c->ev_chunk[k], s));
      GS_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->ev_chunk[k], 0));
      GS_CUDA(c, cudaMemcpyAsync(out_score + row0, d_score + row0, (size_t)np * N, cudaMemcpyDeviceToHost, c->copy_stream));
      GS_CUDA(c, cudaMemcpyAsync(out_sel + row0, d_sel + row0, (size_t)np * N, cudaMemcpyDeviceToHost, c->copy_stream));
    }
    if (out_best)
      GS_CUDA(c, cudaMemcpyAsync(out_best, d_best, sizeof(int32_t) * (size_t)P, cudaMemcpyDeviceToHost, s));
    GS_CUDA(c, cudaStreamSynchronize(s));
    GS_CUDA(c, cudaStreamSynchronize(c->copy_stream));
    return GS_OK;
  }

  rc = score_raw_locked(c, (const uint32_t*)c->b_pod_app.p, (const float*)c->b_pod_slo.p,
                        pod_self ? (const int32_t*)c->b_pod_self.p : nullptr, P, flags, (uint8_t*)c->b_score.p,
                        (uint8_t*)c->b_sel.p, d_hi, d_neg, s);
  if (rc) return rc;
  if (N) {
    cudaError_t e = gs::launch_normalize(P, c->n_nodes, rescale, (uint8_t*)c->b_score.p,
                                         (const unsigned long long*)d_hi, (const unsigned long long*)d_neg,
                                         (int32_t*)c->b_best.p, s);
    if (e != cudaSuccess) return c->cuda_fail(e, "normalize");
  }
  if (out_score && PN) GS_CUDA(c, cudaMemcpyAsync(out_score, c->b_score.p, PN, cudaMemcpyDeviceToHost, s));
  if (out_sel && PN) GS_CUDA(c, cudaMemcpyAsync(out_sel, c->b_sel.p, PN, cudaMemcpyDeviceToHost, s));
  if (out_best) {
    if (N)
      GS_CUDA(c, cudaMemcpyAsync(out_best, c->b_best.p, sizeof(int32_t) * (size_t)P, cudaMemcpyDeviceToHost, s));
    else
      for (int32_t p = 0; p < P; ++p) out_best[p] = -1;
  }
  GS_CUDA(c, cudaStreamSynchronize(s));
  return GS_OK;
}

int gs_score_decisions(gs_ctx* c, const uint32_t* pod_app, const float* pod_slo, const int32_t* pod_self, int32_t P,
                       uint32_t flags, gs_decision* out) {
  static_assert(sizeof(gs_decision) == sizeof(gs::Decision) && sizeof(gs_decision) == 8, "decision layout");
  if (!c) return GS_ERR_INVALID;
  GS_LOCK(c);
  if (P < 0 || P > gs::kMaxPods || (P && (!pod_app || !pod_slo || !out)))
    return c->fail(GS_ERR_INVALID, "gs_score_decisions: null pointer or n_pods outside [0, 2^29]");
  if (!c->have_cluster || !c->have_preds)
    return c->fail(GS_ERR_STATE, "scoring needs gs_upload_cluster and gs_upload_predictions first");
  if (P == 0) return GS_OK;
  GS_CUDA(c, cudaSetDevice(c->device));
  const size_t N = (size_t)c->n_nodes, PN = (size_t)P * N;
  if (pod_self) {
    bool any = false;
    for (int32_t p = 0; p < P && !any; ++p) any = pod_self[p] >= 0;
    if (!any) pod_self = nullptr;
  }
  if (pod_self)
    for (int32_t p = 0; p < P; ++p)
      if (pod_self[p] >= c->n_res) return c->fail(GS_ERR_INVALID, "gs_score_decisions: pod_self[%d] beyond residents", p);
  int rc;
  cudaStream_t s = c->stream;
  if ((rc = c->upload(c->b_pod_app, pod_app, sizeof(uint32_t) * (size_t)P, "pod_app"))) return rc;
  if ((rc = c->upload(c->b_pod_slo, pod_slo, sizeof(float) * (size_t)P, "pod_slo"))) return rc;
  if (pod_self && (rc = c->upload(c->b_pod_self, pod_self, sizeof(int32_t) * (size_t)P, "pod_self"))) return rc;
  if ((rc = c->reserve(c->b_score, PN ? PN : 1, "scores"))) return rc;
  if ((rc = c->reserve(c->b_sel, PN ? PN : 1, "selections"))) return rc;
  if ((rc = c->reserve(c->b_extrema, sizeof(uint64_t) * 2 * (size_t)P, "extrema"))) return rc;
  if ((rc = c->reserve(c->b_misc, sizeof(gs_decision) * (size_t)P, "decisions"))) return rc;
  uint64_t* d_hi = (uint64_t*)c->b_extrema.p;
  uint64_t* d_neg = d_hi + P;
  rc = score_raw_locked(c, (const uint32_t*)c->b_pod_app.p, (const float*)c->b_pod_slo.p,
                        pod_self ? (const int32_t*)c->b_pod_self.p : nullptr, P, flags, (uint8_t*)c->b_score.p,
                        (uint8_t*)c->b_sel.p, d_hi, d_neg, s);
  if (rc) return rc;
  gs::DevTables T = c->tables();
  cudaError_t e = gs::launch_decisions(T, (const uint32_t*)c->b_pod_app.p, (const float*)c->b_pod_slo.p, P,
                                       (const uint8_t*)c->b_sel.p, (const unsigned long long*)d_hi,
                                       (const unsigned long long*)d_neg, (gs::Decision*)c->b_misc.p, s);
  if (e != cudaSuccess) return c->cuda_fail(e, "decisions");
  GS_CUDA(c, cudaMemcpyAsync(out, c->b_misc.p, sizeof(gs_decision) * (size_t)P, cudaMemcpyDeviceToHost, s));
  GS_CUDA(c, cudaStreamSynchronize(s));
  return GS_OK;
}

int gs_decisions_dev(gs_ctx* c, const uint32_t* d_pod_app, const float* d_pod_slo, int32_t n_pods, const uint8_t* d_sel,
                     const uint64_t* d_hi_arg, const uint64_t* d_neg_lo, gs_decision* d_out, void* stream) {
  if (!c) return GS_ERR_INVALID;
  GS_LOCK(c);
  if (n_pods < 0 || (n_pods && (!d_pod_app || !d_pod_slo || !d_sel || !d_hi_arg || !d_neg_lo || !d_out)))
    return c->fail(GS_ERR_INVALID, "gs_decisions_dev: null pointer or negative n_pods");
  if (!c->have_cluster || !c->have_preds)
    return c->fail(GS_ERR_STATE, "gs_decisions_dev needs gs_upload_cluster and gs_upload_predictions first");
  GS_CUDA(c, cudaSetDevice(c->device));
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
  gs::DevTables T = c->tables();
  cudaError_t e = gs::launch_decisions(T, d_pod_app, d_pod_slo, n_pods, d_sel, (const unsigned long long*)d_hi_arg,
                                       (const unsigned long long*)d_neg_lo, (gs::Decision*)d_out, s);
  if (e != cudaSuccess) return c->cuda_fail(e, "decisions");
  return c->note_use(s);
}

int gs_normalize_scores(gs_ctx* c, int64_t* scores, int32_t n) {
  if (!c) return GS_ERR_INVALID;
  GS_LOCK(c);
  if (n < 0 || (n && !scores)) return c->fail(GS_ERR_INVALID, "gs_normalize_scores: null list");
  if (n == 0) return GS_OK;
  GS_CUDA(c, cudaSetDevice(c->device));
  int rc;
  if ((rc = c->upload(c->b_misc, scores, sizeof(int64_t) * (size_t)n, "score list"))) return rc;
  cudaError_t e = gs::launch_normalize_i64((long long*)c->b_misc.p, n, c->stream);
  if (e != cudaSuccess) return c->cuda_fail(e, "normalize_i64");
  GS_CUDA(c, cudaMemcpyAsync(scores, c->b_misc.p, sizeof(int64_t) * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
  GS_CUDA(c, cudaStreamSynchronize(c->stream));
  return GS_OK;
}

int gs_set_profiling(gs_ctx* c, int enabled) {
  if (!c) return GS_ERR_INVALID;
  GS_LOCK(c);
  c->profiling = enabled != 0;
  c->ev_valid = false;
  return GS_OK;
}

int gs_last_score_kernel_ms(gs_ctx* c, float* out_ms) {
  if (!c) return GS_ERR_INVALID;
  GS_LOCK(c);
  if (!out_ms) return c->fail(GS_ERR_INVALID, "gs_last_score_kernel_ms: null output");
  if (!c->profiling || !c->ev_valid) return c->fail(GS_ERR_STATE, "no profiled scoring call to report");
  GS_CUDA(c, cudaSetDevice(c->device));
  GS_CUDA(c, cudaEventSynchronize(c->ev_prof[2]));
  GS_CUDA(c, cudaEventElapsedTime(out_ms, c->ev_prof[1], c->ev_prof[2]));
  return GS_OK;
}

int gs_last_stage_a_ms(gs_ctx* c, float* out_ms) {
  if (!c) return GS_ERR_INVALID;
  GS_LOCK(c);
  if (!out_ms) return c->fail(GS_ERR_INVALID, "gs_last_stage_a_ms: null output");
  if (!c->profiling || !c->ev_valid) return c->fail(GS_ERR_STATE, "no profiled scoring call to report");
  GS_CUDA(c, cudaSetDevice(c->device));
  GS_CUDA(c, cudaEventSynchronize(c->ev_prof[1]));
  GS_CUDA(c, cudaEventElapsedTime(out_ms, c->ev_prof[0], c->ev_prof[1]));
  return GS_OK;
}

int gs_selftest(gs_ctx* c, uint64_t seed, int32_t n_million, uint64_t* out_mismatches) {
  if (!c) return GS_ERR_INVALID;
  GS_LOCK(c);
  if (!out_mismatches || n_million < 1) return c->fail(GS_ERR_INVALID, "gs_selftest: bad argument");
  GS_CUDA(c, cudaSetDevice(c->device));
  int rc;
  if ((rc = c->reserve(c->b_misc, 256, "selftest counter"))) return rc;
  GS_CUDA(c, cudaMemsetAsync(c->b_misc.p, 0, sizeof(unsigned long long), c->stream));
  // 3907 blocks x 256 threads = 1 000 192 threads; iters operands each
  cudaError_t e = gs::launch_selftest(seed, 3907, n_million, (unsigned long long*)c->b_misc.p, c->stream);
  if (e != cudaSuccess) return c->cuda_fail(e, "selftest");
  GS_CUDA(c, cudaMemcpyAsync(out_mismatches, c->b_misc.p, sizeof(uint64_t), cudaMemcpyDeviceToHost, c->stream));
  GS_CUDA(c, cudaStreamSynchronize(c->stream));
  return GS_OK;
}

int gs_selftest_filter(gs_ctx* c, uint64_t seed, int32_t n_million, double* out_max_dev, double* out_delta,
                       uint64_t* out_wrong_floors, uint64_t* out_in_band, uint64_t* out_evaluated) {
  if (!c) return GS_ERR_INVALID;
  GS_LOCK(c);
  if (n_million < 1 || !out_max_dev || !out_delta || !out_wrong_floors || !out_in_band || !out_evaluated)
    return c->fail(GS_ERR_INVALID, "gs_selftest_filter: bad argument");
  GS_CUDA(c, cudaSetDevice(c->device));
  int rc;
  if ((rc = c->reserve(c->b_misc, 256, "selftest counters"))) return rc;
  GS_CUDA(c, cudaMemsetAsync(c->b_misc.p, 0, 4 * sizeof(unsigned long long), c->stream));
  cudaError_t e = gs::launch_selftest_filter(seed, 3907, n_million, (unsigned long long*)c->b_misc.p, c->stream);
  if (e != cudaSuccess) return c->cuda_fail(e, "selftest_filter");
  unsigned long long h[4];
  GS_CUDA(c, cudaMemcpyAsync(h, c->b_misc.p, sizeof h, cudaMemcpyDeviceToHost, c->stream));
  GS_CUDA(c, cudaStreamSynchronize(c->stream));
  memcpy(out_max_dev, &h[0], sizeof(double));
  *out_delta = gs::filter_delta();
  *out_wrong_floors = h[1];
  *out_in_band = h[2];
  *out_evaluated = h[3];
  return GS_OK;
}

int gs_select_mig(gs_ctx* c, const uint32_t* pod_app, const float* pod_slo, int32_t P, uint8_t* out) {
  if (!c) return GS_ERR_INVALID;
  GS_LOCK(c);
  if (P < 0 || (P && (!pod_app || !pod_slo || !out))) return c->fail(GS_ERR_INVALID, "gs_select_mig: null pointer");
  if (!c->have_preds) return c->fail(GS_ERR_STATE, "gs_select_mig before gs_upload_predictions");
  if (P == 0) return GS_OK;
  GS_CUDA(c, cudaSetDevice(c->device));
  int rc;
  if ((rc = c->upload(c->b_pod_app, pod_app, sizeof(uint32_t) * (size_t)P, "pod_app"))) return rc;
  if ((rc = c->upload(c->b_pod_slo, pod_slo, sizeof(float) * (size_t)P, "pod_slo"))) return rc;
  if ((rc = c->reserve(c->b_misc, (size_t)P, "mig choice"))) return rc;
  gs::DevTables T = c->tables();
  cudaError_t e = gs::launch_select_mig(T, (const uint32_t*)c->b_pod_app.p, (const float*)c->b_pod_slo.p, P,
                                        (uint8_t*)c->b_misc.p, c->stream);
  if (e != cudaSuccess) return c->cuda_fail(e, "select_mig");
  GS_CUDA(c, cudaMemcpyAsync(out, c->b_misc.p, (size_t)P, cudaMemcpyDeviceToHost, c->stream));
  GS_CUDA(c, cudaStreamSynchronize(c->stream));
  return GS_OK;
}

}  // extern "C"

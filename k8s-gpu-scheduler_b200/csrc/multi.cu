// multi.cu — node-sharded scoring over several GPUs of one box, behind the C ABI (gs_mctx).
//
// Scores of different nodes are independent (Logic reads only its own node's UUIDs and residents,
// pkg/plugins/gpu_plugin/gpu_plugins.go:536-544), so the nodes are block-partitioned over the
// devices and pods / prediction tables are replicated.  The path's one exchange step:
// NormalizeScore needs every pod's minimum and maximum over ALL nodes (:818-827) and host selection
// the argmax.  Both are kept as max-reductions (see gs_score_raw_dev), so the exchange is:
//
//   device 0 holds the global extrema G[2P], zeroed at the start of the call;
//   every device, after its scoring kernels, runs merge_extrema_kernel: atomicMax of its local
//   extrema straight into G through the peer mapping (NVLink / NVSwitch) — 16 bytes per pod;
//   every device waits for all merges (events) and normalises its own columns reading G over the
//   peer mapping; each device's column block goes to the host with one strided copy.
//
// One host thread drives all devices asynchronously; no torch, no NCCL: a Go caller
// (cmd/scheduler/main.go:20-22 registers the plugin in a single process) can use all GPUs of the box.
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/gpusched.h"
#include "context.h"

namespace {

struct Buf {
  void* p = nullptr;
  size_t cap = 0;
};

struct Dev {
  int device = 0;
  gs_ctx* ctx = nullptr;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev_merged = nullptr;
  int node_begin = 0, node_end = 0;
  Buf app, slo, score, sel, ext, best, dec;
};

}  // namespace

struct gs_mctx {
  std::mutex mu;
  std::string err;
  std::vector<Dev> devs;
  cudaEvent_t ev_zero = nullptr;  // device 0: global extrema cleared
  Buf global_ext;                 // device 0: uint64[2P]
  int n_nodes = 0;
  bool have_cluster = false, have_preds = false;

  int fail(int code, const char* fmt, ...) __attribute__((format(printf, 3, 4))) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    err = buf;
    return code;
  }
  int cuda_fail(cudaError_t e, const char* what) {
    return fail(GS_ERR_CUDA, "%s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
  }
  int reserve(Buf& b, size_t bytes, const char* what) {  // the owning device must be current
    if (bytes <= b.cap) return GS_OK;
    if (b.p) cudaFree(b.p);
    b.p = nullptr;
    b.cap = 0;
    const size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&b.p, want);
    if (e != cudaSuccess) {
      cudaGetLastError();
      return fail(GS_ERR_NOMEM, "cudaMalloc(%zu bytes) for %s: %s", want, what, cudaGetErrorString(e));
    }
    b.cap = want;
    return GS_OK;
  }
};

#define GSM_CUDA(m, call)                                      \
  do {                                                         \
    cudaError_t _e = (call);                                   \
    if (_e != cudaSuccess) return (m)->cuda_fail(_e, #call);   \
  } while (0)

namespace {

// G[i] = max(G[i], local[i]) over the peer mapping; both buffers hold 2 * n_pods values
__global__ void merge_extrema_kernel(const unsigned long long* __restrict__ local, unsigned long long* global_ext,
                                     int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long v = local[i];
  if (v) atomicMax_system(&global_ext[i], v);
}

void free_dev(Dev& d) {
  cudaSetDevice(d.device);
  for (Buf* b : {&d.app, &d.slo, &d.score, &d.sel, &d.ext, &d.best, &d.dec})
    if (b->p) cudaFree(b->p);
  if (d.ev_merged) cudaEventDestroy(d.ev_merged);
  if (d.stream) cudaStreamDestroy(d.stream);
  if (d.ctx) gs_destroy(d.ctx);
}

}  // namespace

extern "C" {

int gs_create_multi(const int* devices, int n_devices, gs_mctx** out) {
  if (!out) return GS_ERR_INVALID;
  *out = nullptr;
  if (!devices || n_devices < 1 || n_devices > 64) return GS_ERR_INVALID;
  gs_mctx* m = new (std::nothrow) gs_mctx();
  if (!m) return GS_ERR_NOMEM;
  m->devs.resize((size_t)n_devices);
  int rc = GS_OK;
  for (int i = 0; i < n_devices && rc == GS_OK; ++i) {
    Dev& d = m->devs[(size_t)i];
    d.device = devices[i];
    for (int k = 0; k < i; ++k)
      if (devices[k] == devices[i]) rc = GS_ERR_INVALID;  // every device once
    if (rc == GS_OK) rc = gs_create(d.device, &d.ctx);
    if (rc == GS_OK && (cudaSetDevice(d.device) != cudaSuccess ||
                        cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking) != cudaSuccess ||
                        cudaEventCreateWithFlags(&d.ev_merged, cudaEventDisableTiming) != cudaSuccess))
      rc = GS_ERR_CUDA;
    // every device reaches device 0's memory (the global extrema)
    if (rc == GS_OK && i > 0) {
      int can = 0;
      if (cudaDeviceCanAccessPeer(&can, d.device, devices[0]) != cudaSuccess || !can) {
        rc = GS_ERR_UNSUPPORTED;
      } else {
        const cudaError_t e = cudaDeviceEnablePeerAccess(devices[0], 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) rc = GS_ERR_CUDA;
        cudaGetLastError();
      }
    }
  }
  if (rc == GS_OK && (cudaSetDevice(devices[0]) != cudaSuccess ||
                      cudaEventCreateWithFlags(&m->ev_zero, cudaEventDisableTiming) != cudaSuccess))
    rc = GS_ERR_CUDA;
  if (rc != GS_OK) {
    cudaGetLastError();
    for (Dev& d : m->devs) free_dev(d);
    delete m;
    return rc;
  }
  *out = m;
  return GS_OK;
}

void gs_destroy_multi(gs_mctx* m) {
  if (!m) return;
  {
    std::lock_guard<std::mutex> lk(m->mu);
    for (Dev& d : m->devs) {
      cudaSetDevice(d.device);
      cudaStreamSynchronize(d.stream);
    }
    cudaSetDevice(m->devs[0].device);
    if (m->global_ext.p) cudaFree(m->global_ext.p);
    if (m->ev_zero) cudaEventDestroy(m->ev_zero);
    for (Dev& d : m->devs) free_dev(d);
  }
  delete m;
}

const char* gs_multi_last_error(const gs_mctx* m) { return m ? m->err.c_str() : "null context"; }
int gs_multi_device_count(const gs_mctx* m) { return m ? (int)m->devs.size() : 0; }

int gs_mupload_predictions(gs_mctx* m, const gs_predictions* preds) {
  if (!m) return GS_ERR_INVALID;
  std::lock_guard<std::mutex> lk(m->mu);
  for (Dev& d : m->devs) {
    const int rc = gs_upload_predictions(d.ctx, preds);
    if (rc != GS_OK) return m->fail(rc, "device %d: %s", d.device, gs_last_error(d.ctx));
  }
  m->have_preds = true;
  return GS_OK;
}

// The node axis is block-partitioned: the first n_nodes % n_devices devices get one node more.
int gs_mupload_cluster(gs_mctx* m, const gs_cluster* cl) {
  if (!m) return GS_ERR_INVALID;
  std::lock_guard<std::mutex> lk(m->mu);
  if (!cl || cl->n_nodes < 0 || cl->max_gpus < 1 || !cl->gpu_off)
    return m->fail(GS_ERR_INVALID, "gs_mupload_cluster: null cluster or bad sizes");
  const int W = (int)m->devs.size(), N = cl->n_nodes, G = cl->max_gpus;
  const int base = N / W, extra = N % W;
  for (int i = 0; i < W; ++i) {
    Dev& d = m->devs[(size_t)i];
    d.node_begin = i * base + std::min(i, extra);
    d.node_end = d.node_begin + base + (i < extra ? 1 : 0);
    const int nl = d.node_end - d.node_begin;
    const uint32_t lo = cl->gpu_off[(size_t)d.node_begin * G], hi = cl->gpu_off[(size_t)d.node_end * G];
    std::vector<uint32_t> off((size_t)nl * G + 1);
    for (size_t s = 0; s < off.size(); ++s) off[s] = cl->gpu_off[(size_t)d.node_begin * G + s] - lo;
    gs_cluster sh = *cl;
    sh.n_nodes = nl;
    sh.node_model = cl->node_model ? cl->node_model + d.node_begin : nullptr;
    sh.node_ngpu = cl->node_ngpu ? cl->node_ngpu + d.node_begin : nullptr;
    sh.gpu_off = off.data();
    sh.res_app = (hi > lo && cl->res_app) ? cl->res_app + lo : cl->res_app;
    sh.res_slo = (hi > lo && cl->res_slo) ? cl->res_slo + lo : cl->res_slo;
    const int rc = gs_upload_cluster(d.ctx, &sh, d.node_begin);
    if (rc != GS_OK) return m->fail(rc, "device %d: %s", d.device, gs_last_error(d.ctx));
  }
  m->n_nodes = N;
  m->have_cluster = true;
  return GS_OK;
}

}  // extern "C"

// ---------------------------------------------------------------- scoring ---------------------

namespace {

// Phase 1 on every device (raw scores + local extrema), merge into device 0's G, phase 2 inputs ready:
// on return every device's stream has waited for all merges.  G = m->global_ext.p (uint64[2P]).
int score_and_merge(gs_mctx* m, const uint32_t* pod_app, const float* pod_slo, int32_t P, uint32_t flags) {
  const size_t ext_bytes = sizeof(uint64_t) * 2 * (size_t)P;
  Dev& d0 = m->devs[0];
  GSM_CUDA(m, cudaSetDevice(d0.device));
  int rc = m->reserve(m->global_ext, ext_bytes, "global extrema");
  if (rc) return rc;
  GSM_CUDA(m, cudaMemsetAsync(m->global_ext.p, 0, ext_bytes, d0.stream));
  GSM_CUDA(m, cudaEventRecord(m->ev_zero, d0.stream));
  for (Dev& d : m->devs) {
    GSM_CUDA(m, cudaSetDevice(d.device));
    const size_t nl = (size_t)(d.node_end - d.node_begin), PN = (size_t)P * nl;
    if ((rc = m->reserve(d.app, sizeof(uint32_t) * (size_t)P, "pod_app"))) return rc;
    if ((rc = m->reserve(d.slo, sizeof(float) * (size_t)P, "pod_slo"))) return rc;
    if ((rc = m->reserve(d.score, PN ? PN : 1, "scores"))) return rc;
    if ((rc = m->reserve(d.sel, PN ? PN : 1, "selections"))) return rc;
    if ((rc = m->reserve(d.ext, ext_bytes, "local extrema"))) return rc;
    GSM_CUDA(m, cudaMemcpyAsync(d.app.p, pod_app, sizeof(uint32_t) * (size_t)P, cudaMemcpyHostToDevice, d.stream));
    GSM_CUDA(m, cudaMemcpyAsync(d.slo.p, pod_slo, sizeof(float) * (size_t)P, cudaMemcpyHostToDevice, d.stream));
    uint64_t* hi = (uint64_t*)d.ext.p;
    rc = gs_score_raw_dev(d.ctx, (const uint32_t*)d.app.p, (const float*)d.slo.p, nullptr, P,
                          flags & (GS_SCORE_DIRECT | GS_SCORE_NO_DEDUPE | GS_SCORE_LANES_NODES), (uint8_t*)d.score.p,
                          (uint8_t*)d.sel.p, hi, hi + P, d.stream);
    if (rc != GS_OK) return m->fail(rc, "device %d: %s", d.device, gs_last_error(d.ctx));
    if (nl == 0) GSM_CUDA(m, cudaMemsetAsync(d.ext.p, 0, ext_bytes, d.stream));  // (gs_score_raw_dev zeroes them too)
    GSM_CUDA(m, cudaStreamWaitEvent(d.stream, m->ev_zero, 0));
    merge_extrema_kernel<<<(2 * P + 255) / 256, 256, 0, d.stream>>>((const unsigned long long*)d.ext.p,
                                                                     (unsigned long long*)m->global_ext.p, 2 * P);
    GSM_CUDA(m, cudaGetLastError());
    GSM_CUDA(m, cudaEventRecord(d.ev_merged, d.stream));
  }
  for (Dev& d : m->devs) {
    GSM_CUDA(m, cudaSetDevice(d.device));
    for (Dev& o : m->devs) GSM_CUDA(m, cudaStreamWaitEvent(d.stream, o.ev_merged, 0));
  }
  return GS_OK;
}

int check_call(gs_mctx* m, const void* a, const void* b, int32_t P, const char* who) {
  if (P < 0 || (P && (!a || !b))) return m->fail(GS_ERR_INVALID, "%s: null pod table or negative n_pods", who);
  if (!m->have_cluster || !m->have_preds)
    return m->fail(GS_ERR_STATE, "%s needs gs_mupload_cluster and gs_mupload_predictions first", who);
  return GS_OK;
}

}  // namespace

extern "C" {

int gs_mscore_batch(gs_mctx* m, const uint32_t* pod_app, const float* pod_slo, int32_t P, uint32_t flags,
                    uint8_t* out_score, uint8_t* out_sel, int32_t* out_best) {
  if (!m) return GS_ERR_INVALID;
  std::lock_guard<std::mutex> lk(m->mu);
  int rc = check_call(m, pod_app, pod_slo, P, "gs_mscore_batch");
  if (rc || P == 0) return rc;
  if (m->n_nodes == 0) {
    if (out_best)
      for (int32_t p = 0; p < P; ++p) out_best[p] = -1;
    return GS_OK;
  }
  if ((rc = score_and_merge(m, pod_app, pod_slo, P, flags))) return rc;
  const uint64_t* G = (const uint64_t*)m->global_ext.p;
  const size_t N = (size_t)m->n_nodes;
  for (Dev& d : m->devs) {
    GSM_CUDA(m, cudaSetDevice(d.device));
    const size_t nl = (size_t)(d.node_end - d.node_begin);
    const bool want_best = out_best && &d == &m->devs[0];
    if (want_best && (rc = m->reserve(d.best, sizeof(int32_t) * (size_t)P, "best"))) return rc;
    // NormalizeScore on this device's columns with the global extrema; device 0 also decodes the best node
    // (with GS_SCORE_RAW only the decode runs: gs_best_dev)
    if (flags & GS_SCORE_NORMALIZE)
      rc = gs_normalize_dev(d.ctx, P, (uint8_t*)d.score.p, G, G + P, want_best ? (int32_t*)d.best.p : nullptr, d.stream);
    else if (want_best)
      rc = gs_best_dev(d.ctx, P, G, G + P, (int32_t*)d.best.p, d.stream);
    if (rc != GS_OK) return m->fail(rc, "device %d: %s", d.device, gs_last_error(d.ctx));
    if (nl && out_score)
      GSM_CUDA(m, cudaMemcpy2DAsync(out_score + d.node_begin, N, d.score.p, nl, nl, (size_t)P, cudaMemcpyDeviceToHost,
                                    d.stream));
    if (nl && out_sel)
      GSM_CUDA(m, cudaMemcpy2DAsync(out_sel + d.node_begin, N, d.sel.p, nl, nl, (size_t)P, cudaMemcpyDeviceToHost,
                                    d.stream));
    if (want_best)
      GSM_CUDA(m, cudaMemcpyAsync(out_best, d.best.p, sizeof(int32_t) * (size_t)P, cudaMemcpyDeviceToHost, d.stream));
  }
  for (Dev& d : m->devs) {
    GSM_CUDA(m, cudaSetDevice(d.device));
    GSM_CUDA(m, cudaStreamSynchronize(d.stream));
  }
  return GS_OK;
}

int gs_mscore_decisions(gs_mctx* m, const uint32_t* pod_app, const float* pod_slo, int32_t P, uint32_t flags,
                        gs_decision* out) {
  if (!m) return GS_ERR_INVALID;
  std::lock_guard<std::mutex> lk(m->mu);
  int rc = check_call(m, pod_app, out, P, "gs_mscore_decisions");
  if (rc || P == 0) return rc;
  if (!pod_slo) return m->fail(GS_ERR_INVALID, "gs_mscore_decisions: null pod_slo");
  if (m->n_nodes == 0) {
    for (int32_t p = 0; p < P; ++p) out[p] = gs_decision{-1, (uint8_t)GS_SEL_NONE, 0, 0, 0};
    return GS_OK;
  }
  if ((rc = score_and_merge(m, pod_app, pod_slo, P, flags))) return rc;
  const uint64_t* G = (const uint64_t*)m->global_ext.p;
  std::vector<std::vector<gs_decision>> per(m->devs.size());
  for (size_t i = 0; i < m->devs.size(); ++i) {
    Dev& d = m->devs[i];
    GSM_CUDA(m, cudaSetDevice(d.device));
    if ((rc = m->reserve(d.dec, sizeof(gs_decision) * (size_t)P, "decisions"))) return rc;
    rc = gs_decisions_dev(d.ctx, (const uint32_t*)d.app.p, (const float*)d.slo.p, P, (const uint8_t*)d.sel.p, G, G + P,
                          (gs_decision*)d.dec.p, d.stream);
    if (rc != GS_OK) return m->fail(rc, "device %d: %s", d.device, gs_last_error(d.ctx));
    per[i].resize((size_t)P);
    GSM_CUDA(m, cudaMemcpyAsync(per[i].data(), d.dec.p, sizeof(gs_decision) * (size_t)P, cudaMemcpyDeviceToHost, d.stream));
  }
  for (Dev& d : m->devs) {
    GSM_CUDA(m, cudaSetDevice(d.device));
    GSM_CUDA(m, cudaStreamSynchronize(d.stream));
  }
  // node and extrema are global on every device; the selection byte and the MPS choice come from the
  // device that owns the chosen node
  for (int32_t p = 0; p < P; ++p) {
    out[p] = per[0][(size_t)p];
    const int node = out[p].node;
    for (size_t i = 0; i < m->devs.size(); ++i)
      if (node >= m->devs[i].node_begin && node < m->devs[i].node_end) out[p] = per[i][(size_t)p];
  }
  return GS_OK;
}

}  // extern "C"

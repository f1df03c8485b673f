// context.h — the per-device context behind gs_ctx.
#pragma once
#include <cstdint>
#include <mutex>
#include <string>
#include <vector>

#include "tables.h"

namespace gs {

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
};

struct Ctx {
  // gs_complete_last_timing: [fits sampled, moments ms, Gram ms, estimator ms] per fit, of the last completion fit
  double fit_timing[4] = {0.0, 0.0, 0.0, 0.0};
  std::mutex mu;
  std::string err;
  int device = 0;
  cudaStream_t stream = nullptr;

  // cluster snapshot
  bool have_cluster = false;
  int n_nodes = 0, max_gpus = 1, redis_reachable = 1, node_base = 0;
  int64_t n_res = 0;
  DevBuf node_model, node_ngpu, gpu_off, res_app, res_slo, node_orig, res_new;
  // host mirror of the snapshot, so that gs_update_node can splice one node's residents in
  std::vector<uint8_t> h_model, h_ngpu;
  std::vector<uint32_t> h_off, h_app;
  std::vector<float> h_slo;
  // prediction tables
  bool have_preds = false;
  int n_apps = 0, n_cols = 0;
  DevBuf conf, interf, interf_t, col_of;
  // derived
  bool prep_dirty = true;
  DevBuf res_rec;
  // scratch of the restructured scoring path
  DevBuf sc_counts, sc_cursor, sc_nchunks, sc_chunks, sc_sorted, sc_stage, sc_counts_all;
  DevBuf sc_bucket, sc_rep, sc_hkeys, sc_hvals;
  bool dedupe = true;       // class dedupe on (GS_NO_DEDUPE=1 in the environment, or GS_SCORE_NO_DEDUPE per call, turns it off)
  bool lanes_nodes = false; // GS_LANES_NODES=1: always the one-thread-per-node kernel
  // host-buffer entry point: second stream + events to overlap the device->host copies with compute
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_chunk[8] = {};
  cudaEvent_t ev_copy_done = nullptr;
  size_t stage_budget = (size_t)4 << 30;  // largest per-batch residents-sum table (GS_STAGE_TABLE_MB)
  // staging for the host-pointer entry points
  DevBuf b_pod_app, b_pod_slo, b_pod_self, b_score, b_sel, b_extrema, b_best, b_misc;

  // instrumentation
  bool profiling = false;
  bool ev_valid = false;
  cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
  cudaEvent_t ev_prof[3] = {};  // before stage A, after it, after the scoring kernel
  bool ev_prof_valid = false;
  // recorded after the last launch of every device-pointer scoring call (on the caller's stream);
  // table uploads wait for it, so that a re-upload never overwrites tables a kernel still reads
  cudaEvent_t ev_last_use = nullptr;
  int note_use(cudaStream_t s);
  int wait_last_use();

  int fail(int code, const char* fmt, ...) __attribute__((format(printf, 3, 4)));
  int cuda_fail(cudaError_t e, const char* what);
  int reserve(DevBuf& b, size_t bytes, const char* what);
  void release(DevBuf& b);
  int upload(DevBuf& b, const void* src, size_t bytes, const char* what);
  DevTables tables() const;
  int ensure_prepared();
  int ensure_scratch(int n_pods);
  PodScratch scratch_view(int n_pods, bool use_dedupe = true) const;
};

cudaError_t launch_normalize_i64(long long* scores, int n, cudaStream_t s);

}  // namespace gs

// the opaque handle of include/gpusched.h
struct gs_ctx : public gs::Ctx {};

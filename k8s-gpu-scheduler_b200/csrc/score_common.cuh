// score_common.cuh — device helpers shared by the scoring kernels.
#pragma once
#include "score_math.cuh"
#include "tables.h"

namespace gs {

// ------------------------------------------------------------------ helpers ----------

// Fold one pod's per-lane (score, node) into per-pod extrema.  key orders by score, then by
// the lowest node index in the CALLER's order (24 bits: shards of up to 16M nodes); 0 = lane
// holds no node.
constexpr uint32_t kNodeKeyMask = 0x00FFFFFFu;

__device__ __forceinline__ uint32_t lane_key(int score, uint32_t node_orig) {
  return (((uint32_t)score << 24) | (kNodeKeyMask - node_orig)) + 1u;
}

__device__ __forceinline__ void publish_extrema(uint32_t key, uint32_t negmax, int node_base, uint32_t pod,
                                                unsigned long long* hi_arg, unsigned long long* neg_lo) {
  uint32_t k = key - 1u;
  uint32_t score = k >> 24;
  uint32_t node = kNodeKeyMask - (k & kNodeKeyMask) + (uint32_t)node_base;
  unsigned long long packed = ((unsigned long long)score << 32) | (unsigned long long)(0xFFFFFFFFu - node);
  atomicMax(&hi_arg[pod], packed);
  atomicMax(&neg_lo[pod], (unsigned long long)negmax);
}

__device__ __forceinline__ bool conf_lookup(const DevTables& T, uint32_t app, int model, int k, float* out) {
  if (app >= (uint32_t)T.n_apps || k < 0 || k > kMaxGpusPerNode) return false;
  int col = T.col_of[(model - 1) * (kMaxGpusPerNode + 1) + k];
  if (col < 0) return false;
  *out = T.conf[(size_t)app * T.n_cols + col];
  return true;
}

// V100 candidate prediction: smallest of 1P/2P/4P above the SLO, gpu_plugins.go:639-651
__device__ __forceinline__ float conf_v100_for(const DevTables& T, uint32_t app, float slo, bool* f6) {
  float c = -1.0f;
  bool any = false;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    float tmp;
    bool ok = conf_lookup(T, app, kModelV100, 1 << i, &tmp);
    any |= ok;
    if (ok && tmp > slo && (c == -1.0f || tmp < c)) c = tmp;
  }
  *f6 = any;
  return c;
}

// Logic() for one (pod, node), op for op, any input (gpu_plugins.go:455-777).  Returns the
// raw score and writes the packed selection byte.  Used by score_direct_kernel for every
// pair and by score_fast_kernel for the pairs outside its streamlined domain.
// `node` is an internal node index; `self` the CALLER's resident index (or -1).
static __device__ __noinline__ int logic_eval(const DevTables& T, uint32_t app, float cur_slo, int32_t self, int node,
                                              unsigned* out_sel) {
  if (self >= 0) self = (int32_t)T.res_new[self];
  int score = 0;
  unsigned sel = kSelNone, flags = 0;
  const size_t A = (size_t)T.n_apps;
  if (T.redis_reachable) {
    const int model = T.node_model[node];
    const int k = T.node_ngpu[node];
    if (model != kModelNone && k > 0) flags |= kFlagEligible;
    if (model != kModelNone) {
      const float inv_cur = __fdiv_rn(1.0f, cur_slo);
      const float* itab = T.interf + (size_t)(model - 1) * A * A;
      for (int g = 0; g < k; ++g) {
        GpuAcc acc = {0.0, 0.0, 0, 0};
        const uint32_t lo = T.gpu_off[(size_t)node * T.max_gpus + g];
        const uint32_t hi = T.gpu_off[(size_t)node * T.max_gpus + g + 1];
        for (uint32_t r = lo; r < hi; ++r) {
          float slo = T.res_slo[r];
          if (slo == 0.0f) continue;
          uint32_t ar = T.res_app[r];
          float conf;
          if (!conf_lookup(T, ar, model, k, &conf)) continue;
          float interf = 0.f;
          for (uint32_t c = lo; c < hi; ++c) {
            if (c == r || (self >= 0 && c == (uint32_t)self)) continue;
            uint32_t ac = T.res_app[c];
            if (ac < (uint32_t)T.n_apps) interf = __fadd_rn(interf, itab[(size_t)ar * A + ac]);
          }
          if (app < (uint32_t)T.n_apps) interf = __fadd_rn(interf, itab[(size_t)ar * A + app]);
          add_term(acc, slo, __fdiv_rn(1.0f, slo), conf, interf);
        }
        float conf = -1.0f;
        if (model == kModelA30) {
          float v;
          bool ok = conf_lookup(T, app, model, k, &v);
          conf = ok ? v : 0.0f;
          if (ok) flags |= kFlagF6;
        } else {
          bool f6;
          conf = conf_v100_for(T, app, cur_slo, &f6);
          if (f6) flags |= kFlagF6;
        }
        if (conf != -1.0f) {
          flags |= kFlagF7;
          float interf = 0.f;
          if (app < (uint32_t)T.n_apps) {
            for (uint32_t c = lo; c < hi; ++c) {
              if (self >= 0 && c == (uint32_t)self) continue;
              uint32_t ac = T.res_app[c];
              if (ac < (uint32_t)T.n_apps) interf = __fadd_rn(interf, itab[(size_t)app * A + ac]);
            }
          }
          add_term(acc, cur_slo, inv_cur, conf, interf);
        }
        int si = score_int(combine(acc.neg_sum, acc.pos_sum, acc.cnt_neg, acc.cnt_pos));
        if (si > score) {
          score = si;
          sel = (unsigned)g;
        }
      }
    }
  }
  *out_sel = sel | flags;
  return score;
}

}  // namespace gs

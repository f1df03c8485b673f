// score_fast.cu — the restructured Score kernel (sm_100a).
//
// Reference: pkg/plugins/gpu_plugin/gpu_plugins.go:562-757 (the per-GPU loop of Logic).
//
// A resident's term depends on the pod being scheduled only through that pod's APP (the
// interference column added last, :607-612), and so does the candidate's interference sum
// (:703-714).  Pods are therefore bucketed by app; a CTA takes one work item = up to
// kChunk pods of one app x kTile consecutive nodes, one thread per node.  Each thread
//   1. builds, once, the running sums of each of its node's GPUs (negative_sum,
//      positive_sum, counts) and the candidate interference, exactly as Logic does;
//   2. folds everything of the weighted combine (:731-743) that does not depend on the
//      candidate's own term into per-(GPU, branch) constants:
//        branch "neg" (candidate's SLO > conf - interf):  tmp = cst + 100*((kk*(N0+v))/c)
//        branch "pos":                                     tmp = cst + 100*((kk*(P0+v))/c)
//      with kk = k or 1-k (1 if the other side is empty), c = the branch's count, cst =
//      the other side's finished product (0 if empty).  These are the SAME IEEE
//      operations in the SAME order as the reference performs — x*1.0 and 0.0+x are exact
//      — so the result is bit-identical;
//   3. streams the bucket's pods through them: per pod and GPU one reciprocal, one
//      division by a small integer and eight other float64 operations remain.
//
// Float32 filter.  The integer score is floor(tmp) and tmp = A + B*v with per-(GPU, branch)
// constants A = cst + 100*kk*base/c, B = 100*kk/c and v = 1/z in (0, 1].  A float32 evaluation of
// A + B*v (one approximate reciprocal, two fused multiply-adds) lies within kFilterErr of the
// float64 value the reference computes (bound derived at kFilterDelta below), so whenever it is
// farther than that from an integer its floor IS the reference's int64(tmp).  Only the few
// (pod, node, GPU) whose estimate falls within kFilterDelta of an integer — about 1.2e-4 of them
// — are evaluated with the exact float64 sequence (for the whole warp, chosen by a vote, so
// there is no divergence).  The float32 constants live in shared memory (one 8-byte slot per
// thread, GPU and branch); the exact path rebuilds kk/cst from the sums (registers, or the
// stage-A table) when it is needed.
//
// Pods without an SLO (SLO 0, 10 % of the synthetic batch; also denormal SLOs) have an
// infinite 1/SLO: the candidate's term is exactly 0 (or NaN when SLO == conf - interf), so
// int64(tmp) per (GPU, branch) is a per-thread constant, precomputed in step 2, and the pod
// costs a handful of float32 compares per GPU.
//
// Pods whose SLO is NaN/Inf, lanes whose float32 term could overflow, and nodes whose
// sums are not finite take the general path (score_common.cuh's op-for-op logic_eval) —
// chosen per warp, so the streamlined loop is divergence-free.
#include "score_fold.cuh"

namespace gs {

template <int GT>
struct FastSmem {
  float2 ab[GT][2][kTile];      // [g][0 = neg, 1 = pos][thread] = float32 {A, B}: tmp ~ A + B*v
  uint16_t zinf[GT][kTile];     // int64(tmp) when the candidate's term is 0: neg | pos << 8
  double2 lut[kLutSize];        // {RN(1/c), (double)c}
  PodRec pods[kStage];
  uint32_t key[kTile / 32][kStage];
  uint32_t neg[kTile / 32][kStage];
};
static_assert(sizeof(FastSmem<kFastMaxGpus>) <= 48 * 1024, "static shared memory");
static_assert(kKinds == 4, "stage_a_kernel reads the four kind counts of an app as one uint4");
static_assert(kStage == kTile, "a CTA stages its pods with one thread each");

// ------------------------------------------------------------------ stage A ----------

// The residents' sums of every (app bucket, GPU slot) that has pods in this batch.  A warp
// takes one GPU slot with its 32 lanes on 32 consecutive apps: the slot's resident list is
// the same for all lanes (uniform trip count, broadcast loads) and both interference reads
// are contiguous in the app (row r of interf, row r of its transpose).  Results go through a
// padded shared tile so that the table is written, and later read, slot-contiguous.
constexpr int kStageTile = 32;

__global__ void __launch_bounds__(256) stage_a_kernel(DevTables T, const uint32_t* __restrict__ counts, StageA out) {
  __shared__ double s_n0[kStageTile][kStageTile + 1], s_p0[kStageTile][kStageTile + 1];
  __shared__ float s_ci[kStageTile][kStageTile + 1];
  __shared__ uint32_t s_cnt[kStageTile][kStageTile + 1];
  // per slot of the tile: resident range and model (0 = nothing to sum), worked out once by
  // the first warp instead of by every lane of the warp that owns the slot
  __shared__ uint32_t s_lo[kStageTile], s_hi[kStageTile];
  __shared__ int s_model[kStageTile];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int slots = T.n_nodes * T.max_gpus;
  const int s0 = blockIdx.x * kStageTile, a0 = blockIdx.y * kStageTile;
  const int app_idx = a0 + lane;  // bucket n_apps = pods whose app is unknown
  const bool app_known = app_idx < T.n_apps;
  bool wanted = false;  // does any (app, kind) bucket of this app hold pods?
  if (app_idx <= T.n_apps) {
    const uint4 c4 = *reinterpret_cast<const uint4*>(counts + (size_t)app_idx * kKinds);
    wanted = (c4.x | c4.y | c4.z | c4.w) != 0u;
  }
  if (warp == 0) {
    const int slot = s0 + lane;
    int model = kModelNone;
    uint32_t lo = 0, hi = 0;
    if (slot < slots && T.redis_reachable) {
      const int node = slot / T.max_gpus, g = slot - node * T.max_gpus;
      if (g < (int)T.node_ngpu[node]) model = T.node_model[node];
      lo = T.gpu_off[slot];
      hi = T.gpu_off[slot + 1];
    }
    s_model[lane] = model;
    s_lo[lane] = lo;
    s_hi[lane] = hi;
  }
  if (!__syncthreads_or(wanted)) return;  // no pod of these 32 apps in the batch
  const size_t A = (size_t)T.n_apps;
#pragma unroll 1
  for (int j = 0; j < kStageTile / 8; ++j) {
    const int sl = warp * (kStageTile / 8) + j;
    GpuAcc acc = {0.0, 0.0, 0, 0};
    float ci = 0.f;
    const int model = s_model[sl];
    if (model != kModelNone) {
      const float* itab = T.interf + (size_t)(model - 1) * A * A + app_idx;
      const float* itab_t = T.interf_t + (size_t)(model - 1) * A * A + app_idx;
      const uint32_t lo = s_lo[sl], hi = s_hi[sl];
      for (uint32_t r = lo; r < hi; ++r) {
        const uint32_t ar = T.res_app[r];
        const float4 rec = T.res_rec[r];
        if (app_known && ar < (uint32_t)T.n_apps) ci = __fadd_rn(ci, itab_t[(size_t)ar * A]);
        if (rec.x != 0.0f) {
          float interf = rec.w;
          if (app_known) interf = __fadd_rn(interf, itab[(size_t)ar * A]);
          add_term(acc, rec.x, rec.y, rec.z, interf);
        }
      }
    }
    s_n0[lane][sl] = acc.neg_sum;
    s_p0[lane][sl] = acc.pos_sum;
    s_ci[lane][sl] = ci;
    s_cnt[lane][sl] = (uint32_t)acc.cnt_neg | ((uint32_t)acc.cnt_pos << 16);
  }
  __syncthreads();
#pragma unroll
  for (int pass = 0; pass < kStageTile / 8; ++pass) {
    const int al = pass * 8 + warp;
    if (a0 + al <= T.n_apps && s0 + lane < slots) {
      const size_t o = (size_t)(a0 + al) * out.stride + (size_t)(s0 + lane);
      out.n0[o] = s_n0[al][lane];
      out.p0[o] = s_p0[al][lane];
      out.ci[o] = s_ci[al][lane];
      out.cnt[o] = s_cnt[al][lane];
    }
  }
}

// interf_t[m][b][a] = interf[m][a][b]
__global__ void __launch_bounds__(256) transpose_interf_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                              int n) {
  __shared__ float tile[32][33];
  const size_t plane = (size_t)blockIdx.z * n * n;
  int x = blockIdx.x * 32 + (threadIdx.x & 31), y0 = blockIdx.y * 32 + (threadIdx.x >> 5);
  for (int j = 0; j < 32; j += 8)
    if (x < n && y0 + j < n) tile[(threadIdx.x >> 5) + j][threadIdx.x & 31] = in[plane + (size_t)(y0 + j) * n + x];
  __syncthreads();
  x = blockIdx.y * 32 + (threadIdx.x & 31);
  y0 = blockIdx.x * 32 + (threadIdx.x >> 5);
  for (int j = 0; j < 32; j += 8)
    if (x < n && y0 + j < n) out[plane + (size_t)(y0 + j) * n + x] = tile[threadIdx.x & 31][(threadIdx.x >> 5) + j];
}

cudaError_t launch_transpose_interf(const float* interf, float* interf_t, int n_apps, cudaStream_t s) {
  dim3 grid((n_apps + 31) / 32, (n_apps + 31) / 32, 2);
  note_launch();
  transpose_interf_kernel<<<grid, 256, 0, s>>>(interf, interf_t, n_apps);
  return cudaGetLastError();
}

// ------------------------------------------------------------------ main kernel ------

// a[g] for a run-time g on an array the compiler keeps in registers (a select chain instead of
// local memory); only the rarely taken exact path uses these
template <typename V, int GT>
__device__ __forceinline__ V dyn_at(const V (&a)[GT], int g) {
  V x = a[0];
#pragma unroll
  for (int i = 1; i < GT; ++i) x = g == i ? a[i] : x;
  return x;
}

// GT = GPU slots the unrolled loops cover: the smallest of 2, 4, 8 that holds max_gpus, so that
// clusters with few GPUs per node do not pay for eight.
template <bool kTable, int GT>
__global__ void __launch_bounds__(kTile, GS_MIN_CTAS) score_fast_kernel(DevTables T, StageA stage, const PodRec* __restrict__ pods,
                                                             const ChunkDesc* __restrict__ chunks,
                                                             const uint32_t* __restrict__ n_chunks,
                                                             uint8_t* __restrict__ out_score,
                                                             uint8_t* __restrict__ out_sel,
                                                             unsigned long long* hi_arg,
                                                             unsigned long long* neg_lo, int skip_plain) {
  __shared__ FastSmem<GT> sm_storage;
  FastSmem<GT>* sm = &sm_storage;
  if (blockIdx.x >= *n_chunks) return;
  const int tid = threadIdx.x;
  const ChunkDesc cd = chunks[blockIdx.x];
  (void)skip_plain;
  if (tid < kLutSize) sm->lut[tid] = make_double2(tid ? __ddiv_rn(1.0, (double)tid) : 0.0, (double)tid);

  const uint32_t tile_node0 = blockIdx.y * kTile;
  const int node = (int)tile_node0 + tid;
  const bool in_range = node < T.n_nodes;
  const uint32_t norig = in_range ? T.node_orig[node] : 0u;  // the caller's index of this node
  const uint32_t app = cd.app;
  const bool app_known = app < (uint32_t)T.n_apps;
  const size_t A = (size_t)T.n_apps;

  int model = kModelNone, k = 0;
  if (in_range && T.redis_reachable) {
    model = T.node_model[node];
    k = T.node_ngpu[node];
  }
  const bool scored = model != kModelNone && k > 0;  // the GPU loop of Logic runs at least once
  unsigned node_flags = scored ? kFlagEligible : 0u;

  float conf_a30 = 0.0f, ci_max = 0.0f;
  int nocand = (int)kSelNone;  // score << 8 | sel for pods without a candidate term (:667-696)
  bool thr_general = false;    // this node's sums/counts fall outside the streamlined domain

  // Steps 1+2.  GPUs a node does not have (and lanes without a node) keep constants that
  // evaluate to tmp = 0 + 100*((0*(0+v))/1) = 0: they can never win (a GPU needs
  // int64(tmp) >= 1), so the streamlined loop needs no guard.
  // per-GPU state: the candidate's interference sum in registers; the residents' sums and counts
  // are only needed again by the exact path — with the stage-A table it re-reads them there, so
  // these arrays are dead after the set-up; without it they stay in registers
  double base_n[GT], base_p[GT];
  float cand_interf[GT];
  uint32_t cnts[GT];  // cnt_neg | cnt_pos << 16
#pragma unroll
  for (int g = 0; g < GT; ++g) {
    base_n[g] = 0.0;
    base_p[g] = 0.0;
    cand_interf[g] = 0.f;
    cnts[g] = 0u;
    // estimate 0.5: score 0, never near an integer (GPUs the node lacks, lanes without a node)
    sm->ab[g][0][tid] = make_float2(0.5f, 0.f);
    sm->ab[g][1][tid] = make_float2(0.5f, 0.f);
    sm->zinf[g][tid] = 0;
  }
  __syncthreads();  // lut
  int nc_score = 0;
  unsigned nc_sel = kSelNone;
  size_t st_row = 0;  // this thread's row of the stage-A table (kTable, scored lanes)
  // The streamlined domain of one GPU: non-negative finite sums of ordinary size (they are sums
  // of fewer than 64 values in (0, 1]), a finite candidate interference, counts inside the table.
  auto in_domain = [&](const GpuAcc& acc, float ci) -> bool {
    return acc.neg_sum >= 0.0 && acc.neg_sum <= 1.0e6 && acc.pos_sum >= 0.0 && acc.pos_sum <= 1.0e6 &&
           fabsf(ci) <= 3.0e38f && acc.cnt_neg + acc.cnt_pos + 1 < kLutSize;  // false for NaN
  };
  // Fold one GPU's sums into what the pod loop needs, inside the streamlined domain: the same
  // operations in the same order as combine() / branch_const(), every division through the table
  // and none of them guarded.  An empty side needs no special case: its k comes out as exactly
  // 1 or 0 (c/c and 0/c are exact) and its product as exactly 0 (lut[0] = {0, 0} turns m/0 into
  // 0), and x*1.0, 0.0+x are exact — so the values equal the guarded form's literals 1.0 and 0.0.
  auto fold_fast = [&](int g, const GpuAcc& acc, float ci) {
    const double2* lut = sm->lut;
    const int n = acc.cnt_neg, p = acc.cnt_pos;
    const double ns = acc.neg_sum, ps = acc.pos_sum;
    const double2 Ln = lut[n], Lp = lut[p], Lnp = lut[n + p], Lnp1 = lut[n + p + 1], Ln1 = lut[n + 1], Lp1 = lut[p + 1];
    auto dv = [](double m, const double2& L) { return div_small_int(m, L.x, L.y); };
    // residents only (no candidate term, :667-696)
    const double k0 = dv((double)n, Lnp);
    const double tmp0 = __dadd_rn(__dmul_rn(100.0, dv(__dmul_rn(__dsub_rn(1.0, k0), ps), Lp)),
                                  __dmul_rn(100.0, dv(__dmul_rn(k0, ns), Ln)));
    const int si = score_int(tmp0);
    if (si > nc_score) {
      nc_score = si;
      nc_sel = (unsigned)g;
    }
    // candidate on the negative side: k = (n+1)/(n+1+p)
    const double kk_n = dv((double)(n + 1), Lnp1);
    const double cst_n = __dmul_rn(100.0, dv(__dmul_rn(__dsub_rn(1.0, kk_n), ps), Lp));
    // candidate on the positive side: k = n/(n+p+1)
    const double kq_p = dv((double)n, Lnp1);
    const double kk_p = __dsub_rn(1.0, kq_p);
    const double cst_p = __dmul_rn(100.0, dv(__dmul_rn(kq_p, ns), Ln));
    // a candidate term of exactly 0 on either side (pods without an SLO)
    const double tn = __dadd_rn(cst_n, __dmul_rn(100.0, dv(__dmul_rn(kk_n, __dadd_rn(ns, 0.0)), Ln1)));
    const double tp = __dadd_rn(cst_p, __dmul_rn(100.0, dv(__dmul_rn(kk_p, __dadd_rn(ps, 0.0)), Lp1)));
    sm->zinf[g][tid] = (uint16_t)((unsigned)score_int(tn) | ((unsigned)score_int(tp) << 8));
    // float32 filter: tmp = A + B*v with A = tmp at v = 0 (tn, tp) and B = 100*kk/c; (0, 0) =
    // "always take the exact path" outside the range the error bound assumes (see kFilterDelta)
    const double Bn = 100.0 * kk_n * Ln1.x, Bp = 100.0 * kk_p * Lp1.x;
    const bool okn = tn >= 0.0 && Bn >= 0.0 && tn + Bn <= kFilterMaxSum, okp = tp >= 0.0 && Bp >= 0.0 && tp + Bp <= kFilterMaxSum;
    sm->ab[g][0][tid] = okn ? make_float2((float)tn, (float)Bn) : make_float2(0.f, 0.f);
    sm->ab[g][1][tid] = okp ? make_float2((float)tp, (float)Bp) : make_float2(0.f, 0.f);
    ci_max = fmaxf(ci_max, fabsf(ci));
  };
  // The same outside the domain (guarded divisions, IEEE division where the table does not reach):
  // the residents-only score and the zero-term scores stay exact, candidates of this node go
  // through logic_eval.
  auto fold_general = [&](int g, const GpuAcc& acc, float ci) {
    const double2* lut = sm->lut;
    const int cn0 = acc.cnt_neg, cp0 = acc.cnt_pos;
    const int si = score_int(combine_lut(acc.neg_sum, acc.pos_sum, cn0, cp0, lut));
    if (si > nc_score) {
      nc_score = si;
      nc_sel = (unsigned)g;
    }
    const double2 kn = branch_const(true, acc.neg_sum, acc.pos_sum, cn0, cp0, lut);  // {kk, cst}
    const double2 kp = branch_const(false, acc.neg_sum, acc.pos_sum, cn0, cp0, lut);
    const double tn = __dadd_rn(kn.y, __dmul_rn(100.0, div_count(__dmul_rn(kn.x, __dadd_rn(acc.neg_sum, 0.0)),
                                                                cn0 + 1, lut)));
    const double tp = __dadd_rn(kp.y, __dmul_rn(100.0, div_count(__dmul_rn(kp.x, __dadd_rn(acc.pos_sum, 0.0)),
                                                                cp0 + 1, lut)));
    sm->zinf[g][tid] = (uint16_t)((unsigned)score_int(tn) | ((unsigned)score_int(tp) << 8));
    sm->ab[g][0][tid] = make_float2(0.f, 0.f);
    sm->ab[g][1][tid] = make_float2(0.f, 0.f);
    ci_max = fmaxf(ci_max, fabsf(ci));
    thr_general = true;
  };
  if (scored) {
    if (model == kModelA30) {
      float v;
      bool ok = conf_lookup(T, app, model, k, &v);
      conf_a30 = ok ? v : 0.0f;  // missing key -> Go zero value, gpu_plugins.go:637
      if (ok) node_flags |= kFlagF6;
    }
    if (kTable) {
      // sums written by stage_a_kernel, slot-contiguous (coalesced across the tile): all loads
      // first, then the folds as eight independent chains
      st_row = (size_t)(app_known ? app : (uint32_t)T.n_apps) * stage.stride + (size_t)node * T.max_gpus;
#pragma unroll
      for (int g = 0; g < GT; ++g) {
        if (g < k) {
          // read once: streaming loads, so that the table does not push half-written output
          // sectors out of L2
          base_n[g] = __ldcs(stage.n0 + st_row + g);
          base_p[g] = __ldcs(stage.p0 + st_row + g);
          cand_interf[g] = __ldcs(stage.ci + st_row + g);
          cnts[g] = __ldcs(stage.cnt + st_row + g);
        }
      }
      bool dom = true;
#pragma unroll
      for (int g = 0; g < GT; ++g) {
        if (g < k) {
          const GpuAcc acc = {base_n[g], base_p[g], (int)(cnts[g] & 0xFFFFu), (int)(cnts[g] >> 16)};
          dom = dom && in_domain(acc, cand_interf[g]);
        }
      }
      if (dom) {
#pragma unroll
        for (int g = 0; g < GT; ++g) {
          if (g < k) {
            const GpuAcc acc = {base_n[g], base_p[g], (int)(cnts[g] & 0xFFFFu), (int)(cnts[g] >> 16)};
            fold_fast(g, acc, cand_interf[g]);
          }
        }
      } else {
#pragma unroll 1
        for (int g = 0; g < k; ++g) {  // rare: one rolled copy of the guarded code
          const uint32_t cn = stage.cnt[st_row + g];
          const GpuAcc acc = {stage.n0[st_row + g], stage.p0[st_row + g], (int)(cn & 0xFFFFu), (int)(cn >> 16)};
          fold_general(g, acc, stage.ci[st_row + g]);
        }
      }
    } else {
      // walk the resident lists (a rolled loop: its results go through thread-local arrays)
      const float* itab = T.interf + (size_t)(model - 1) * A * A;
      double l_bn[GT], l_bp[GT];
      float l_ci[GT];
      uint32_t l_cnt[GT];
#pragma unroll
      for (int g = 0; g < GT; ++g) {
        l_bn[g] = 0.0;
        l_bp[g] = 0.0;
        l_ci[g] = 0.f;
        l_cnt[g] = 0u;
      }
      const uint32_t* offp = T.gpu_off + (size_t)node * T.max_gpus;
      uint32_t lo = offp[0];
#pragma unroll 1
      for (int g = 0; g < k; ++g) {
        GpuAcc acc = {0.0, 0.0, 0, 0};
        float ci = 0.f;
        const uint32_t hi = offp[g + 1];
        for (uint32_t r = lo; r < hi; ++r) {
          const uint32_t ar = T.res_app[r];
          const float4 rec = T.res_rec[r];
          if (app_known && ar < (uint32_t)T.n_apps) ci = __fadd_rn(ci, itab[(size_t)app * A + ar]);
          if (rec.x != 0.0f) {
            float interf = rec.w;
            if (app_known) interf = __fadd_rn(interf, itab[(size_t)ar * A + app]);
            add_term(acc, rec.x, rec.y, rec.z, interf);
          }
        }
        lo = hi;
        if (in_domain(acc, ci))
          fold_fast(g, acc, ci);
        else
          fold_general(g, acc, ci);
        l_bn[g] = acc.neg_sum;
        l_bp[g] = acc.pos_sum;
        l_ci[g] = ci;
        l_cnt[g] = (uint32_t)acc.cnt_neg | ((uint32_t)acc.cnt_pos << 16);
      }
#pragma unroll
      for (int g = 0; g < GT; ++g) {
        base_n[g] = l_bn[g];
        base_p[g] = l_bp[g];
        cand_interf[g] = l_ci[g];
        cnts[g] = l_cnt[g];
      }
    }
    nocand = (nc_score << 8) | (int)nc_sel;
  }

  const int warp = tid >> 5;
  const unsigned char* lut_bytes = reinterpret_cast<const unsigned char*>(sm->lut);
  for (uint32_t stage0 = 0; stage0 < cd.len; stage0 += kStage) {
    const int len = (int)min((uint32_t)kStage, cd.len - stage0);
    __syncthreads();  // previous stage fully consumed (and, first time, step 2's shared writes)
    if (tid < len) sm->pods[tid] = pods[cd.start + stage0 + tid];
    __syncthreads();
    for (int i = 0; i < len; ++i) {
      const PodRec pr = sm->pods[i];
      const uint32_t orig = pr.orig & kPodIndexMask;
      float conf = conf_a30;
      unsigned flags = node_flags;
      if (scored && model == kModelV100) {
        conf = pr.conf_v100;
        if (pr.orig & kPodF6) flags |= kFlagF6;
      }
      const bool has_cand = scored && conf != -1.0f;
      int packed = nocand;  // also right for !scored lanes: score 0, nothing selected
      // nodes are grouped by model (gs_upload_cluster), so whole warps have no candidate term:
      // warps of nodes without model/GPUs, and V100 warps when no configuration exceeds the SLO
      if (!__any_sync(0xFFFFFFFFu, has_cand)) {
        // nothing to evaluate: every lane keeps its residents-only result
      } else if (pr.orig & kPodInfInv) {
        // t = +-Inf * (slo - d): 1/(1+Inf) = 0 on either side, NaN when slo - d is 0 or NaN
        int score = 0;
        unsigned sel = kSelNone;
#pragma unroll
        for (int g = 0; g < GT; ++g) {
          const float d = __fsub_rn(conf, cand_interf[g]);
          const float e = __fsub_rn(pr.slo, d);
          const unsigned z = sm->zinf[g][tid];
          int si = pr.slo > d ? (int)(z & 0xFFu) : (int)(z >> 8);
          if (!(e < 0.0f || e > 0.0f)) si = 0;  // NaN term: int64(NaN) never wins
          if (si > score) {
            score = si;
            sel = (unsigned)g;
          }
        }
        if (has_cand) packed = (score << 8) | (int)sel;
      } else {
        // May any lane's float32 term leave the finite range?  |t| <= |1/slo|*(|slo|+|conf|+max|ci|).
        const float t_bound = fabsf(pr.inv_slo) * (fabsf(pr.slo) + fabsf(conf) + ci_max);
        const bool lane_general = has_cand && (thr_general || !(t_bound < 1.0e38f));
        if (!((pr.orig & kPodGeneral) || __any_sync(0xFFFFFFFFu, lane_general))) {
          // running maximum of key = int64(tmp) * 8 + (7 - g): the highest score wins, the
          // first GPU among equals (the reference's strict `>` in visiting order)
          // both branch records as 32-bit shared addresses in registers: one select + one load with
          // an immediate offset per GPU
          uint32_t ab_neg = (uint32_t)__cvta_generic_to_shared(&sm->ab[0][0][tid]);
          uint32_t ab_pos = (uint32_t)__cvta_generic_to_shared(&sm->ab[0][1][tid]);
          asm volatile("" : "+r"(ab_neg), "+r"(ab_pos));  // keep both materialised (else: select + add per GPU)
          // float32 part for two GPUs at a time with packed FADD2/FMUL2 (each half rounds like the
          // scalar instruction): -d = interf - conf (exactly -(conf - interf)), slo - d = slo + (-d)
          const float2 nconf2 = make_float2(-conf, -conf);
          const float2 slo2 = make_float2(pr.slo, pr.slo), inv2 = make_float2(pr.inv_slo, pr.inv_slo);
          // float32 estimate of tmp: returns floor(est), *r = est - floor(est) in [0, 1)
          auto estimate = [&](int g, bool neg, float t, float* r) -> float {
            const float z = filter_z(t, neg);
            float2 ab;
            asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];"
                         : "=f"(ab.x), "=f"(ab.y)
                         : "r"((neg ? ab_neg : ab_pos) + (uint32_t)(g * 2 * kTile * 8)));  // folds into [R+imm]
            const float est = filter_est(ab.x, ab.y, z);
            const float fl = filter_floor(est);                        // floor for 0 <= est < 2^22, exact
            *r = est - fl;                                             // exact
            return fl;
          };
          float fbest = 0.f, rlo = 0.5f, rhi = 0.5f;
#pragma unroll
          for (int gp = 0; gp < GT / 2; ++gp) {
            const float2 nd2 = __fadd2_rn(make_float2(cand_interf[2 * gp], cand_interf[2 * gp + 1]), nconf2);
            const float2 t2 = __fmul2_rn(inv2, __fadd2_rn(slo2, nd2));
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              const int g = 2 * gp + h;
              float r;
              const float fl = estimate(g, pr.slo > -(h ? nd2.y : nd2.x), h ? t2.y : t2.x, &r);
              rlo = fminf(rlo, r);
              rhi = fmaxf(rhi, r);
              fbest = fmaxf(fbest, fmaf(fl, 8.0f, (float)(GT - 1 - g)));
            }
          }
          int best = (int)fbest;
          const bool unsure = has_cand && !(rlo >= kFilterDelta && rhi <= 1.0f - kFilterDelta);
          if (__any_sync(0xFFFFFFFFu, unsure)) {
            // Some estimate of this warp lies within kFilterDelta of an integer: redo the pod GPU by
            // GPU, with the exact float64 sequence wherever a lane is unsure about that GPU.
            best = 0;
            const double2* lut = sm->lut;
#pragma unroll 1
            for (int g = 0; g < GT; ++g) {
              const float nd = __fadd_rn(dyn_at(cand_interf, g), -conf);
              const float t = __fmul_rn(pr.inv_slo, __fadd_rn(pr.slo, nd));
              const bool neg = pr.slo > -nd;
              float r;
              int si = (int)estimate(g, neg, t, &r);
              const bool near = has_cand && !(r >= kFilterDelta && r <= 1.0f - kFilterDelta);
              if (__any_sync(0xFFFFFFFFu, near)) {
                double bn, bp;
                uint32_t cn;
                if (kTable) {
                  const bool ld = scored && g < k;
                  bn = ld ? stage.n0[st_row + g] : 0.0;
                  bp = ld ? stage.p0[st_row + g] : 0.0;
                  cn = ld ? stage.cnt[st_row + g] : 0u;
                } else {
                  bn = dyn_at(base_n, g);
                  bp = dyn_at(base_p, g);
                  cn = dyn_at(cnts, g);
                }
                const int cn0 = (int)(cn & 0xFFFFu), cp0 = (int)(cn >> 16);
                double w = fabs((double)t);
                if (neg) w = __dadd_rn(w, 1.0);
                if (neg) w = __dmul_rn(w, w);
                const double v = rcp_ge1(__dadd_rn(1.0, w));
                const double2 kc = branch_const(neg, bn, bp, cn0, cp0, lut);
                const double2 rc = lut[min(neg ? cn0 + 1 : cp0 + 1, kLutSize - 1)];  // lanes outside the domain: discarded
                const double S = __dadd_rn(neg ? bn : bp, v);
                const double q = div_small_int(__dmul_rn(kc.x, S), rc.x, rc.y);
                si = floor_nonneg(__dadd_rn(kc.y, __dmul_rn(100.0, q)));
                if (!(scored && g < k)) si = 0;  // a GPU this lane's node does not have never wins
              }
              best = max(best, si * 8 + (GT - 1 - g));
            }
          }
          const int score = best >> 3;
          const unsigned sel = score > 0 ? (unsigned)(GT - 1 - (best & 7)) : kSelNone;
          if (has_cand) packed = (score << 8) | (int)sel;
        } else if (has_cand) {
          unsigned selbyte;
          const int sc = logic_eval(T, app, pr.slo, -1, node, &selbyte);
          packed = (sc << 8) | (int)(selbyte & 31u);
        }
      }
      if (has_cand) flags |= kFlagF7;
      const int score = packed >> 8;
      if (in_range) {
        const size_t o = (size_t)orig * T.n_nodes + norig;
        out_score[o] = (uint8_t)score;
        out_sel[o] = (uint8_t)((unsigned)(packed & 0xFF) | flags);
      }
      uint32_t key = in_range ? lane_key(score, norig) : 0u;
      uint32_t ngm = in_range ? (uint32_t)(100 - score) : 0u;
      key = __reduce_max_sync(0xFFFFFFFFu, key);
      ngm = __reduce_max_sync(0xFFFFFFFFu, ngm);
      if ((tid & 31) == 0) {
        sm->key[warp][i] = key;
        sm->neg[warp][i] = ngm;
      }
    }
    __syncthreads();
    if (tid < len) {
      uint32_t key = sm->key[0][tid], ngm = sm->neg[0][tid];
#pragma unroll
      for (int w = 1; w < kTile / 32; ++w) {
        key = max(key, sm->key[w][tid]);
        ngm = max(ngm, sm->neg[w][tid]);
      }
      publish_extrema(key, ngm, T.node_base, sm->pods[tid].orig & kPodIndexMask, hi_arg, neg_lo);
    }
  }
}

// stage A for every app bucket that `counts` (pods per bucket) marks as present
cudaError_t launch_stage_a(const DevTables& T, const uint32_t* counts, const StageA& stage, cudaStream_t s) {
  const int slots = T.n_nodes * T.max_gpus;
  if (slots == 0 || !stage.n0) return cudaSuccess;
  dim3 ga((unsigned)((slots + kStageTile - 1) / kStageTile), (unsigned)((T.n_apps + 1 + kStageTile - 1) / kStageTile));
  note_launch();
  stage_a_kernel<<<ga, 256, 0, s>>>(T, counts, stage);
  return cudaGetLastError();
}

// the scoring kernel over the work items that launch_pod_bucketing left in `sc`
cudaError_t launch_score_main(const DevTables& T, const PodScratch& sc, uint8_t* score, uint8_t* sel,
                              unsigned long long* hi_arg, unsigned long long* neg_lo, cudaStream_t s, bool skip_plain) {
  dim3 grid(sc.max_chunks, (unsigned)((T.n_nodes + kTile - 1) / kTile));
#define GS_LAUNCH_FAST(TABLE, GT)                                                                                  \
  score_fast_kernel<TABLE, GT><<<grid, kTile, 0, s>>>(T, sc.stage, sc.sorted, sc.chunks, sc.n_chunks, score, sel, \
                                                      hi_arg, neg_lo, skip_plain ? 1 : 0)
  const bool table = sc.stage.n0 != nullptr;
  note_launch();
  if (T.max_gpus <= 2) {
    if (table) GS_LAUNCH_FAST(true, 2); else GS_LAUNCH_FAST(false, 2);
  } else if (T.max_gpus <= 4) {
    if (table) GS_LAUNCH_FAST(true, 4); else GS_LAUNCH_FAST(false, 4);
  } else {
    if (table) GS_LAUNCH_FAST(true, 8); else GS_LAUNCH_FAST(false, 8);
  }
#undef GS_LAUNCH_FAST
  return cudaGetLastError();
}

// the lanes-are-pods kernel needs the per-batch table (here: in its folded form) and at most 8 GPU
// slots; everything else (small batches, tables beyond the budget, the no-allocation entry point)
// runs the one-thread-per-node kernel
bool use_pods_kernel(const DevTables& T, const PodScratch& sc, bool lanes_nodes) {
  return !lanes_nodes && sc.stage.n0 != nullptr && T.max_gpus <= kFastMaxGpus;
}

cudaError_t launch_stage_any(const DevTables& T, const uint32_t* counts, const PodScratch& sc, cudaStream_t s,
                             bool lanes_nodes) {
  return use_pods_kernel(T, sc, lanes_nodes) ? launch_stage_fold(T, counts, sc.stage, s)
                                             : launch_stage_a(T, counts, sc.stage, s);
}

cudaError_t launch_score_any(const DevTables& T, const PodScratch& sc, uint8_t* score, uint8_t* sel,
                             unsigned long long* hi_arg, unsigned long long* neg_lo, cudaStream_t s, bool lanes_nodes) {
  if (use_pods_kernel(T, sc, lanes_nodes)) return launch_score_pods(T, sc, score, sel, hi_arg, neg_lo, s);
  return launch_score_main(T, sc, score, sel, hi_arg, neg_lo, s, false);
}

// prof (optional) = three events: before the stage kernel, between it and the scoring kernel, after
// the scoring kernel (the class expansion that follows is outside the bracket)
cudaError_t launch_score_fast(const DevTables& T, const uint32_t* pod_app, const float* pod_slo, int n_pods,
                              const PodScratch& sc, uint8_t* score, uint8_t* sel, unsigned long long* hi_arg,
                              unsigned long long* neg_lo, cudaStream_t s, const cudaEvent_t* prof, bool lanes_nodes) {
  if (n_pods == 0 || T.n_nodes == 0) return cudaSuccess;
  cudaError_t e = launch_pod_bucketing(T, pod_app, pod_slo, n_pods, sc, s);
  if (e != cudaSuccess) return e;
  if (prof && (e = cudaEventRecord(prof[0], s)) != cudaSuccess) return e;
  if ((e = launch_stage_any(T, sc.counts, sc, s, lanes_nodes)) != cudaSuccess) return e;
  if (prof && (e = cudaEventRecord(prof[1], s)) != cudaSuccess) return e;
  if ((e = launch_score_any(T, sc, score, sel, hi_arg, neg_lo, s, lanes_nodes)) != cudaSuccess) return e;
  if (prof && (e = cudaEventRecord(prof[2], s)) != cudaSuccess) return e;
  return launch_expand_classes(T, sc.rep, n_pods, score, sel, hi_arg, neg_lo, s);
}

// ------------------------------------------------------------------ self test --------

// Compares the streamlined arithmetic with CUDA's IEEE intrinsics on pseudo-random
// operands from the domains the kernel uses them on; counts mismatches.
__global__ void selftest_kernel(unsigned long long seed, int iters, unsigned long long* mismatches) {
  unsigned long long x = seed + 0x9E3779B97F4A7C15ull * (blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x + 1);
  unsigned long long bad = 0;
  for (int it = 0; it < iters; ++it) {
    x ^= x << 13;
    x ^= x >> 7;
    x ^= x << 17;
    // z >= 1: random mantissa, exponent 0..258 (the kernel's z is below 2^257)
    unsigned long long mant = x & 0x000FFFFFFFFFFFFFull;
    unsigned long long ez = 1023ull + ((x >> 52) % 259ull);
    double z = __longlong_as_double((long long)((ez << 52) | mant));
    if (__double_as_longlong(rcp_ge1(z)) != __double_as_longlong(__drcp_rn(z))) ++bad;
    // m in (2^-300, 2^10), c in 1..63
    unsigned long long y = x * 0xD1342543DE82EF95ull + 1ull;
    unsigned long long em = 723ull + ((y >> 52) % 310ull);
    double m = __longlong_as_double((long long)((em << 52) | (y & 0x000FFFFFFFFFFFFFull)));
    int c = 1 + (int)((y >> 40) % 63ull);
    double cd = (double)c;
    if (__double_as_longlong(div_small_int(m, __ddiv_rn(1.0, cd), cd)) != __double_as_longlong(__ddiv_rn(m, cd))) ++bad;
    // floor of v in [0, 128)
    double v = (double)(y >> 11) * (1.0 / 70368744177664.0);  // 2^53 / 2^46 = 128
    if (floor_nonneg(v) != __double2int_rz(v)) ++bad;
  }
  if (bad) atomicAdd(mismatches, bad);
}

// The float32 filter against the exact float64 sequence, on the device (the hardware's own
// rcp.approx, not an emulation).  Per sample: residents' counts and sums of one GPU are drawn,
// folded into {A, B, kk, cst} exactly as fold_fast does, and one candidate term t is evaluated
// both ways.  Even iterations draw |t| log-uniformly over 2^-20 .. 2^10; odd iterations aim tmp
// at an integer +- 1e-4 (solving for t), which is where a wrong floor could hide.
//   out[0] = max |est - tmp| (bits of a non-negative double), out[1] = samples whose estimate is
//   outside the +-kFilterDelta band yet floors differently from tmp (must be 0), out[2] = samples
//   inside the band (these take the exact path in the kernel), out[3] = samples evaluated.
__global__ void selftest_filter_kernel(unsigned long long seed, int iters, unsigned long long* out) {
  unsigned long long x = seed + 0x9E3779B97F4A7C15ull * (blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x + 1);
  unsigned long long wrong = 0, band = 0, done = 0;
  double maxdev = 0.0;
  auto next = [&]() {
    x ^= x << 13;
    x ^= x >> 7;
    x ^= x << 17;
    return x;
  };
  auto unit = [&]() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); };  // [0, 1)
  auto dv = [](double m, int c) {
    if (c == 0) return 0.0;  // lut[0] = {0, 0}
    const double cd = (double)c;
    return div_small_int(m, __ddiv_rn(1.0, cd), cd);
  };
  for (int it = 0; it < iters; ++it) {
    const unsigned long long r0 = next();
    const bool small = (r0 & 15) < 11;
    const int n = small ? (int)((r0 >> 8) % 5) : (int)((r0 >> 8) % 31);
    const int p = small ? (int)((r0 >> 16) % 5) : (int)((r0 >> 16) % 31);
    const bool neg = (r0 >> 32) & 1;
    const double ns = (double)n * (1.0 - unit()), ps = (double)p * (1.0 - unit());
    // fold_fast's constants
    const double kk_n = dv((double)(n + 1), n + p + 1);
    const double cst_n = __dmul_rn(100.0, dv(__dmul_rn(__dsub_rn(1.0, kk_n), ps), p));
    const double kq_p = dv((double)n, n + p + 1);
    const double kk_p = __dsub_rn(1.0, kq_p);
    const double cst_p = __dmul_rn(100.0, dv(__dmul_rn(kq_p, ns), n));
    const double tn = __dadd_rn(cst_n, __dmul_rn(100.0, dv(__dmul_rn(kk_n, __dadd_rn(ns, 0.0)), n + 1)));
    const double tp = __dadd_rn(cst_p, __dmul_rn(100.0, dv(__dmul_rn(kk_p, __dadd_rn(ps, 0.0)), p + 1)));
    const double Bn = 100.0 * kk_n * __ddiv_rn(1.0, (double)(n + 1)), Bp = 100.0 * kk_p * __ddiv_rn(1.0, (double)(p + 1));
    const double A = neg ? tn : tp, B = neg ? Bn : Bp;
    if (!(A >= 0.0 && B >= 0.0 && A + B <= kFilterMaxSum)) continue;  // the kernel switches the filter off here
    double at;  // |t|
    if ((it & 1) == 0) {
      at = exp2(-20.0 + 30.0 * unit()) * (1.0 + unit());
    } else {
      const double vmax = neg ? 0.5 : 1.0;
      const double lo = ceil(A), hi = floor(A + B * vmax);
      if (!(hi >= lo) || B <= 0.0) continue;
      const double k = lo + floor(unit() * (hi - lo + 1.0));
      const unsigned sel = (unsigned)(next() & 7);
      const double eps = sel == 0 ? 0.0 : (sel == 1 ? 1e-7 : (sel == 2 ? -1e-7 : (unit() - 0.5) * 2e-4));
      double v = (k + eps - A) / B;
      if (!(v > 1e-6)) v = 1e-6;
      if (v > vmax) v = vmax;
      const double z = 1.0 / v;
      at = neg ? sqrt(z - 1.0) - 1.0 : z - 1.0;
      if (!(at >= 0.0)) at = 0.0;
    }
    const float t = (next() & 1) ? (float)at : -(float)at;
    if (!(fabsf(t) < 1.0e18f)) continue;
    // exact float64 sequence (the kernel's exact path)
    double w = fabs((double)t);
    if (neg) w = __dadd_rn(w, 1.0);
    if (neg) w = __dmul_rn(w, w);
    const double v = rcp_ge1(__dadd_rn(1.0, w));
    const int c = neg ? n + 1 : p + 1;
    const double S = __dadd_rn(neg ? ns : ps, v);
    const double q = dv(__dmul_rn(neg ? kk_n : kk_p, S), c);
    const double tmp = __dadd_rn(neg ? cst_n : cst_p, __dmul_rn(100.0, q));
    const int si = floor_nonneg(tmp);
    // float32 estimate
    const float est = filter_est((float)A, (float)B, filter_z(t, neg));
    const float fl = filter_floor(est);
    const float r = est - fl;
    const double dev = fabs((double)est - tmp);
    maxdev = dev > maxdev ? dev : maxdev;
    ++done;
    if (r >= kFilterDelta && r <= 1.0f - kFilterDelta) {
      if ((int)fl != si) ++wrong;
    } else {
      ++band;
    }
  }
  atomicMax(&out[0], (unsigned long long)__double_as_longlong(maxdev));
  if (wrong) atomicAdd(&out[1], wrong);
  atomicAdd(&out[2], band);
  atomicAdd(&out[3], done);
}

cudaError_t launch_selftest_filter(unsigned long long seed, int blocks, int iters, unsigned long long* d_out,
                                   cudaStream_t s) {
  note_launch();
  selftest_filter_kernel<<<blocks, 256, 0, s>>>(seed, iters, d_out);
  return cudaGetLastError();
}

double filter_delta() { return (double)kFilterDelta; }

cudaError_t launch_selftest(unsigned long long seed, int blocks, int iters, unsigned long long* d_mismatches,
                            cudaStream_t s) {
  note_launch();
  selftest_kernel<<<blocks, 256, 0, s>>>(seed, iters, d_mismatches);
  return cudaGetLastError();
}

}  // namespace gs

// score_pods.cu — the lanes-are-pods Score kernel (sm_100a).
//
// Reference: pkg/plugins/gpu_plugin/gpu_plugins.go:562-757 (the per-GPU loop of Logic).
//
// Same restructuring and the same float32 filter as score_fast.cu (read its header first); what
// changes is who owns what.  A CTA is ONE warp and takes one work item (up to kChunk pods of one
// app) x one slice of kSlice = 32 consecutive nodes:
//
//   phase 1  lane = node.  Each lane folds its node's eight GPUs (sums from the stage-A table) into
//            the float32 filter constants {A, B} per (GPU, branch), exactly as score_fast_kernel
//            does, and leaves them in shared memory — once per (work item, node).
//   phase 2  lane = (pod, node): the warp takes kPW = 8 pods x kNW = 4 nodes per pass, so that small
//            app buckets (most buckets hold fewer than 32 pods) waste few lanes.  A node's
//            constants are read as shared-memory broadcasts (four distinct records per load); the
//            per-pod values — SLO, |1/SLO|, the V100 configuration — stay in registers.
//
// Why this is cheaper than one thread per node:
//   * the evaluation loop runs over a LIST of the slice's nodes that need one — nodes without a
//     model or UUIDs (score 0, :558) and, for pods that no V100 configuration satisfies (:667), all
//     V100 nodes are answered from a per-slice template row instead;
//   * per-pod extrema stay in registers over the slice (no REDUX per pod, one pair of atomics per
//     (pod, slice)), and results go through an 8 x 32-byte staging tile and leave as full 32-byte
//     sectors instead of byte stores;
//   * two GPUs are evaluated per step with packed float32 instructions (FADD2/FMUL2/FFMA2), and the
//     branch is resolved by evaluating tmp = A + B*v for both branches in one FFMA2 and selecting.
//
// Work items hold pods of kind kKindPlain followed by pods of kind kKindNoV100 (ChunkDesc::kind
// carries the number of the former); pods without an SLO and pods with a non-finite SLO are
// separate work items that the one-thread-per-node kernel serves (they are few: after class
// dedupe at most one pod without an SLO per app).
#include "score_fold.cuh"

namespace gs {

// ------------------------------------------------------------------ stage: sums + fold ----

// stage_a_kernel (score_fast.cu) followed by the fold, for the lanes-are-pods kernel: a warp takes
// one GPU slot with its lanes on 32 consecutive apps (the slot's resident list is warp-uniform, both
// interference reads are contiguous), sums the residents' terms op for op (add_term) and folds the
// sums at once into what the pod loop needs (fold_fast_fn / fold_general_fn) — once per (app, slot)
// instead of once per (work item, node) inside the scoring kernel.  Written slot-contiguous through
// a padded shared tile.
constexpr int kFoldTile = 32;

__global__ void __launch_bounds__(256) stage_fold_kernel(DevTables T, const uint32_t* __restrict__ counts, FoldTab out) {
  __shared__ float4 s_ab[kFoldTile][kFoldTile + 1];
  __shared__ float s_ci[kFoldTile][kFoldTile + 1];
  __shared__ uint32_t s_pk[kFoldTile][kFoldTile + 1];
  __shared__ uint32_t s_lo[kFoldTile], s_hi[kFoldTile];
  __shared__ int s_model[kFoldTile];
  __shared__ double2 s_lut[kLutSize];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int slots = T.n_nodes * T.max_gpus;
  const int s0 = blockIdx.x * kFoldTile, a0 = blockIdx.y * kFoldTile;
  const int app_idx = a0 + lane;  // bucket n_apps = pods whose app is unknown
  const bool app_known = app_idx < T.n_apps;
  bool wanted = false;
  if (app_idx <= T.n_apps) {
    const uint4 c4 = *reinterpret_cast<const uint4*>(counts + (size_t)app_idx * kKinds);
    wanted = (c4.x | c4.y | c4.z | c4.w) != 0u;
  }
  if (threadIdx.x < kLutSize)
    s_lut[threadIdx.x] = make_double2(threadIdx.x ? __ddiv_rn(1.0, (double)threadIdx.x) : 0.0, (double)threadIdx.x);
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
  for (int j = 0; j < kFoldTile / 8; ++j) {
    const int sl = warp * (kFoldTile / 8) + j;
    GpuAcc acc = {0.0, 0.0, 0, 0};
    float ci = 0.f;
    const int model = s_model[sl];
    // a GPU the node does not have: estimate 0.5 on both sides — score 0, never near an integer
    float4 ab = make_float4(0.5f, 0.f, 0.5f, 0.f);
    uint32_t pk = 0u;
    if (model != kModelNone && s_lo[sl] == s_hi[sl]) {
      // a GPU without residents (warp-uniform): the fold of empty sums — tmp = 0 + 100*((1*(0+v))/1) on
      // either side, nothing without a candidate
      ab = make_float4(0.f, 100.f, 0.f, 100.f);
    } else if (model != kModelNone) {
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
      FoldOut o;
      uint32_t general = 0u;
      if (fold_in_domain(acc, ci)) {
        o = fold_fast_fn(acc, s_lut);
      } else {
        o = fold_general_fn(acc, s_lut);
        general = kFoldGeneral;
      }
      ab = make_float4(o.An, o.Bn, o.Ap, o.Bp);
      pk = (uint32_t)o.si | ((uint32_t)o.z_neg << 8) | ((uint32_t)o.z_pos << 16) | general;
    }
    s_ab[lane][sl] = ab;
    s_ci[lane][sl] = ci;
    s_pk[lane][sl] = pk;
  }
  __syncthreads();
#pragma unroll
  for (int pass = 0; pass < kFoldTile / 8; ++pass) {
    const int al = pass * 8 + warp;
    if (a0 + al <= T.n_apps && s0 + lane < slots) {
      const size_t o = (size_t)(a0 + al) * out.stride + (size_t)(s0 + lane);
      out.ab[o] = s_ab[al][lane];
      out.ci[o] = s_ci[al][lane];
      out.pk[o] = s_pk[al][lane];
    }
  }
}

cudaError_t launch_stage_fold(const DevTables& T, const uint32_t* counts, const StageA& stage, cudaStream_t s) {
  const int slots = T.n_nodes * T.max_gpus;
  if (slots == 0 || !stage.n0) return cudaSuccess;
  dim3 ga((unsigned)((slots + kFoldTile - 1) / kFoldTile), (unsigned)((T.n_apps + 1 + kFoldTile - 1) / kFoldTile));
  note_launch();
  stage_fold_kernel<<<ga, 256, 0, s>>>(T, counts, fold_view(stage));
  return cudaGetLastError();
}

// ------------------------------------------------------------------ scoring kernel --------

#ifndef GS_PODS_MIN_CTAS
#define GS_PODS_MIN_CTAS 20
#endif
constexpr int kPW = 8;         // pods per pass
constexpr int kNW = 32 / kPW;  // nodes per pass

struct NodeHdr {
  uint32_t nocand;  // score << 8 | selected GPU of the residents alone (:667-696)
  uint32_t flags;   // bits 0..7 selection-byte flags of the node; kHdrV100; bits 16..23 = GPUs of the node;
                    // bits 24..31 = the flags of a pod with a candidate term here (node flags | F6 on V100 | F7)
  float bound;      // |conf| (A30) + max_g |ci_g|: the node's part of the bound on |t|
  uint32_t keybase;  // lane_key(0, node): the node's part of the per-pod (score, lowest node) key
};
constexpr uint32_t kHdrV100 = 1u << 8;

// Row strides are padded so that the records of any four nodes out of eight consecutive ones (what a
// pass reads side by side) start in different banks: 80 B (20 words: 20k mod 32 is a different
// multiple of 4 for k = 0..7), 40 B (10 words), and 40-byte staging rows for the eight pods.
template <int GT>
struct PodsSmem {
  float4 abn[kSlice][GT / 2 + 1];  // per GPU pair {An[g], An[g+1], Bn[g], Bn[g+1]}
  float4 abp[kSlice][GT / 2 + 1];  // the same for the positive branch
  float st[kSlice][GT + 2];        // A30: ci - conf (= -(conf - ci), :716);  V100: ci (the pod's conf is subtracted per lane)
  uint32_t pk[kSlice][GT];         // FoldTab::pk of the node's GPUs (scores with a zero candidate term, for pods without an SLO)
  NodeHdr hdr[kSlice];
  double2 lut[kLutSize];
  uint8_t list[2][kSlice];         // [V100 nodes wanted?] nodes of the slice to evaluate (streamlined domain)
  uint8_t glist[2][kSlice];        // ... nodes to evaluate op for op (sums outside the domain)
  uint8_t tmpl_score[2][kSlice], tmpl_sel[2][kSlice];  // results of the nodes that are not evaluated
  uint8_t stage_score[kPW][kSlice + 8], stage_sel[kPW][kSlice + 8];
  uint32_t n_list[2], n_glist[2], key_c[2], neg_c[2];  // extrema of the template nodes
};

// The residents' sums of one GPU for a candidate of app `app`, op for op as stage_fold_kernel forms
// them (same order, same rounding).  Only the exact path needs them again, so they are recomputed
// there instead of being kept in a second table.
__device__ __forceinline__ GpuAcc resident_sums(const DevTables& T, uint32_t app, int node, int g) {
  GpuAcc acc = {0.0, 0.0, 0, 0};
  const int model = T.node_model[node];
  const size_t A = (size_t)T.n_apps;
  const bool app_known = app < (uint32_t)T.n_apps;
  const float* itab = T.interf + (size_t)(model - 1) * A * A;
  const uint32_t lo = T.gpu_off[(size_t)node * T.max_gpus + g], hi = T.gpu_off[(size_t)node * T.max_gpus + g + 1];
  for (uint32_t r = lo; r < hi; ++r) {
    const float4 rec = T.res_rec[r];
    if (rec.x != 0.0f) {
      float interf = rec.w;
      if (app_known) interf = __fadd_rn(interf, itab[(size_t)T.res_app[r] * A + app]);
      add_term(acc, rec.x, rec.y, rec.z, interf);
    }
  }
  return acc;
}

// One (pod, node, GPU) step of the exact float64 sequence, for the lanes whose float32 estimate is
// within kFilterDelta of an integer (same arithmetic as score_fast_kernel's exact path).
__device__ __forceinline__ int exact_gpu(const DevTables& T, uint32_t app, int node, int g, int k, float t, bool neg,
                                         const double2* lut) {
  if (g >= k) return 0;  // a GPU the node does not have never wins
  const GpuAcc acc = resident_sums(T, app, node, g);
  const double bn = acc.neg_sum, bp = acc.pos_sum;
  const int cn0 = acc.cnt_neg, cp0 = acc.cnt_pos;
  double w = fabs((double)t);
  if (neg) w = __dadd_rn(w, 1.0);
  if (neg) w = __dmul_rn(w, w);
  const double v = rcp_ge1(__dadd_rn(1.0, w));
  const double2 kc = branch_const(neg, bn, bp, cn0, cp0, lut);
  const double2 rc = lut[min(neg ? cn0 + 1 : cp0 + 1, kLutSize - 1)];
  const double S = __dadd_rn(neg ? bn : bp, v);
  const double q = div_small_int(__dmul_rn(kc.x, S), rc.x, rc.y);
  return floor_nonneg(__dadd_rn(kc.y, __dmul_rn(100.0, q)));
}

// Redo one pass GPU by GPU: float32 estimate first, the exact sequence for the lanes whose estimate
// is near an integer.  Returns max over g of int64(tmp) * 8 + (GT - 1 - g).
template <int GT>
static __device__ __noinline__ int redo_pass(const PodsSmem<GT>* sm, const DevTables& T, uint32_t app, int node, int n,
                                             int k, float s, float arinv, float negconf, bool has_cand) {
  int best = 0;
#pragma unroll 1
  for (int g = 0; g < GT; ++g) {
    const float nd = __fadd_rn(sm->st[n][g], negconf);
    const float e = __fadd_rn(s, nd);
    const float t = __fmul_rn(arinv, e);
    const bool neg = e > 0.0f;
    const float4 ab = neg ? sm->abn[n][g >> 1] : sm->abp[n][g >> 1];
    const float A = (g & 1) ? ab.y : ab.x, B = (g & 1) ? ab.w : ab.z;
    const float est = filter_est(A, B, filter_z(t, neg));
    const float fl = filter_floor(est);
    const float r = est - fl;
    int si = (int)fl;
    if (has_cand && !(r >= kFilterDelta && r <= 1.0f - kFilterDelta)) si = exact_gpu(T, app, node, g, k, t, neg, sm->lut);
    best = max(best, si * 8 + (GT - 1 - g));
  }
  return best;
}

template <int GT>
__global__ void __launch_bounds__(32, GS_PODS_MIN_CTAS) score_pods_kernel(DevTables T, FoldTab fold, const PodRec* __restrict__ pods,
                                                             const ChunkDesc* __restrict__ chunks,
                                                             uint32_t* __restrict__ n_chunks,
                                                             uint8_t* __restrict__ out_score,
                                                             uint8_t* __restrict__ out_sel, unsigned long long* hi_arg,
                                                             unsigned long long* neg_lo) {
  __shared__ __align__(16) PodsSmem<GT> sm_storage;
  PodsSmem<GT>* sm = &sm_storage;
  const int lane = threadIdx.x;
  for (int i = lane; i < kLutSize; i += 32) sm->lut[i] = make_double2(i ? __ddiv_rn(1.0, (double)i) : 0.0, (double)i);
  const uint32_t n_slices = (uint32_t)((T.n_nodes + kSlice - 1) / kSlice);
  const uint32_t n_items = n_chunks[0] * n_slices;
  const size_t N = (size_t)T.n_nodes;
  const bool vec8 = (N % 8 == 0) && ((((uintptr_t)out_score) | ((uintptr_t)out_sel)) % 8 == 0);
  const int pi = lane & (kPW - 1), q = lane >> 3;
  // Persistent: the warps of the grid take (work item, node slice) pairs from a counter until none is
  // left (n_chunks[1], zeroed by pod_scan_kernel) — consecutive pairs are consecutive slices of one
  // work item, so the fold table is read, and the result rows are written, in address order.
#pragma unroll 1
  for (;;) {
    uint32_t item = 0;
    if (lane == 0) item = atomicAdd(&n_chunks[1], 1u);
    item = __shfl_sync(0xFFFFFFFFu, item, 0);
    if (item >= n_items) break;
    const ChunkDesc cd = chunks[item / n_slices];
    const int node0 = (int)(item % n_slices) * kSlice;
    const int n_in = min(kSlice, T.n_nodes - node0);
    const uint32_t app = cd.app;
    const bool app_known = app < (uint32_t)T.n_apps;
    __syncwarp();  // the previous pair's reads of the shared tables are done (and, first time, the lut is written)

    // -------------------------------------------------------------- phase 1: lane = node
    {
      const int node = node0 + lane;
      const bool in_range = lane < n_in;
      int model = kModelNone, k = 0;
      if (in_range && T.redis_reachable) {
        model = T.node_model[node];
        k = T.node_ngpu[node];
      }
      const bool scored = model != kModelNone && k > 0;
      unsigned node_flags = scored ? kFlagEligible : 0u;
      float conf_a30 = 0.0f, ci_max = 0.0f;
      bool general = false;
      int nc_score = 0;
      unsigned nc_sel = kSelNone;
      float4 ab[GT];
      float ci[GT];
      uint32_t pk[GT];
#pragma unroll
      for (int g = 0; g < GT; ++g) {
        ab[g] = make_float4(0.5f, 0.f, 0.5f, 0.f);  // estimate 0.5: score 0, never near an integer
        ci[g] = 0.f;
        pk[g] = 0u;
      }
      if (scored) {
        if (model == kModelA30) {
          float v;
          const bool ok = conf_lookup(T, app, model, k, &v);
          conf_a30 = ok ? v : 0.0f;  // missing key -> Go zero value, gpu_plugins.go:637
          if (ok) node_flags |= kFlagF6;
        }
        const size_t row = (size_t)(app_known ? app : (uint32_t)T.n_apps) * fold.stride + (size_t)node * T.max_gpus;
#pragma unroll
        for (int g = 0; g < GT; ++g) {
          if (g < k) {  // read once: streaming loads
            ab[g] = __ldcs(fold.ab + row + g);
            ci[g] = __ldcs(fold.ci + row + g);
            pk[g] = __ldcs(fold.pk + row + g);
          }
        }
#pragma unroll
        for (int g = 0; g < GT; ++g) {
          if (g < k) {
            const int si = (int)(pk[g] & 0xFFu);
            if (si > nc_score) {  // strict: the first GPU among equals (:691-694)
              nc_score = si;
              nc_sel = (unsigned)g;
            }
            general = general || (pk[g] & kFoldGeneral) != 0u;
            ci_max = fmaxf(ci_max, fabsf(ci[g]));
          }
        }
      }
      const float negconf = model == kModelA30 ? -conf_a30 : -0.0f;
#pragma unroll
      for (int gp = 0; gp < GT / 2; ++gp) {
        sm->abn[lane][gp] = make_float4(ab[2 * gp].x, ab[2 * gp + 1].x, ab[2 * gp].y, ab[2 * gp + 1].y);
        sm->abp[lane][gp] = make_float4(ab[2 * gp].z, ab[2 * gp + 1].z, ab[2 * gp].w, ab[2 * gp + 1].w);
      }
#pragma unroll
      for (int g = 0; g < GT; ++g) {
        sm->st[lane][g] = __fadd_rn(ci[g], negconf);  // x + (-0) = x: V100 rows keep ci
        sm->pk[lane][g] = pk[g];
      }
      NodeHdr h;
      h.nocand = ((uint32_t)nc_score << 8) | nc_sel;
      h.flags = node_flags | (model == kModelV100 ? kHdrV100 : 0u) | ((uint32_t)k << 16) |
                ((node_flags | (model == kModelV100 ? kFlagF6 : 0u) | kFlagF7) << 24);
      h.bound = (model == kModelA30 ? fabsf(conf_a30) : 0.0f) + ci_max;
      h.keybase = lane_key(0, (uint32_t)node);
      sm->hdr[lane] = h;
      // evaluation lists and template rows: [0] every node with a GPU loop, [1] A30 nodes only (pods that
      // no V100 configuration satisfies have no candidate term on V100 nodes, :667).
      // F6 on V100 nodes = some 1P/2P/4P_V100 column exists for this app (gpu_plugins.go:639-651)
      bool f6_app = false;
      {
        float tmp;
        for (int i = 0; i < 3; ++i) f6_app |= conf_lookup(T, app, kModelV100, 1 << i, &tmp);
      }
      const bool is_v = model == kModelV100;
#pragma unroll
      for (int li = 0; li < 2; ++li) {
        const bool cand = scored && !(li == 1 && is_v);
        const unsigned m_fast = __ballot_sync(0xFFFFFFFFu, cand && !general);
        const unsigned m_gen = __ballot_sync(0xFFFFFFFFu, cand && general);
        const unsigned lt = (1u << lane) - 1u;
        if (cand && !general) sm->list[li][__popc(m_fast & lt)] = (uint8_t)lane;
        if (cand && general) sm->glist[li][__popc(m_gen & lt)] = (uint8_t)lane;
        const bool tmpl = in_range && !cand;
        sm->tmpl_score[li][lane] = (uint8_t)nc_score;
        sm->tmpl_sel[li][lane] = (uint8_t)(nc_sel | node_flags | ((is_v && scored && f6_app) ? kFlagF6 : 0u));
        uint32_t key = tmpl ? lane_key(nc_score, (uint32_t)node) : 0u;
        uint32_t ng = tmpl ? (uint32_t)(100 - nc_score) : 0u;
        key = __reduce_max_sync(0xFFFFFFFFu, key);
        ng = __reduce_max_sync(0xFFFFFFFFu, ng);
        if (lane == 0) {
          sm->n_list[li] = __popc(m_fast);
          sm->n_glist[li] = __popc(m_gen);
          sm->key_c[li] = key;
          sm->neg_c[li] = ng;
        }
      }
    }
    // the slice's largest node bound: lets whole passes skip the per-lane overflow check
    float slice_bound = sm->hdr[lane].bound;  // own write, visible without a barrier
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) slice_bound = fmaxf(slice_bound, __shfl_xor_sync(0xFFFFFFFFu, slice_bound, d));
    if (!(slice_bound <= 3.0e38f)) slice_bound = __int_as_float(0x7F800000);  // NaN anywhere -> +Inf: always check
    __syncwarp();

    // -------------------------------------------------------------- phase 2: lane = (pod, node)
    const float2 one2 = make_float2(1.0f, 1.0f), mone2 = make_float2(-1.0f, -1.0f);
    const float2 magic2 = make_float2(8388608.0f, 8388608.0f), nmagic2 = make_float2(-8388607.5f, -8388607.5f);
    const float2 eight2 = make_float2(8.0f, 8.0f);
    // the work item's pods by kind: [0, e0) plain, [e0, e1) no V100 configuration, [e1, e2) no SLO, [e2, len) general
    const uint32_t e0 = cd.kind & 0x1FFu, e1 = e0 + ((cd.kind >> 9) & 0x1FFu), e2 = e1 + ((cd.kind >> 18) & 0x1FFu);
#pragma unroll 1
    for (int kind = 0; kind < kKinds; ++kind) {
    const uint32_t seg0 = kind == 0 ? 0u : (kind == 1 ? e0 : (kind == 2 ? e1 : e2));
    const uint32_t seg_end = kind == 0 ? e0 : (kind == 1 ? e1 : (kind == 2 ? e2 : cd.len));
    const int li = kind == kKindNoV100 ? 1 : 0;
#pragma unroll 1
    // Pods without an SLO are few (after dedupe one per app): their passes take ONE pod x 32 nodes instead
    // of 8 pods x 4 nodes, so that no lane idles; pk/qk/nwk are the lane's pod row, node lane and nodes per pass
    const bool pod_major = kind == kKindInfInv;
    const int pk = pod_major ? 0 : pi, qk = pod_major ? lane : q, nwk = pod_major ? 32 : kNW;
    for (uint32_t base = seg0; base < seg_end; base += (pod_major ? 1 : kPW)) {  // a pass never mixes kinds
    const bool valid = base + pk < seg_end;
    const PodRec pr = pods[cd.start + (valid ? base + pk : base)];
    const uint32_t orig = pr.orig & kPodIndexMask;
    const float s = pr.slo, arinv = fabsf(pr.inv_slo);
    const float ncv = kind == kKindNoV100 ? -0.0f : -pr.conf_v100;  // that kind never evaluates a V100 node
    // may any lane's float32 term leave the finite range on some node of the slice?
    const bool slow = __any_sync(0xFFFFFFFFu, !(arinv * (fabsf(s) + fabsf(pr.conf_v100) + slice_bound) < 1.0e38f));
    // template rows into the staging tile: lane -> pod row lane >> 2, 8 nodes each
    {
      const int row = lane >> 2, c8 = (lane & 3) * 8;
      *reinterpret_cast<uint2*>(&sm->stage_score[row][c8]) = *reinterpret_cast<const uint2*>(&sm->tmpl_score[li][c8]);
      *reinterpret_cast<uint2*>(&sm->stage_sel[row][c8]) = *reinterpret_cast<const uint2*>(&sm->tmpl_sel[li][c8]);
    }
    __syncwarp();
    uint32_t hi = 0u, ng = 0u;
    const int nL = (int)sm->n_list[li];
    if (kind <= kKindNoV100) {
#pragma unroll 1
    for (int it = 0; it < nL; it += kNW) {
      const bool active = it + q < nL;
      const int n = sm->list[li][active ? it + q : it];
      const NodeHdr h = sm->hdr[n];
      const bool is_v = (h.flags & kHdrV100) != 0u;
      const bool has_cand = active && valid;
      const float negconf = is_v ? ncv : -0.0f;
      const float2 negconf2 = make_float2(negconf, negconf), s2 = make_float2(s, s), ar2 = make_float2(arinv, arinv);
      int best;
      bool lane_general = false;
      if (slow) lane_general = has_cand && !(arinv * (fabsf(s) + (is_v ? fabsf(pr.conf_v100) : 0.0f) + h.bound) < 1.0e38f);
      if (slow && __any_sync(0xFFFFFFFFu, lane_general)) {
        // rare: some lane's term may overflow — every candidate lane of this pass goes op for op
        best = 0;
        if (has_cand) {
          unsigned selbyte;
          const int sc = logic_eval(T, app, s, -1, node0 + n, &selbyte);
          const unsigned g = selbyte & 31u;
          best = sc > 0 ? sc * 8 + (GT - 1 - (int)g) : 0;
        }
      } else {
        float fbest = 0.f, rdev = 0.f;  // rdev = max over GPUs of |frac(est) - 0.5|
#pragma unroll
        for (int gp = 0; gp < GT / 2; ++gp) {
          const float2 st2 = *reinterpret_cast<const float2*>(&sm->st[n][2 * gp]);
          const float4 abn = sm->abn[n][gp], abp = sm->abp[n][gp];
          const float2 e2 = __fadd2_rn(s2, __fadd2_rn(st2, negconf2));  // slo - (conf - interf), :716
          const float2 t2 = __fmul2_rn(ar2, e2);                        // |1/slo| * (slo - d): |t| with the sign of e
          const float a0 = 1.0f + fabsf(t2.x), a1 = 1.0f + fabsf(t2.y);
          const float2 zn2 = __ffma2_rn(make_float2(a0, a1), make_float2(a0, a1), one2);
          const bool neg0 = e2.x > 0.0f, neg1 = e2.y > 0.0f;           // SLO > conf - interf
          float v0, v1;
          asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(v0) : "f"(neg0 ? zn2.x : a0));
          asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(v1) : "f"(neg1 ? zn2.y : a1));
          const float2 v2 = make_float2(v0, v1);
          const float2 estn = __ffma2_rn(make_float2(abn.z, abn.w), v2, make_float2(abn.x, abn.y));
          const float2 estp = __ffma2_rn(make_float2(abp.z, abp.w), v2, make_float2(abp.x, abp.y));
          const float2 est = make_float2(neg0 ? estn.x : estp.x, neg1 ? estn.y : estp.y);
          // flh = floor(est) + 0.5 (exact for 0 <= est < 2^22: 2^23 - 0.5 is a float), rc = frac(est) - 0.5 (exact)
          const float2 flh = __fadd2_rn(__fadd2_rz(est, magic2), nmagic2);
          const float2 rc = __ffma2_rn(flh, mone2, est);
          // key = floor(est) * 8 + (GT - 1 - g) = flh * 8 + (GT - 5 - g)
          const float2 key = __ffma2_rn(flh, eight2, make_float2((float)(GT - 5 - 2 * gp), (float)(GT - 6 - 2 * gp)));
          rdev = fmaxf(fmaxf(rdev, fabsf(rc.x)), fabsf(rc.y));
          fbest = fmaxf(fmaxf(fbest, key.x), key.y);
        }
        best = __float_as_int(fbest + 8388608.0f) & 0x7FFFFF;  // fbest is a small non-negative integer: no F2I
        const bool unsure = has_cand && !(rdev <= 0.5f - kFilterDelta);  // some estimate within kFilterDelta of an integer
        if (__any_sync(0xFFFFFFFFu, unsure)) {
          best = redo_pass<GT>(sm, T, app, node0 + n, n, (int)((h.flags >> 16) & 0xFFu), s, arinv, negconf, has_cand);
        }
      }
      const int score = best >> 3;
      const unsigned sel = score > 0 ? (unsigned)(GT - 1 - (best & 7)) : kSelNone;
      if (has_cand) {
        sm->stage_score[pi][n] = (uint8_t)score;
        sm->stage_sel[pi][n] = (uint8_t)(sel | (h.flags >> 24));
        hi = max(hi, ((uint32_t)score << 24) + h.keybase);
        ng = max(ng, (uint32_t)(100 - score));
      }
    }
    } else if (kind == kKindInfInv) {
      // 1/SLO is infinite: t = +-Inf * (slo - d), so the term is exactly 0 on either side (1/(1+Inf)), or
      // NaN when slo - d is 0 or NaN — int64(tmp) per (GPU, side) is the fold's z_neg / z_pos
#pragma unroll 1
      for (int it = 0; it < nL; it += nwk) {
        if (it + qk < nL && valid) {
          const int n = sm->list[0][it + qk];
          const NodeHdr h = sm->hdr[n];
          const bool is_v = (h.flags & kHdrV100) != 0u;
          const bool has_cand = !is_v || pr.conf_v100 != -1.0f;
          const float negconf = is_v ? -pr.conf_v100 : -0.0f;
          int score = (int)(h.nocand >> 8);
          unsigned sel = h.nocand & 0xFFu;
          if (has_cand) {
            score = 0;
            sel = kSelNone;
#pragma unroll
            for (int g = 0; g < GT; ++g) {
              const float nd = __fadd_rn(sm->st[n][g], negconf);  // -(conf - interf)
              const float e = __fadd_rn(s, nd);                   // slo - (conf - interf)
              const uint32_t z = sm->pk[n][g];
              int si = s > -nd ? (int)((z >> 8) & 0xFFu) : (int)((z >> 16) & 0xFFu);
              if (!(e < 0.0f || e > 0.0f)) si = 0;  // NaN term: int64(NaN) never wins
              if (si > score) {
                score = si;
                sel = (unsigned)g;
              }
            }
          }
          sm->stage_score[pk][n] = (uint8_t)score;
          sm->stage_sel[pk][n] = (uint8_t)(sel | (h.flags & 0xFFu) | ((is_v && (pr.orig & kPodF6)) ? kFlagF6 : 0u) |
                                           (has_cand ? kFlagF7 : 0u));
          hi = max(hi, lane_key(score, (uint32_t)(node0 + n)));
          ng = max(ng, (uint32_t)(100 - score));
        }
      }
    } else {
      // SLO is NaN or infinite (or its reciprocal is 0): every node with a GPU loop, op for op
#pragma unroll 1
      for (int it = 0; it < nL; it += kNW) {
        if (it + q < nL && valid) {
          const int n = sm->list[0][it + q];
          unsigned selbyte;
          const int sc = logic_eval(T, app, s, -1, node0 + n, &selbyte);
          sm->stage_score[pi][n] = (uint8_t)sc;
          sm->stage_sel[pi][n] = (uint8_t)selbyte;
          hi = max(hi, lane_key(sc, (uint32_t)(node0 + n)));
          ng = max(ng, (uint32_t)(100 - sc));
        }
      }
    }
    // nodes whose sums fall outside the streamlined domain: op for op
    const int nG = (int)sm->n_glist[li];
#pragma unroll 1
    for (int it = 0; it < nG; it += nwk) {
      if (it + qk < nG && valid) {
        const int n = sm->glist[li][it + qk];
        unsigned selbyte;
        const int sc = logic_eval(T, app, s, -1, node0 + n, &selbyte);
        sm->stage_score[pk][n] = (uint8_t)sc;
        sm->stage_sel[pk][n] = (uint8_t)selbyte;
        hi = max(hi, lane_key(sc, (uint32_t)(node0 + n)));
        ng = max(ng, (uint32_t)(100 - sc));
      }
    }
    if (pod_major) {  // all 32 lanes hold nodes of the one pod
#pragma unroll
      for (int dd = 1; dd < 8; dd <<= 1) {
        hi = max(hi, __shfl_xor_sync(0xFFFFFFFFu, hi, dd));
        ng = max(ng, __shfl_xor_sync(0xFFFFFFFFu, ng, dd));
      }
    }
    // per-pod extrema: the four node lanes of a pod, plus the template nodes
    hi = max(hi, __shfl_xor_sync(0xFFFFFFFFu, hi, 8));
    ng = max(ng, __shfl_xor_sync(0xFFFFFFFFu, ng, 8));
    hi = max(hi, __shfl_xor_sync(0xFFFFFFFFu, hi, 16));
    ng = max(ng, __shfl_xor_sync(0xFFFFFFFFu, ng, 16));
    hi = max(hi, sm->key_c[li]);
    ng = max(ng, sm->neg_c[li]);
    if ((pod_major ? lane == 0 : q == 0) && valid) publish_extrema(hi, ng, T.node_base, orig, hi_arg, neg_lo);
    __syncwarp();
    // write the tile out: lane -> pod row lane >> 2, nodes [8 * (lane & 3), +8)
    {
      const int row = lane >> 2, c8 = (lane & 3) * 8;
      uint32_t orow = __shfl_sync(0xFFFFFFFFu, orig, row);             // lanes 0..7 hold the eight pods
      bool rvalid = __shfl_sync(0xFFFFFFFFu, (int)valid, row) != 0;
      if (pod_major) {                                                  // one pod: staging row 0 only
        orow = orig;
        rvalid = row == 0;
      }
      if (rvalid && c8 < n_in) {
        const size_t o = (size_t)orow * N + (size_t)(node0 + c8);
        if (vec8) {
          *reinterpret_cast<uint2*>(out_score + o) = *reinterpret_cast<const uint2*>(&sm->stage_score[row][c8]);
          *reinterpret_cast<uint2*>(out_sel + o) = *reinterpret_cast<const uint2*>(&sm->stage_sel[row][c8]);
        } else {
          const int m = min(8, n_in - c8);
          for (int j = 0; j < m; ++j) {
            out_score[o + j] = sm->stage_score[row][c8 + j];
            out_sel[o + j] = sm->stage_sel[row][c8 + j];
          }
        }
      }
    }
    __syncwarp();
    }  // passes of one kind
    }  // kinds
  }  // (work item, slice) pairs
}

// every work item through the lanes-are-pods kernel; the grid is sized to fill the GPU and the warps
// share the (work item, slice) pairs through a counter
cudaError_t launch_score_pods(const DevTables& T, const PodScratch& sc, uint8_t* score, uint8_t* sel,
                              unsigned long long* hi_arg, unsigned long long* neg_lo, cudaStream_t s) {
  static int n_sm = 0;
  if (n_sm == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n_sm <= 0) n_sm = 148;
  }
  const unsigned grid = (unsigned)n_sm * (unsigned)GS_PODS_MIN_CTAS;  // as many one-warp CTAs as stay resident
  const FoldTab fold = fold_view(sc.stage);
  note_launch();
  if (T.max_gpus <= 2)
    score_pods_kernel<2><<<grid, 32, 0, s>>>(T, fold, sc.sorted, sc.chunks, sc.n_chunks, score, sel, hi_arg, neg_lo);
  else if (T.max_gpus <= 4)
    score_pods_kernel<4><<<grid, 32, 0, s>>>(T, fold, sc.sorted, sc.chunks, sc.n_chunks, score, sel, hi_arg, neg_lo);
  else
    score_pods_kernel<8><<<grid, 32, 0, s>>>(T, fold, sc.sorted, sc.chunks, sc.n_chunks, score, sel, hi_arg, neg_lo);
  return cudaGetLastError();
}

}  // namespace gs

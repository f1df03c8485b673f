// score_kernels.cu — sm_100a kernels for Score -> Logic -> NormalizeScore.
//
// Reference: pkg/plugins/gpu_plugin/gpu_plugins.go:455-777 (Logic), :816-841
// (NormalizeScore), :366-399 (MIG choice in reconfigure).
//
// Two scoring kernels produce identical bytes:
//   score_direct_kernel  one thread per (pod,node), walks the GPU and resident lists
//                        exactly like Logic() does; handles every case (self-exclusion,
//                        up to 31 GPUs per node).
//   score_fast_kernel    the restructured path.  A resident's term depends on the pod
//                        being scheduled only through that pod's APP (its interference
//                        column, :607-612), so pods are bucketed by app; one thread owns a
//                        (node, app) pair, builds the node's per-GPU running sums
//                        (negative_sum, positive_sum, counts, candidate interference)
//                        ONCE in registers, then streams the bucket's pods through them.
//                        Per pod and GPU only the candidate's own term and the weighted
//                        combine remain.
#include "score_common.cuh"
#include "context.h"

namespace gs {

// ------------------------------------------------------------------ resident prep ----

// Per resident r of GPU slot s: the part of its term that does not depend on the pod
// being scheduled.  gpu_plugins.go:572 (SLO==0 skip), :583-587 (conf column), :595-606
// (co-resident interference, float32, in resident order).
__global__ void resident_prep_kernel(DevTables T, float4* __restrict__ res_rec) {
  int slot = blockIdx.x * blockDim.x + threadIdx.x;
  if (slot >= T.n_nodes * T.max_gpus) return;
  int node = slot / T.max_gpus;
  int model = T.node_model[node];
  int k = T.node_ngpu[node];
  uint32_t lo = T.gpu_off[slot], hi = T.gpu_off[slot + 1];
  int col = (model != kModelNone && k <= kMaxGpusPerNode) ? T.col_of[(model - 1) * (kMaxGpusPerNode + 1) + k] : -1;
  size_t A = (size_t)T.n_apps;
  for (uint32_t r = lo; r < hi; ++r) {
    uint32_t app = T.res_app[r];
    float slo = T.res_slo[r];
    bool valid = model != kModelNone && slo != 0.0f && app < (uint32_t)T.n_apps && col >= 0;
    float4 rec = make_float4(0.f, 0.f, 0.f, 0.f);
    if (valid) {
      const float* row = T.interf + (size_t)(model - 1) * A * A + (size_t)app * A;
      float S = 0.f;
      for (uint32_t c = lo; c < hi; ++c) {
        if (c == r) continue;
        uint32_t ac = T.res_app[c];
        if (ac < (uint32_t)T.n_apps) S = __fadd_rn(S, row[ac]);
      }
      rec = make_float4(slo, __fdiv_rn(1.0f, slo), T.conf[(size_t)app * T.n_cols + col], S);
    }
    res_rec[r] = rec;
  }
}

cudaError_t launch_resident_prep(const DevTables& T, float4* res_rec, cudaStream_t s) {
  int slots = T.n_nodes * T.max_gpus;
  if (slots == 0) return cudaSuccess;
  note_launch();
  resident_prep_kernel<<<(slots + 127) / 128, 128, 0, s>>>(T, res_rec);
  return cudaGetLastError();
}

// ------------------------------------------------------------------ direct kernel ----

// Logic() for one (pod,node); grid = (pods, node tiles).
__global__ void __launch_bounds__(kTile) score_direct_kernel(DevTables T, const uint32_t* __restrict__ pod_app,
                                                            const float* __restrict__ pod_slo,
                                                            const int32_t* __restrict__ pod_self,
                                                            uint8_t* __restrict__ out_score,
                                                            uint8_t* __restrict__ out_sel,
                                                            unsigned long long* hi_arg,
                                                            unsigned long long* neg_lo) {
  __shared__ uint32_t s_key[kTile / 32], s_neg[kTile / 32];
  const uint32_t p = blockIdx.x;
  const int node = blockIdx.y * kTile + threadIdx.x;
  const bool in_range = node < T.n_nodes;
  int score = 0;
  uint32_t norig = 0;
  if (in_range) {
    unsigned selbyte;
    norig = T.node_orig[node];
    score = logic_eval(T, pod_app[p], pod_slo[p], pod_self ? pod_self[p] : -1, node, &selbyte);
    size_t o = (size_t)p * T.n_nodes + norig;
    out_score[o] = (uint8_t)score;
    out_sel[o] = (uint8_t)selbyte;
  }
  // per-pod extrema over this tile
  uint32_t key = in_range ? lane_key(score, norig) : 0u;
  uint32_t neg = in_range ? (uint32_t)(100 - score) : 0u;
  key = __reduce_max_sync(0xFFFFFFFFu, key);
  neg = __reduce_max_sync(0xFFFFFFFFu, neg);
  if ((threadIdx.x & 31) == 0) {
    s_key[threadIdx.x >> 5] = key;
    s_neg[threadIdx.x >> 5] = neg;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int w = 1; w < kTile / 32; ++w) {
      key = max(key, s_key[w]);
      neg = max(neg, s_neg[w]);
    }
    publish_extrema(key, neg, T.node_base, p, hi_arg, neg_lo);
  }
}

cudaError_t launch_score_direct(const DevTables& T, const uint32_t* pod_app, const float* pod_slo,
                                const int32_t* pod_self, int n_pods, uint8_t* score, uint8_t* sel,
                                unsigned long long* hi_arg, unsigned long long* neg_lo, cudaStream_t s) {
  if (n_pods == 0 || T.n_nodes == 0) return cudaSuccess;
  dim3 grid((unsigned)n_pods, (unsigned)((T.n_nodes + kTile - 1) / kTile));
  note_launch();
  score_direct_kernel<<<grid, kTile, 0, s>>>(T, pod_app, pod_slo, pod_self, score, sel, hi_arg, neg_lo);
  return cudaGetLastError();
}

// ------------------------------------------------------------------ class dedupe -----

// Scores depend on a pending pod only through (app, SLO) (gpu_plugins.go:626-651 candidate
// configuration, :698-723 candidate term, :607-612 interference column) — pods of one Deployment
// share both (deploy/busybox/busybox.yaml:26-28).  Pods with the same app and the same SLO BITS form
// a class; only the lowest-index pod of a class is scored, the others receive copies of its rows.
// Open addressing with linear probing; the table is at least twice the batch, so probes are short.
constexpr unsigned long long kEmptyKey = ~0ull;

__device__ __forceinline__ uint32_t hash_key(unsigned long long k) {
  k ^= k >> 33;
  k *= 0xFF51AFD7ED558CCDull;
  k ^= k >> 33;
  k *= 0xC4CEB9FE1A85EC53ull;
  k ^= k >> 33;
  return (uint32_t)k;
}

// pass 1: rep[p] = slot of the pod's class in the table; hvals[slot] = lowest pod index of the class
__global__ void class_insert_kernel(const uint32_t* __restrict__ pod_app, const float* __restrict__ pod_slo, int n_pods,
                                    unsigned long long* hkeys, uint32_t* hvals, uint32_t mask, uint32_t* rep) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_pods) return;
  const unsigned long long key = ((unsigned long long)pod_app[p] << 32) | (unsigned long long)__float_as_uint(pod_slo[p]);
  if (key == kEmptyKey) {  // the one key that collides with the empty marker stays its own class
    rep[p] = kNoBucket;
    return;
  }
  uint32_t h = hash_key(key) & mask;
  for (;;) {
    const unsigned long long prev = atomicCAS(&hkeys[h], kEmptyKey, key);
    if (prev == kEmptyKey || prev == key) break;
    h = (h + 1) & mask;
  }
  atomicMin(&hvals[h], (uint32_t)p);
  rep[p] = h;
}

// pass 2: rep[p] = representative pod of p's class
__global__ void class_lookup_kernel(int n_pods, const uint32_t* __restrict__ hvals, uint32_t* rep) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_pods) return;
  const uint32_t slot = rep[p];
  rep[p] = slot == kNoBucket ? (uint32_t)p : hvals[slot];
}

// Copy a representative's finished rows (raw score, selection byte) and extrema to every other pod
// of its class: one warp per pod, widest vector the two rows' alignment allows.
__device__ __forceinline__ void copy_row(uint8_t* __restrict__ dst, const uint8_t* __restrict__ src, int n, int lane) {
  const uintptr_t a = (uintptr_t)dst, b = (uintptr_t)src;
  const int rel = (int)((a ^ b) & 15);
  if ((rel & 3) != 0) {
    for (int i = lane; i < n; i += 32) dst[i] = src[i];
    return;
  }
  const int w = rel == 0 ? 16 : (rel == 8 ? 8 : 4);  // bytes per vector both pointers can be aligned to together
  int head = (int)((w - (a & (w - 1))) & (w - 1));
  if (head > n) head = n;
  for (int i = lane; i < head; i += 32) dst[i] = src[i];
  const int nvec = (n - head) / w;
  if (w == 16) {
    const uint4* s4 = (const uint4*)(src + head);
    uint4* d4 = (uint4*)(dst + head);
    for (int i = lane; i < nvec; i += 32) d4[i] = s4[i];
  } else if (w == 8) {
    const uint2* s2 = (const uint2*)(src + head);
    uint2* d2 = (uint2*)(dst + head);
    for (int i = lane; i < nvec; i += 32) d2[i] = s2[i];
  } else {
    const uint32_t* s1 = (const uint32_t*)(src + head);
    uint32_t* d1 = (uint32_t*)(dst + head);
    for (int i = lane; i < nvec; i += 32) d1[i] = s1[i];
  }
  for (int i = head + nvec * w + lane; i < n; i += 32) dst[i] = src[i];
}

__global__ void __launch_bounds__(128) expand_classes_kernel(int n_pods, int n_nodes, const uint32_t* __restrict__ rep,
                                                            uint8_t* score, uint8_t* sel, unsigned long long* hi_arg,
                                                            unsigned long long* neg_lo) {
  const int lane = threadIdx.x & 31;
  const int p = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (p >= n_pods) return;
  const uint32_t r = rep[p];
  if (r == (uint32_t)p) return;
  copy_row(score + (size_t)p * n_nodes, score + (size_t)r * n_nodes, n_nodes, lane);
  copy_row(sel + (size_t)p * n_nodes, sel + (size_t)r * n_nodes, n_nodes, lane);
  if (lane == 0) {
    hi_arg[p] = hi_arg[r];
    neg_lo[p] = neg_lo[r];
  }
}

cudaError_t launch_expand_classes(const DevTables& T, const uint32_t* rep, int n_pods, uint8_t* score, uint8_t* sel,
                                  unsigned long long* hi_arg, unsigned long long* neg_lo, cudaStream_t s) {
  if (!rep || n_pods == 0) return cudaSuccess;
  note_launch();
  expand_classes_kernel<<<(n_pods + 3) / 4, 128, 0, s>>>(n_pods, T.n_nodes, rep, score, sel, hi_arg, neg_lo);
  return cudaGetLastError();
}

// ------------------------------------------------------------------ pod bucketing ----

// the fields of a pod that the scoring kernels need, and its kind
__device__ __forceinline__ PodRec make_pod_rec(const DevTables& T, uint32_t app, float slo, uint32_t p, int* kind) {
  bool f6;
  PodRec rec;
  rec.slo = slo;
  rec.inv_slo = __fdiv_rn(1.0f, slo);
  rec.conf_v100 = conf_v100_for(T, app, slo, &f6);
  // plain (finite SLO with a finite non-zero reciprocal), infinite reciprocal (SLO missing = 0, or
  // denormal), or general
  const bool inf_inv = fabsf(rec.inv_slo) > 3.4e38f;                            // +-Inf; false for NaN
  const bool plain = fabsf(slo) <= 3.4e38f && rec.inv_slo != 0.0f && !inf_inv;  // false for NaN, +-Inf
  rec.orig = p | (f6 ? kPodF6 : 0u) | (inf_inv ? kPodInfInv : (plain ? 0u : kPodGeneral));
  *kind = inf_inv ? kKindInfInv : (!plain ? kKindGeneral : (rec.conf_v100 == -1.0f ? kKindNoV100 : kKindPlain));
  return rec;
}

// bucket of every pod that is scored itself (class representatives, or all pods without dedupe)
__global__ void pod_hist_kernel(DevTables T, const uint32_t* __restrict__ pod_app, const float* __restrict__ pod_slo,
                                int n_pods, const uint32_t* __restrict__ rep, uint32_t* counts,
                                uint32_t* __restrict__ bucket) {
  int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_pods) return;
  if (rep && rep[p] != (uint32_t)p) {
    bucket[p] = kNoBucket;
    return;
  }
  int kind;
  const uint32_t a = pod_app[p];
  (void)make_pod_rec(T, a, pod_slo[p], (uint32_t)p, &kind);
  const int b = bucket_of(a, T.n_apps, kind);
  bucket[p] = (uint32_t)b;
  atomicAdd(&counts[b], 1u);
}

// which apps occur in a batch at all (bucket kind 0 of each app): stage A's input when a batch is scored in slices
__global__ void app_presence_kernel(const uint32_t* __restrict__ pod_app, int n_pods, int n_apps, uint32_t* counts) {
  int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_pods) return;
  atomicAdd(&counts[bucket_of(pod_app[p], n_apps, 0)], 1u);
}

// Single CTA: exclusive scan of the bucket sizes -> scatter cursors and chunk descriptors.
__global__ void __launch_bounds__(1024) pod_scan_kernel(int n_apps, const uint32_t* __restrict__ counts,
                                                        uint32_t* __restrict__ cursor,
                                                        ChunkDesc* __restrict__ chunks, uint32_t* n_chunks,
                                                        uint32_t max_chunks) {
  __shared__ uint32_t s_warp_c[32], s_warp_n[32];
  __shared__ uint32_t s_carry_c, s_carry_n;
  if (threadIdx.x == 0) {
    s_carry_c = 0;
    s_carry_n = 0;
  }
  __syncthreads();
  const int n_buckets = (n_apps + 1) * kKinds;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int base = 0; base < n_buckets; base += 1024) {
    int b = base + threadIdx.x;
    uint32_t c = b < n_buckets ? counts[b] : 0u;
    // work items: the four kind buckets of an app are neighbours in the sorted order and share work
    // items (the app's first bucket emits them)
    const int kind_b = b % kKinds;
    uint32_t ck[kKinds] = {0u, 0u, 0u, 0u};
    uint32_t c_items = 0u;
    if (kind_b == 0 && b < n_buckets)
      for (int q = 0; q < kKinds; ++q) {
        ck[q] = counts[b + q];
        c_items += ck[q];
      }
    uint32_t n = (c_items + kChunk - 1) / kChunk;
    uint32_t ic = c, in = n;  // inclusive warp scans
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      uint32_t tc = __shfl_up_sync(0xFFFFFFFFu, ic, d);
      uint32_t tn = __shfl_up_sync(0xFFFFFFFFu, in, d);
      if (lane >= d) {
        ic += tc;
        in += tn;
      }
    }
    if (lane == 31) {
      s_warp_c[warp] = ic;
      s_warp_n[warp] = in;
    }
    __syncthreads();
    if (warp == 0) {
      uint32_t wc = s_warp_c[lane], wn = s_warp_n[lane];
      uint32_t xc = wc, xn = wn;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        uint32_t tc = __shfl_up_sync(0xFFFFFFFFu, xc, d);
        uint32_t tn = __shfl_up_sync(0xFFFFFFFFu, xn, d);
        if (lane >= d) {
          xc += tc;
          xn += tn;
        }
      }
      s_warp_c[lane] = xc - wc;  // exclusive warp offsets
      s_warp_n[lane] = xn - wn;
    }
    __syncthreads();
    uint32_t off_c = s_carry_c + s_warp_c[warp] + ic - c;
    uint32_t off_n = s_carry_n + s_warp_n[warp] + in - n;
    if (b < n_buckets) {
      cursor[b] = off_c;
      const int ai = b / kKinds;
      uint32_t app = ai < n_apps ? (uint32_t)ai : kAppUnknown;
      for (uint32_t j = 0; j < n; ++j) {
        if (off_n + j < max_chunks) {
          uint32_t len = c_items - j * kChunk;
          len = len < (uint32_t)kChunk ? len : (uint32_t)kChunk;
          // pods of each kind inside [j * kChunk, j * kChunk + len) of the app's range
          const uint32_t lo = j * kChunk, hi = lo + len;
          uint32_t inside[kKinds], pos = 0u;
          for (int q = 0; q < kKinds; ++q) {
            const uint32_t a = pos > lo ? pos : lo, e = pos + ck[q] < hi ? pos + ck[q] : hi;
            inside[q] = e > a ? e - a : 0u;
            pos += ck[q];
          }
          chunks[off_n + j] = ChunkDesc{app, off_c + lo, len, pack_kinds(inside[0], inside[1], inside[2])};
        }
      }
    }
    __syncthreads();
    if (threadIdx.x == 1023) {
      s_carry_c = off_c + c;
      s_carry_n = off_n + n;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    n_chunks[0] = s_carry_n < max_chunks ? s_carry_n : max_chunks;
    n_chunks[1] = 0u;  // work counter of the persistent scoring kernel
  }
}

__global__ void pod_scatter_kernel(DevTables T, const uint32_t* __restrict__ pod_app,
                                   const float* __restrict__ pod_slo, int n_pods,
                                   const uint32_t* __restrict__ bucket, uint32_t* cursor,
                                   PodRec* __restrict__ sorted) {
  int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_pods) return;
  const uint32_t b = bucket[p];
  if (b == kNoBucket) return;
  int kind;
  const PodRec rec = make_pod_rec(T, pod_app[p], pod_slo[p], (uint32_t)p, &kind);
  sorted[atomicAdd(&cursor[b], 1u)] = rec;
}

// apps present in a batch (counts must hold n_buckets + 1 words; zeroed here)
cudaError_t launch_pod_histogram(const uint32_t* pod_app, int n_pods, int n_apps, uint32_t* counts, cudaStream_t s) {
  cudaError_t e = cudaMemsetAsync(counts, 0, sizeof(uint32_t) * (size_t)(n_buckets_for(n_apps) + 1), s);
  if (e != cudaSuccess) return e;
  note_launch();
  app_presence_kernel<<<(n_pods + 255) / 256, 256, 0, s>>>(pod_app, n_pods, n_apps, counts);
  return cudaGetLastError();
}

// Class dedupe (if the scratch carries a table), then bucket the scored pods by (app, kind):
// histogram -> scan (cursors + work items) -> scatter.
cudaError_t launch_pod_bucketing(const DevTables& T, const uint32_t* pod_app, const float* pod_slo, int n_pods,
                                 const PodScratch& sc, cudaStream_t s) {
  cudaError_t e = cudaMemsetAsync(sc.counts, 0, sizeof(uint32_t) * (size_t)(n_buckets_for(T.n_apps) + 1), s);
  if (e != cudaSuccess) return e;
  const int pb = (n_pods + 255) / 256;
  if (sc.rep) {
    if ((e = cudaMemsetAsync(sc.hkeys, 0xFF, sizeof(unsigned long long) * (size_t)sc.hash_size, s)) != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(sc.hvals, 0xFF, sizeof(uint32_t) * (size_t)sc.hash_size, s)) != cudaSuccess) return e;
    note_launch();
    class_insert_kernel<<<pb, 256, 0, s>>>(pod_app, pod_slo, n_pods, sc.hkeys, sc.hvals, sc.hash_size - 1, sc.rep);
    note_launch();
    class_lookup_kernel<<<pb, 256, 0, s>>>(n_pods, sc.hvals, sc.rep);
  }
  note_launch();
  pod_hist_kernel<<<pb, 256, 0, s>>>(T, pod_app, pod_slo, n_pods, sc.rep, sc.counts, sc.bucket);
  note_launch();
  pod_scan_kernel<<<1, 1024, 0, s>>>(T.n_apps, sc.counts, sc.cursor, sc.chunks, sc.n_chunks, sc.max_chunks);
  note_launch();
  pod_scatter_kernel<<<pb, 256, 0, s>>>(T, pod_app, pod_slo, n_pods, sc.bucket, sc.cursor, sc.sorted);
  return cudaGetLastError();
}

// ------------------------------------------------------------------ normalise --------

// NormalizeScore (gpu_plugins.go:816-841) over raw scores in place: one warp per pod.
// Raw scores lie in [0,100], so the integer rescale (s-lo)*100/(hi-lo) is a 101-entry
// table per pod, built with the same truncating integer division.
__global__ void __launch_bounds__(128) normalize_kernel(int n_pods, int n_nodes, int rescale,
                                                       uint8_t* __restrict__ score,
                                                       const unsigned long long* __restrict__ hi_arg,
                                                       const unsigned long long* __restrict__ neg_lo,
                                                       int32_t* __restrict__ best) {
  __shared__ uint8_t s_lut[4][128];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int p = blockIdx.x * 4 + warp;
  if (p >= n_pods) return;
  const unsigned long long ha = hi_arg[p];
  const int hi = (int)(ha >> 32);
  const int lo = 100 - (int)neg_lo[p];
  if (lane == 0 && best) best[p] = (int32_t)(0xFFFFFFFFu - (uint32_t)(ha & 0xFFFFFFFFull));
  if (!rescale) return;
  const int range = hi - lo;
  uint8_t* lut = s_lut[warp];
  for (int s = lane; s < 128; s += 32) {
    int v = 0;
    if (range != 0 && s >= lo && s <= hi) v = ((s - lo) * 100) / range;
    lut[s] = (uint8_t)v;
  }
  __syncwarp();
  uint8_t* row = score + (size_t)p * n_nodes;
  // scalar head up to 16-byte alignment, vector body, scalar tail
  uintptr_t addr = (uintptr_t)row;
  int head = (int)((16 - (addr & 15)) & 15);
  if (head > n_nodes) head = n_nodes;
  for (int i = lane; i < head; i += 32) row[i] = lut[row[i] & 127];
  int nvec = (n_nodes - head) / 16;
  uint4* v4 = (uint4*)(row + head);
  for (int i = lane; i < nvec; i += 32) {
    uint4 v = v4[i];
    uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      uint32_t x = w[j];
      w[j] = (uint32_t)lut[x & 127] | ((uint32_t)lut[(x >> 8) & 127] << 8) | ((uint32_t)lut[(x >> 16) & 127] << 16) |
             ((uint32_t)lut[(x >> 24) & 127] << 24);
    }
    v4[i] = make_uint4(w[0], w[1], w[2], w[3]);
  }
  for (int i = head + nvec * 16 + lane; i < n_nodes; i += 32) row[i] = lut[row[i] & 127];
}

cudaError_t launch_normalize(int n_pods, int n_nodes, int rescale, uint8_t* score,
                             const unsigned long long* hi_arg, const unsigned long long* neg_lo, int32_t* best,
                             cudaStream_t s) {
  if (n_pods == 0) return cudaSuccess;
  note_launch();
  normalize_kernel<<<(n_pods + 3) / 4, 128, 0, s>>>(n_pods, n_nodes, rescale, score, hi_arg, neg_lo, best);
  return cudaGetLastError();
}

// NormalizeScore on one NodeScoreList of int64 scores, literally gpu_plugins.go:816-841.
__global__ void __launch_bounds__(1024) normalize_i64_kernel(long long* __restrict__ scores, int n) {
  __shared__ long long s_hi[32], s_lo[32];
  long long hi = -0x7FFFFFFFFFFFFFFFLL, lo = 0x7FFFFFFFFFFFFFFFLL;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    long long v = scores[i];
    hi = v > hi ? v : hi;
    lo = v < lo ? v : lo;
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    long long oh = __shfl_xor_sync(0xFFFFFFFFu, hi, d), ol = __shfl_xor_sync(0xFFFFFFFFu, lo, d);
    hi = oh > hi ? oh : hi;
    lo = ol < lo ? ol : lo;
  }
  if ((threadIdx.x & 31) == 0) {
    s_hi[threadIdx.x >> 5] = hi;
    s_lo[threadIdx.x >> 5] = lo;
  }
  __syncthreads();
  hi = s_hi[0];
  lo = s_lo[0];
  for (int w = 1; w < (int)(blockDim.x >> 5); ++w) {
    hi = s_hi[w] > hi ? s_hi[w] : hi;
    lo = s_lo[w] < lo ? s_lo[w] : lo;
  }
  const long long old_range = hi - lo, new_range = 100;
  for (int i = threadIdx.x; i < n; i += blockDim.x)
    scores[i] = old_range == 0 ? 0 : ((scores[i] - lo) * new_range / old_range);
}

cudaError_t launch_normalize_i64(long long* scores, int n, cudaStream_t s) {
  if (n == 0) return cudaSuccess;
  note_launch();
  normalize_i64_kernel<<<1, 1024, 0, s>>>(scores, n);
  return cudaGetLastError();
}

// ------------------------------------------------------------------ decisions --------

// What the host needs of a scheduling decision, per pod: the chosen node, the selection byte there
// (Logic's {nodeName: selectedUUID} record, gpu_plugins.go:759-768, which PostBind reads back at
// :894-895), the raw score extrema, and which V100 configuration Logic recorded as MPS_<node>
// (:640-665; PostBind turns it into the MPS limits at :896-903).
__global__ void decisions_kernel(DevTables T, const uint32_t* __restrict__ pod_app, const float* __restrict__ pod_slo,
                                 int n_pods, const uint8_t* __restrict__ sel,
                                 const unsigned long long* __restrict__ hi_arg,
                                 const unsigned long long* __restrict__ neg_lo, Decision* __restrict__ out) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_pods) return;
  Decision d;
  d.node = -1;
  d.sel = (uint8_t)kSelNone;
  d.raw_hi = d.raw_lo = d.mps = 0;
  if (T.n_nodes > 0) {
    const unsigned long long ha = hi_arg[p];
    const uint32_t node = 0xFFFFFFFFu - (uint32_t)(ha & 0xFFFFFFFFull);  // global index
    const int local = (int)node - T.node_base;
    d.node = (int32_t)node;
    d.raw_hi = (uint8_t)(ha >> 32);
    d.raw_lo = (uint8_t)(100 - (int)neg_lo[p]);
    if (local >= 0 && local < T.n_nodes) {  // on a node shard the winner may live on another rank
      d.sel = sel[(size_t)p * T.n_nodes + local];
      if (T.redis_reachable && T.node_model[local] == kModelV100 && T.node_ngpu[local] > 0) {
        // confIndex starts as "1P_V100" and moves to the smallest prediction above the SLO
        const uint32_t app = pod_app[p];
        const float slo = pod_slo[p];
        float c = -1.0f;
        int which = 1;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          float tmp;
          if (conf_lookup(T, app, kModelV100, 1 << i, &tmp) && tmp > slo && (c == -1.0f || tmp < c)) {
            c = tmp;
            which = 1 << i;
          }
        }
        d.mps = (uint8_t)which;
      }
    }
  }
  out[p] = d;
}

cudaError_t launch_decisions(const DevTables& T, const uint32_t* pod_app, const float* pod_slo, int n_pods,
                             const uint8_t* sel, const unsigned long long* hi_arg, const unsigned long long* neg_lo,
                             Decision* out, cudaStream_t s) {
  if (n_pods == 0) return cudaSuccess;
  note_launch();
  decisions_kernel<<<(n_pods + 255) / 256, 256, 0, s>>>(T, pod_app, pod_slo, n_pods, sel, hi_arg, neg_lo, out);
  return cudaGetLastError();
}

// ------------------------------------------------------------------ MIG choice -------

// reconfigure()'s arithmetic, gpu_plugins.go:366-399; one thread per pod.
__global__ void select_mig_kernel(DevTables T, const uint32_t* __restrict__ pod_app,
                                  const float* __restrict__ pod_slo, int n_pods, uint8_t* __restrict__ out) {
  int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_pods) return;
  uint32_t app = pod_app[p];
  float slo = pod_slo[p];
  int opt = 0;
  double max_score = 0.0;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    float conf;
    if (!conf_lookup(T, app, kModelA30, 1 << i, &conf)) continue;
    double sc;
    if (slo > conf) {
      double td = (double)__fadd_rn(__fsub_rn(slo, conf), 1.0f);
      sc = __ddiv_rn(100.0, __dadd_rn(1.0, __dmul_rn(td, td)));
    } else {
      sc = __ddiv_rn(100.0, __dadd_rn(1.0, (double)__fsub_rn(slo, conf)));
    }
    if (sc >= max_score) {
      max_score = sc;
      opt = (1 << i) - 1;
    }
  }
  if (opt == 3) opt = 2;
  out[p] = (uint8_t)opt;
}

cudaError_t launch_select_mig(const DevTables& T, const uint32_t* pod_app, const float* pod_slo, int n_pods,
                              uint8_t* out, cudaStream_t s) {
  if (n_pods == 0) return cudaSuccess;
  note_launch();
  select_mig_kernel<<<(n_pods + 255) / 256, 256, 0, s>>>(T, pod_app, pod_slo, n_pods, out);
  return cudaGetLastError();
}

}  // namespace gs

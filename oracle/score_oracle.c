/*
 * score_oracle.c — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement, in plain C, of the scoring hot path of dimgatz98/k8s-gpu-scheduler:
 * Logic() / NormalizeScore() / the arithmetic of reconfigure() in
 * pkg/plugins/gpu_plugin/gpu_plugins.go.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this file's library; the
 * product library (libgpusched.so) never does.
 *
 * PARITY UNPINNED for the Go side: the reference has no test, golden vector or
 * known-answer fixture for this path (no *_test.go next to gpu_plugins.go), and neither
 * a Go toolchain nor the k8s.io/kubernetes v1.21.0 module (go.mod:67) exists on this
 * machine, so the reference itself could not be run.  The restatement below follows the
 * Go source statement by statement (line numbers cited at each step) and is pinned only
 * by hand-derived known answers in tests/test_oracle_kat.py.
 *
 * Determinisation (the reference is non-deterministic):
 *  (i)   no shuffle: GPUs are visited in Redis array order (replaces the time-seeded
 *        rand.Shuffle at gpu_plugins.go:560-561; only ties between GPUs are affected);
 *  (ii)  residents are visited in snapshot (CSR) order (replaces Go map order at
 *        :571,:595,:703), which fixes the float32/float64 summation order;
 *  (iii) pod-name -> table-row/column substring matching (:600-605,:607-612,:708-713,
 *        recom_server.py:67-71) is resolved by the host to exact app ids; an app that
 *        matches nothing is ORC_APP_UNKNOWN;
 *  (iv)  best node = highest normalised score, lowest node index on ties (the framework
 *        uses reservoir sampling).
 * Floating point is pinned to Go on amd64: every float32 operation rounds to float32,
 * float64 to float64, no fused multiply-add (build with -ffp-contract=off, never
 * -ffast-math); math.Pow(x, 2) rounds once, like x*x; int64(NaN) compares below any
 * score, as CVTTSD2SQ's 0x8000000000000000 does.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_MODEL_NONE 0
#define ORC_MODEL_A30 1
#define ORC_MODEL_V100 2
#define ORC_APP_UNKNOWN 0xFFFFFFFFu
#define ORC_MAX_GPUS 31
#define ORC_SEL_NONE 31

typedef struct {
  int32_t n_nodes;
  int32_t max_gpus;
  const uint8_t* node_model;
  const uint8_t* node_ngpu;
  const uint32_t* gpu_off;
  const uint32_t* res_app;
  const float* res_slo;
  int32_t redis_reachable;
} orc_cluster;

typedef struct {
  int32_t n_apps;
  int32_t n_cols;
  const float* conf;     /* [n_apps x n_cols] */
  const float* interf;   /* [2 x n_apps x n_apps] */
  const int32_t* col_of; /* [2 x (ORC_MAX_GPUS+1)] */
} orc_preds;

/* confPredictions[fmt.Sprintf("%dP_%s", k, model)] — returns ok like a Go map lookup.
 * gpu_plugins.go:583-584, 636-637; empty map for an unknown app :322-325. */
static int conf_lookup(const orc_preds* pr, uint32_t app, int model, int k, float* out) {
  if (app == ORC_APP_UNKNOWN || app >= (uint32_t)pr->n_apps) return 0;
  if (k < 0 || k > ORC_MAX_GPUS) return 0;
  int32_t col = pr->col_of[(model - 1) * (ORC_MAX_GPUS + 1) + k];
  if (col < 0) return 0;
  *out = pr->conf[(size_t)app * pr->n_cols + col];
  return 1;
}

/* first key of tmpInterference contained in the co-located pod's name, resolved to an
 * app id by rule (iii).  gpu_plugins.go:600-605, 607-612, 708-713. */
static int interf_lookup(const orc_preds* pr, int model, uint32_t row_app, uint32_t col_app,
                         float* out) {
  if (row_app == ORC_APP_UNKNOWN || row_app >= (uint32_t)pr->n_apps) return 0;
  if (col_app == ORC_APP_UNKNOWN || col_app >= (uint32_t)pr->n_apps) return 0;
  size_t A = (size_t)pr->n_apps;
  *out = pr->interf[(size_t)(model - 1) * A * A + (size_t)row_app * A + col_app];
  return 1;
}

/* the two-branch SLO term, gpu_plugins.go:616-622 and :717-723 */
static void slo_term(float SLO, float confPrediction, float interference, double* negative_sum,
                     double* positive_sum, int* count_negative, int* count_positive) {
  float d = confPrediction - interference;
  float t = (1 / SLO) * (SLO - d);  /* float32 throughout */
  double x = fabs((double)t);
  if (SLO > d) {
    *count_negative += 1;
    double y = x + 1;
    double y2 = y * y;  /* math.Pow(y, 2) */
    *negative_sum += 1 / (1 + y2);
  } else {
    *count_positive += 1;
    *positive_sum += 1 / (1 + x);
  }
}

/* gpu_plugins.go:675-688 and :731-743 */
static double combine(double negative_sum, double positive_sum, int count_negative,
                      int count_positive) {
  double factor = 100.0;
  double tmpScore = 0;
  if (count_positive > 0 && count_negative > 0) {
    double k = (double)count_negative / (double)(count_negative + count_positive);
    double a = (1 - k) * positive_sum;
    a = a / (double)count_positive;
    a = factor * a;
    double b = k * negative_sum;
    b = b / (double)count_negative;
    b = factor * b;
    tmpScore = a + b;
  } else if (count_negative > 0) {
    tmpScore = factor * (negative_sum / (double)count_negative);
  } else if (count_positive > 0) {
    tmpScore = factor * (positive_sum / (double)count_positive);
  }
  return tmpScore;
}

/* Go int64(float64) on amd64; NaN/out of range -> INT64_MIN */
static int64_t go_int64(double v) {
  if (!(v == v) || v >= 9223372036854775808.0 || v < -9223372036854775808.0) return INT64_MIN;
  return (int64_t)v;
}

#define ORC_F6 (1u << 5)
#define ORC_F7 (1u << 6)
#define ORC_ELIGIBLE (1u << 7)

/* Logic(), gpu_plugins.go:455-777, for one (pod, node).  Returns the raw score and
 * writes the packed selection byte (same layout as GS_SEL_* in include/gpusched.h). */
int64_t orc_logic(const orc_cluster* cl, const orc_preds* pr, uint32_t pod_app, float currSLO,
                  int32_t pod_self, int32_t node, uint8_t* out_sel) {
  int64_t score = 0; /* framework.MinNodeScore, :457 */
  int sel = ORC_SEL_NONE;
  unsigned flags = 0;
  if (!cl->redis_reachable) { /* :508-527: returns 0 whatever DCGM says */
    *out_sel = (uint8_t)ORC_SEL_NONE;
    return 0;
  }
  int model = cl->node_model[node]; /* :477-499 */
  int n_uuids = cl->node_ngpu[node]; /* :536-542 */
  if (model != ORC_MODEL_NONE && n_uuids > 0) flags |= ORC_ELIGIBLE;
  if (model != ORC_MODEL_NONE) { /* :558 */
    int G = cl->max_gpus;
    for (int g = 0; g < n_uuids; ++g) { /* :562, rule (i) */
      double negative_sum = 0, positive_sum = 0;
      int count_negative = 0, count_positive = 0;
      uint32_t lo = cl->gpu_off[(size_t)node * G + g];
      uint32_t hi = cl->gpu_off[(size_t)node * G + g + 1];
      for (uint32_t r = lo; r < hi; ++r) { /* :571, rule (ii) */
        float SLO = cl->res_slo[r];
        if (SLO == 0) continue; /* :572 */
        float confPrediction;
        if (!conf_lookup(pr, cl->res_app[r], model, n_uuids, &confPrediction)) continue; /* :583-587 */
        float interference = 0; /* :589 */
        for (uint32_t c = lo; c < hi; ++c) { /* :595 */
          if (c == r || (pod_self >= 0 && c == (uint32_t)pod_self)) continue; /* :596 */
          float val;
          if (interf_lookup(pr, model, cl->res_app[r], cl->res_app[c], &val)) interference += val;
        }
        {
          float val; /* :607-612: the pod being scheduled, added last */
          if (interf_lookup(pr, model, cl->res_app[r], pod_app, &val)) interference += val;
        }
        slo_term(SLO, confPrediction, interference, &negative_sum, &positive_sum, &count_negative,
                 &count_positive); /* :616-622 */
      }
      /* the pod being scheduled, :626-666 */
      float confPrediction = -1;
      if (model == ORC_MODEL_A30) {
        float v = 0; /* missing key -> zero value, NOT -1 (:637) */
        int ok = conf_lookup(pr, pod_app, model, n_uuids, &v);
        confPrediction = ok ? v : 0.0f;
        if (ok) flags |= ORC_F6;
      } else { /* V100, :638-651 */
        static const int parts[3] = {1, 2, 4};
        for (int i = 0; i < 3; ++i) {
          float tmp;
          int ok = conf_lookup(pr, pod_app, model, parts[i], &tmp);
          if (ok) flags |= ORC_F6;
          if (ok && tmp > currSLO && (confPrediction == -1 || tmp < confPrediction)) confPrediction = tmp;
        }
      }
      if (confPrediction == -1) { /* :667-696 */
        double tmpScore = combine(negative_sum, positive_sum, count_negative, count_positive);
        if (go_int64(tmpScore) > score) {
          score = go_int64(tmpScore);
          sel = g;
        }
        continue;
      }
      flags |= ORC_F7;
      float interference = 0; /* :698 */
      for (uint32_t c = lo; c < hi; ++c) { /* :703 */
        if (pod_self >= 0 && c == (uint32_t)pod_self) continue; /* :704 */
        float val;
        if (interf_lookup(pr, model, pod_app, cl->res_app[c], &val)) interference += val;
      }
      slo_term(currSLO, confPrediction, interference, &negative_sum, &positive_sum, &count_negative,
               &count_positive); /* :717-723 */
      double tmpScore = combine(negative_sum, positive_sum, count_negative, count_positive); /* :731-743 */
      if (go_int64(tmpScore) > score) { /* :752-755 */
        score = go_int64(tmpScore);
        sel = g;
      }
    }
  }
  *out_sel = (uint8_t)(sel | flags);
  return score; /* :776 */
}

/* NormalizeScore(), gpu_plugins.go:816-841, on one pod's NodeScoreList */
void orc_normalize(int64_t* scores, int32_t n) {
  int64_t highest = -INT64_MAX; /* -math.MaxInt64 */
  int64_t lowest = INT64_MAX;
  for (int32_t i = 0; i < n; ++i) {
    if (scores[i] > highest) highest = scores[i];
    if (scores[i] < lowest) lowest = scores[i];
  }
  int64_t oldRange = highest - lowest;
  int64_t newRange = 100 - 0; /* framework.MaxNodeScore - framework.MinNodeScore */
  for (int32_t i = 0; i < n; ++i) {
    if (oldRange == 0) {
      scores[i] = 0;
    } else {
      scores[i] = ((scores[i] - lowest) * newRange / oldRange) + 0;
    }
  }
}

/* Whole batch: Score for every (pod,node), optional NormalizeScore per pod, best node by
 * rule (iv).  OpenMP over pods (the framework itself fans Score out over 16 workers).
 * out_score/out_sel are [n_pods x n_nodes] bytes; out_best may be NULL. */
void orc_score_batch(const orc_cluster* cl, const orc_preds* pr, const uint32_t* pod_app,
                     const float* pod_slo, const int32_t* pod_self, int32_t n_pods, int normalize,
                     int32_t node_base, uint8_t* out_score, uint8_t* out_sel, int32_t* out_best,
                     int n_threads) {
  int32_t N = cl->n_nodes;
#ifdef _OPENMP
  if (n_threads > 0) omp_set_num_threads(n_threads);
#else
  (void)n_threads;
#endif
#pragma omp parallel
  {
    int64_t* row = N > 0 ? (int64_t*)__builtin_malloc(sizeof(int64_t) * (size_t)N) : 0;
#pragma omp for schedule(dynamic, 8)
    for (int32_t p = 0; p < n_pods; ++p) {
      int32_t self = pod_self ? pod_self[p] : -1;
      for (int32_t n = 0; n < N; ++n) {
        row[n] = orc_logic(cl, pr, pod_app[p], pod_slo[p], self, n, &out_sel[(size_t)p * N + n]);
      }
      int32_t best = -1;
      int64_t best_s = -1;
      for (int32_t n = 0; n < N; ++n) {
        if (row[n] > best_s) {
          best_s = row[n];
          best = n;
        }
      }
      if (normalize) orc_normalize(row, N);
      for (int32_t n = 0; n < N; ++n) out_score[(size_t)p * N + n] = (uint8_t)row[n];
      if (out_best) out_best[p] = best < 0 ? -1 : best + node_base;
    }
    __builtin_free(row);
  }
}

/* reconfigure()'s choice of MIG configuration, gpu_plugins.go:366-399.  Returns the
 * index into configs = {"all-4g.24gb","all-2g.12gb","all-1g.6gb"} (:52). */
int orc_select_mig(const orc_preds* pr, uint32_t pod_app, float SLO) {
  static const int parts[3] = {1, 2, 4};
  int optimalConfiguration = 0;
  double maxScore = 0;
  for (int i = 0; i < 3; ++i) {
    float confPrediction;
    if (!conf_lookup(pr, pod_app, ORC_MODEL_A30, parts[i], &confPrediction)) continue; /* :371-374 */
    double score;
    if (SLO > confPrediction) {
      float t = SLO - confPrediction + 1; /* float32, :376 */
      double td = (double)t;
      score = 100 / (1 + td * td);
    } else {
      float t = SLO - confPrediction; /* :386 — no abs: may be negative or +Inf */
      score = 100 / (1 + (double)t);
    }
    if (score >= maxScore) { /* :377,:387 */
      maxScore = score;
      optimalConfiguration = parts[i] - 1;
    }
  }
  if (optimalConfiguration == 3) optimalConfiguration -= 1; /* :397-399 */
  return optimalConfiguration;
}

void orc_select_mig_batch(const orc_preds* pr, const uint32_t* pod_app, const float* pod_slo,
                          int32_t n_pods, uint8_t* out) {
  for (int32_t p = 0; p < n_pods; ++p) out[p] = (uint8_t)orc_select_mig(pr, pod_app[p], pod_slo[p]);
}

int orc_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

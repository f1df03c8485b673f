/*
 * score_one_pod.c — the C ABI of libgpusched.so from plain C: two nodes with two GPUs each, one
 * pending pod, raw and normalised scores and the selected GPU per node.  This is the call
 * sequence the cgo binding of INTEGRATION.md makes (upload tables once, score per cycle).
 *
 *   gcc -std=c99 -Iinclude examples/score_one_pod.c -Lk8s-gpu-scheduler_b200 -lgpusched \
 *       -Wl,-rpath,$PWD/k8s-gpu-scheduler_b200 -o /tmp/score_one_pod && /tmp/score_one_pod
 *
 * Needs a CUDA device: there is no CPU fallback (gs_create fails without one).
 */
#include <stdio.h>
#include <string.h>

#include "gpusched.h"

#define CHECK(call)                                                              \
  do {                                                                           \
    int rc_ = (call);                                                            \
    if (rc_ != GS_OK) {                                                          \
      fprintf(stderr, "%s -> %d: %s\n", #call, rc_, gs_last_error(ctx));         \
      gs_destroy(ctx);                                                           \
      return 1;                                                                  \
    }                                                                            \
  } while (0)

int main(void) {
  gs_ctx* ctx = NULL;
  if (gs_create(0, &ctx) != GS_OK) {
    fprintf(stderr, "gs_create failed (no usable CUDA device?)\n");
    return 2;
  }
  printf("%s\n", gs_version());

  /* two apps, columns 1P_A30, 2P_A30, 1P_V100, 2P_V100 */
  enum { A = 2, K = 4, G = 2, N = 2 };
  const float conf[A * K] = {100.f, 60.f, 140.f, 90.f, /* app 0 */
                             200.f, 120.f, 260.f, 150.f /* app 1 */};
  float interf[2 * A * A];
  int32_t col_of[2 * (GS_MAX_GPUS_PER_NODE + 1)];
  int i;
  for (i = 0; i < 2 * A * A; ++i) interf[i] = 5.f + (float)i;
  for (i = 0; i < 2 * (GS_MAX_GPUS_PER_NODE + 1); ++i) col_of[i] = -1;
  col_of[0 * (GS_MAX_GPUS_PER_NODE + 1) + 1] = 0; /* 1P_A30 */
  col_of[0 * (GS_MAX_GPUS_PER_NODE + 1) + 2] = 1; /* 2P_A30 */
  col_of[1 * (GS_MAX_GPUS_PER_NODE + 1) + 1] = 2; /* 1P_V100 */
  col_of[1 * (GS_MAX_GPUS_PER_NODE + 1) + 2] = 3; /* 2P_V100 */
  gs_predictions preds;
  memset(&preds, 0, sizeof preds);
  preds.n_apps = A;
  preds.n_cols = K;
  preds.conf = conf;
  preds.interf = interf;
  preds.col_of = col_of;
  CHECK(gs_upload_predictions(ctx, &preds));

  /* node 0: A30 with 2 UUIDs, one resident of app 1 (SLO 300) on its first GPU;
     node 1: V100 with 2 UUIDs, residents of app 0 (SLO 150) and app 1 (SLO 250) on its second */
  const uint8_t node_model[N] = {GS_MODEL_A30, GS_MODEL_V100};
  const uint8_t node_ngpu[N] = {2, 2};
  const uint32_t gpu_off[N * G + 1] = {0, 1, 1, 1, 3};
  const uint32_t res_app[3] = {1, 0, 1};
  const float res_slo[3] = {300.f, 150.f, 250.f};
  gs_cluster cl;
  memset(&cl, 0, sizeof cl);
  cl.n_nodes = N;
  cl.max_gpus = G;
  cl.node_model = node_model;
  cl.node_ngpu = node_ngpu;
  cl.gpu_off = gpu_off;
  cl.res_app = res_app;
  cl.res_slo = res_slo;
  cl.redis_reachable = 1;
  CHECK(gs_upload_cluster(ctx, &cl, 0));

  /* one pending pod of app 0 with SLO 180 */
  const uint32_t pod_app[1] = {0};
  const float pod_slo[1] = {180.f};
  uint8_t raw[N], norm[N], sel[N];
  int32_t best[1];
  CHECK(gs_score_batch(ctx, pod_app, pod_slo, NULL, 1, GS_SCORE_RAW, raw, sel, NULL));
  CHECK(gs_score_batch(ctx, pod_app, pod_slo, NULL, 1, GS_SCORE_NORMALIZE, norm, sel, best));
  for (i = 0; i < N; ++i) {
    printf("node %d: Score %3u  NormalizeScore %3u  ", i, raw[i], norm[i]);
    if (GS_SEL_GPU(sel[i]) == GS_SEL_NONE)
      printf("no GPU selected");
    else
      printf("GPU #%u", GS_SEL_GPU(sel[i]));
    printf("  (eligible %u, F6 %u, F7 %u)\n", GS_SEL_ELIGIBLE(sel[i]), GS_SEL_F6(sel[i]), GS_SEL_F7(sel[i]));
  }
  printf("best node: %d\n", (int)best[0]);
  gs_destroy(ctx);
  return 0;
}

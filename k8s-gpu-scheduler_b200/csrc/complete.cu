// complete.cu — recommender completion (IterativeImputer + BayesianRidge) entry points.
// Placeholder bodies until the CUDA path lands: they fail loudly, never fall back.
#include "../../include/gpusched.h"

struct gs_model {
  int d;
};

extern "C" {
int gs_complete_fit(gs_ctx*, const double*, int32_t, int32_t, gs_model** out) {
  if (out) *out = nullptr;
  return GS_ERR_UNSUPPORTED;
}
int gs_complete_transform(gs_ctx*, const gs_model*, const double*, int32_t, double*) { return GS_ERR_UNSUPPORTED; }
int gs_model_info(const gs_model*, int32_t*, int32_t*, int32_t*) { return GS_ERR_UNSUPPORTED; }
int gs_model_step(const gs_model*, int32_t, int32_t*, double*, double*) { return GS_ERR_UNSUPPORTED; }
void gs_model_destroy(gs_model*) {}
}

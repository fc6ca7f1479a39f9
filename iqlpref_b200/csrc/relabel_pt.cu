// Preference-transformer relabel (placeholder; kernels land in a later commit).
#include "iql_step.cuh"
using namespace iqlb;
extern "C" int iqlb_pt_param_count(const iqlb_pt_desc* d, int64_t* n_floats) {
  (void)d; (void)n_floats;
  return fail(IQLB_ERR_UNSUPPORTED, "PT relabel not built yet%s", "");
}
extern "C" int iqlb_relabel_pt(const iqlb_pt_desc*, const float*, const float*, const float*, int64_t, const int32_t*,
                               const int32_t*, int64_t, float*, void*) {
  return fail(IQLB_ERR_UNSUPPORTED, "PT relabel not built yet%s", "");
}
extern "C" int iqlb_pt_forward(const iqlb_pt_desc*, const float*, const float*, const float*, const int32_t*,
                               const float*, int64_t, float*, void*) {
  return fail(IQLB_ERR_UNSUPPORTED, "PT forward not built yet%s", "");
}

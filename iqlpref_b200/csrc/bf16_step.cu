// bf16 tcgen05/TMEM path of the IQL step (precision == 1).  Placeholder until the UMMA kernels land.
#include "iql_step.cuh"

namespace iqlb {
int bf16_create(iqlb_agents*) { return fail(IQLB_ERR_UNSUPPORTED, "bf16 tcgen05 path not built yet%s", ""); }
void bf16_destroy(iqlb_agents*) {}
int bf16_step_gemms_fwd(iqlb_agents*, cudaStream_t) { return IQLB_ERR_UNSUPPORTED; }
int bf16_step_gemms_bwd(iqlb_agents*, cudaStream_t) { return IQLB_ERR_UNSUPPORTED; }
}  // namespace iqlb

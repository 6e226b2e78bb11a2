// placeholder until the tcgen05 kernel lands
#include "engine.h"
namespace gp {
int gemm_probe_locked(Ctx*, Device*, uint32_t, uint64_t, uint32_t, uint32_t, uint32_t, gp_gemm_result*) { set_error("gemm probe not built"); return GP_ERR_UNSUPPORTED; }
int gemm_host_locked(Ctx*, Device*, const uint16_t*, const uint16_t*, float*, uint32_t, uint32_t, uint32_t) { set_error("gemm probe not built"); return GP_ERR_UNSUPPORTED; }
}

// placeholder until the peer kernels land
#include "engine.h"
namespace gp {
int p2p_probe(Ctx*, int, size_t, int, double*) { set_error("p2p probe not built"); return GP_ERR_UNSUPPORTED; }
int allreduce_probe(Ctx*, int, size_t, uint32_t, uint32_t, double*, double*) { set_error("allreduce probe not built"); return GP_ERR_UNSUPPORTED; }
}

// rk_ivf.cu — IVFFlat (placeholder until the list-probe path lands).
#include "rk_host.h"

int rk_ivf_search(rk_index *, Workspace &, size_t, const float *, const uint32_t *, const rk_scan_opts *, int,
                  const RkSelectParams &)
{
    return rk_fail(RK_ERR_UNSUPPORTED, "IVFFlat scan not built yet");
}
extern "C" int rk_build_ivf(rk_index *, int, uint64_t) { return rk_fail(RK_ERR_UNSUPPORTED, "rk_build_ivf not built yet"); }
extern "C" int rk_build_ivf_from_centers(rk_index *, int, const float *) { return rk_fail(RK_ERR_UNSUPPORTED, "not built yet"); }
extern "C" int rk_ivf_info(const rk_index *, int *, float *, uint32_t *) { return rk_fail(RK_ERR_UNSUPPORTED, "not built yet"); }

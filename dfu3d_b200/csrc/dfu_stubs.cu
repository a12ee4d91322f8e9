// temporary: entry points not implemented yet
#include "dfu_common.cuh"
extern "C" int dfu_sor_filter(dfu_ctx* ctx, void*, const double*, int, const int64_t*, int, double, uint8_t*, double*, double*) { return dfu_set_err(ctx, DFU_ERR_UNSUPPORTED, "not implemented"); }

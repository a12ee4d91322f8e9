// temporary: entry points not implemented yet
#include "dfu_common.cuh"
extern "C" int dfu_fuse_instances_sync(dfu_ctx* ctx, const dfu_inst_params*, int, const double*, int, double*, int64_t, int64_t*, int64_t*) { return dfu_set_err(ctx, DFU_ERR_UNSUPPORTED, "not implemented"); }
extern "C" int dfu_radius_filter(dfu_ctx* ctx, void*, const double*, int, const int64_t*, const double*, const int32_t*, uint8_t*, int32_t*) { return dfu_set_err(ctx, DFU_ERR_UNSUPPORTED, "not implemented"); }
extern "C" int dfu_sor_filter(dfu_ctx* ctx, void*, const double*, int, const int64_t*, int, double, uint8_t*, double*, double*) { return dfu_set_err(ctx, DFU_ERR_UNSUPPORTED, "not implemented"); }
extern "C" int dfu_nn_gate(dfu_ctx* ctx, void*, const double*, int64_t, const double*, int64_t, double, uint8_t*) { return dfu_set_err(ctx, DFU_ERR_UNSUPPORTED, "not implemented"); }
extern "C" int dfu_lshape_fit_sync(dfu_ctx* ctx, void*, const double*, int64_t, double, double, const double*, int, int32_t*, double*, int64_t, int64_t*) { return dfu_set_err(ctx, DFU_ERR_UNSUPPORTED, "not implemented"); }

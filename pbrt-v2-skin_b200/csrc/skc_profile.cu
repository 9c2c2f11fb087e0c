// skc_profile.cu -- MPC_ComputeDiffusionProfile on the device (batched).  Under construction.
#include "skc_internal.h"

extern "C" int skc_profile_batch(skc_ctx *ctx, const skc_layer_spec *specs, const uint32_t *layersPerSet,
                                 const skc_mpc_options *opts, uint32_t nSets, int resample,
                                 uint32_t *lenOut, float *d2, float *R, float *T, float *totals) {
    (void)specs; (void)layersPerSet; (void)opts; (void)nSets; (void)resample; (void)lenOut; (void)d2; (void)R; (void)T; (void)totals;
    return skc_fail(ctx, "skc_profile_batch: not implemented yet");
}

// mpc_shim.cpp -- include/mpc_cuda.h: the reference's MPC_* entry points on top of the device profile generator.
// Reference: multipole/MultipoleProfileCalculator/MultipoleProfileCalculator.{h,cpp} (:294-346 compute, :404-448 resample,
// :451-461 free / clear).  Built as libMultipoleProfileCalculator.so.
#include <mutex>
#include <string>
#include <vector>
#include "../../../include/mpc_cuda.h"
#include "../../../include/skin_cuda.h"

namespace {
std::mutex g_lock;            // one context, one stream: calls are serialised
skc_ctx *g_ctx = nullptr;
std::string g_error;

skc_ctx *context() {          // g_lock held
    if (!g_ctx) {
        g_ctx = skc_create(-1);
        if (!g_ctx) g_error = skc_last_error(nullptr);
    }
    return g_ctx;
}
bool failed(int rc) {
    if (rc) g_error = skc_last_error(g_ctx);
    return rc != 0;
}
void adopt(MPC_Output *o, uint32 length, float *d2, float *R, float *T) {
    delete[] o->pDistanceSquared; delete[] o->pReflectance; delete[] o->pTransmittance;
    o->length = length; o->pDistanceSquared = d2; o->pReflectance = R; o->pTransmittance = T;
}
// new arrays for `queries` squared distances, resampled on the device from o's current arrays
void resampleAt(MPC_Output *o, std::vector<float> &queries) {
    std::lock_guard<std::mutex> hold(g_lock);
    skc_ctx *ctx = context();
    if (!ctx) return;
    const uint32 n = (uint32)queries.size();
    float *d2 = new float[n], *R = new float[n], *T = new float[n];
    if (failed(skc_profile_resample(ctx, o->pDistanceSquared, o->pReflectance, o->pTransmittance, o->length, queries.data(), n, R, T))) {
        delete[] d2; delete[] R; delete[] T;
        return;
    }
    for (uint32 i = 0; i < n; ++i) d2[i] = queries[i];
    adopt(o, n, d2, R, T);
}
}  // namespace

void MPC_CreateOutput(uint32 length, MPC_Output **oppOutput) {
    if (!oppOutput) return;
    MPC_Output *o = new MPC_Output();
    o->length = length;
    o->pDistanceSquared = new float[length]();
    o->pReflectance = new float[length]();
    o->pTransmittance = new float[length]();
    o->totalReflectance = o->totalTransmittance = 0.f;
    *oppOutput = o;
}

void MPC_ComputeDiffusionProfile(uint32 numLayers, const MPC_LayerSpec *pLayerSpecs, const MPC_Options *pOptions, MPC_Output **oppOutput) {
    if (!oppOutput || !numLayers) return;                      // MultipoleProfileCalculator.cpp:297
    *oppOutput = nullptr;
    std::lock_guard<std::mutex> hold(g_lock);
    skc_ctx *ctx = context();
    if (!ctx) return;
    std::vector<skc_layer_spec> specs(numLayers);
    for (uint32 l = 0; l < numLayers; ++l) {
        specs[l].ior = pLayerSpecs[l].ior; specs[l].thickness = pLayerSpecs[l].thickness;
        specs[l].mua = pLayerSpecs[l].mua; specs[l].musp = pLayerSpecs[l].musp; specs[l].g_HG = pLayerSpecs[l].g_HG;
    }
    skc_mpc_options opt;
    opt.desiredStepSize = pOptions->desiredStepSize; opt.desiredLength = pOptions->desiredLength; opt.lerpOnThinSlab = pOptions->lerpOnThinSlab ? 1 : 0;
    uint32_t len = 0;
    if (failed(skc_profile_batch(ctx, specs.data(), &numLayers, &opt, 1, 0, &len, nullptr, nullptr, nullptr, nullptr))) return;
    MPC_Output *o = new MPC_Output();
    o->length = len;
    o->pDistanceSquared = new float[len]; o->pReflectance = new float[len]; o->pTransmittance = new float[len];
    float totals[2] = {0.f, 0.f};
    if (failed(skc_profile_batch(ctx, specs.data(), &numLayers, &opt, 1, 0, &len, o->pDistanceSquared, o->pReflectance, o->pTransmittance, totals))) {
        delete[] o->pDistanceSquared; delete[] o->pReflectance; delete[] o->pTransmittance; delete o;
        return;
    }
    o->totalReflectance = totals[0]; o->totalTransmittance = totals[1];
    *oppOutput = o;
}

void MPC_ResampleForUniformDistanceSquaredDistribution(MPC_Output *pOutput, uint32 target) {
    if (!pOutput || !pOutput->length) return;
    const float extent = pOutput->pDistanceSquared[pOutput->length - 1];
    const uint32 targetLength = target ? target : pOutput->length * 2;      // Nyquist rate (:407)
    std::vector<float> q(targetLength);
    for (uint32 i = 0; i < targetLength; ++i) q[i] = (float)i * extent / (float)(targetLength - 1);
    resampleAt(pOutput, q);
}

void MPC_ResampleDistribution(MPC_Output *pOutput, uint32 nSamplePoints, float *samplePoints) {
    if (!pOutput || !pOutput->length) return;
    std::vector<float> q(nSamplePoints);
    for (uint32 i = 0; i < nSamplePoints; ++i) q[i] = samplePoints[i] * samplePoints[i];
    resampleAt(pOutput, q);
}

void MPC_FreeOutput(MPC_Output *pOutput) {
    if (!pOutput) return;
    delete[] pOutput->pDistanceSquared; delete[] pOutput->pReflectance; delete[] pOutput->pTransmittance;
    delete pOutput;
}

void MPC_ClearCache() {
    std::lock_guard<std::mutex> hold(g_lock);
    if (g_ctx) { skc_destroy(g_ctx); g_ctx = nullptr; }
}

const char *MPC_LastError() { return g_error.c_str(); }

// skc_context.cu -- context lifetime, error reporting, scratch memory, stage timing.
#include <cstdlib>
#include <cstring>
#include <stdarg.h>
#include "skc_common.cuh"
#include "../host/skin_tables.h"      // kRGBIllum2Spect30: data exported from the reference (skc_upload_envmap)

static std::string g_createError;

int skc_fail(skc_ctx *ctx, const char *fmt, ...) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    if (ctx) ctx->err = buf; else g_createError = buf;
    return 1;
}

int skc_check(skc_ctx *ctx, cudaError_t e, const char *what) {
    if (e == cudaSuccess) return 0;
    return skc_fail(ctx, "CUDA error %s (%s) at %s", cudaGetErrorName(e), cudaGetErrorString(e), what);
}

void *skc_scratch(skc_ctx *ctx, int slot, size_t bytes) {
    if (ctx->scratchBytes[slot] >= bytes && ctx->scratch[slot]) return ctx->scratch[slot];
    if (ctx->scratch[slot]) { cudaStreamSynchronize(ctx->stream); cudaFree(ctx->scratch[slot]); ctx->scratch[slot] = nullptr; }
    size_t want = bytes + bytes / 8 + 256;
    if (skc_check(ctx, cudaMalloc(&ctx->scratch[slot], want), "scratch cudaMalloc")) { ctx->scratchBytes[slot] = 0; return nullptr; }
    ctx->scratchBytes[slot] = want;
    return ctx->scratch[slot];
}

void skc_stage_begin(skc_ctx *ctx) { if (!ctx->stageOff) cudaEventRecord(ctx->ev0, ctx->stream); }
void skc_stage_end(skc_ctx *ctx, int stage) {
    if (ctx->stageOff) return;
    cudaEventRecord(ctx->ev1, ctx->stream);
    cudaEventSynchronize(ctx->ev1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    if (ctx->stageAccumulate) ctx->stageMs[stage] += ms; else ctx->stageMs[stage] = ms;
}

cudaEvent_t skc_event(skc_ctx *ctx, size_t i) {
    while (ctx->evPool.size() <= i) {
        cudaEvent_t e = nullptr;
        if (skc_check(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming), "cudaEventCreate")) return nullptr;
        ctx->evPool.push_back(e);
    }
    return ctx->evPool[i];
}

extern "C" skc_ctx *skc_create(int device) {
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        skc_fail(nullptr, "skc_create: no CUDA device (%s); this library has no CPU fallback",
                 e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
        return nullptr;
    }
    if (device < 0) { if (cudaGetDevice(&device) != cudaSuccess) device = 0; }
    if (device >= count) { skc_fail(nullptr, "skc_create: device %d out of range (%d devices)", device, count); return nullptr; }
    if (cudaSetDevice(device) != cudaSuccess) { skc_fail(nullptr, "skc_create: cudaSetDevice(%d) failed", device); return nullptr; }
    skc_ctx *ctx = new skc_ctx();
    ctx->device = device;
    if (cudaDeviceGetAttribute(&ctx->smCount, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || ctx->smCount <= 0) ctx->smCount = 148;
    // Stream priorities (lower number = served first).  The main stream carries the critical path (Preprocess, Mo, film) and
    // gets the highest priority; the second stream (skc_frame_prefetch, the octree topology of skc_preprocess) the lowest,
    // so that its blocks only fill what the main stream's kernels leave free.  Measured on skin.scene, frame with prefetch:
    // equal priorities 123.7 ms, second stream first 123.8 ms, main stream first 121.7 ms (staged flow: 126.5 ms).
    int prioLeast = 0, prioGreatest = 0;
    cudaDeviceGetStreamPriorityRange(&prioLeast, &prioGreatest);
    int prioMain = getenv("SKC_MAIN_PRIO") ? atoi(getenv("SKC_MAIN_PRIO")) : prioGreatest, prioAux = getenv("SKC_AUX_PRIO") ? atoi(getenv("SKC_AUX_PRIO")) : prioLeast;
    if (cudaStreamCreateWithPriority(&ctx->stream, cudaStreamNonBlocking, prioMain) != cudaSuccess ||
        cudaStreamCreateWithFlags(&ctx->copyStream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithPriority(&ctx->auxStream, cudaStreamNonBlocking, prioAux) != cudaSuccess ||
        cudaStreamCreateWithPriority(&ctx->prefetchStream, cudaStreamNonBlocking, prioAux) != cudaSuccess ||
        cudaEventCreate(&ctx->defEv0) != cudaSuccess || cudaEventCreate(&ctx->defEv1) != cudaSuccess ||
        cudaEventCreate(&ctx->ev0) != cudaSuccess || cudaEventCreate(&ctx->ev1) != cudaSuccess) {
        skc_fail(nullptr, "skc_create: stream/event creation failed: %s", cudaGetErrorString(cudaGetLastError()));
        delete ctx;
        return nullptr;
    }
    if (const char *k = getenv("SKC_MO_KERNEL")) ctx->moPacket = strcmp(k, "warp") != 0;
    if (cudaDeviceGetAttribute(&ctx->smemOptin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device) != cudaSuccess) ctx->smemOptin = 48 * 1024;
    if (const char *k = getenv("SKC_MO_GROUP")) ctx->moGroup = atoi(k);
    if (const char *k = getenv("SKC_MO_PACKET_HITS")) { int v = atoi(k); ctx->moPacketHits = (v == 8 || v == 16 || v == 32) ? v : 0; }
    if (const char *k = getenv("SKC_MO_WARPS")) ctx->moWarps = atoi(k);
    if (const char *k = getenv("SKC_BVH_WIDE")) ctx->bvhWideOn = atoi(k) != 0;
    if (const char *k = getenv("SKC_MO_TABLE")) ctx->moTable = !strcmp(k, "exact") ? 0 : !strcmp(k, "compact") ? 2 : 1;
    if (const char *k = getenv("SKC_MO_TOL")) { float t = (float)atof(k); if (t > 0.f) ctx->mocTol = t; }
    if (const char *k = getenv("SKC_MO_SORT")) ctx->moSort = atoi(k);
    if (const char *k = getenv("SKC_MO_SPLIT")) ctx->moSplit = atoi(k) == 1 ? 1 : atoi(k) == 2 ? 2 : atoi(k) == 4 ? 4 : 0;
    return ctx;
}

extern "C" void skc_destroy(skc_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    void *ptrs[] = {ctx->meshIdx, ctx->bvhNodes, ctx->bvhWide, ctx->bvhPrims, ctx->primInfo, ctx->meshP, ctx->meshN, ctx->meshUV, ctx->spheres,
                    ctx->profPairs, ctx->profRcp, ctx->rhoHd, ctx->mocRows, ctx->mocHeads, ctx->mocLam, ctx->mocHeadOff, ctx->lights, ctx->ipPosArea, ctx->ipE,
                    ctx->otGeo, ctx->otEt, ctx->resPts, ctx->resRays, ctx->otPointIdx, ctx->otLeafPosArea, ctx->albedoFactors, ctx->envStore, ctx->alphaTex, ctx->alphaUV, ctx->albedoTexStore, ctx->bumpTexStore};
    for (void *p : ptrs) if (p) cudaFree(p);
    skc_frame_release(ctx);
    for (int i = 0; i < 24; ++i) if (ctx->scratch[i]) cudaFree(ctx->scratch[i]);
    cudaEventDestroy(ctx->ev0); cudaEventDestroy(ctx->ev1);
    if (ctx->timer0) { cudaEventDestroy(ctx->timer0); cudaEventDestroy(ctx->timer1); }
    for (cudaEvent_t e : ctx->evPool) cudaEventDestroy(e);
    cudaStreamSynchronize(ctx->copyStream);
    cudaStreamDestroy(ctx->copyStream);
    cudaStreamSynchronize(ctx->auxStream);
    cudaStreamDestroy(ctx->auxStream);
    cudaStreamSynchronize(ctx->prefetchStream);
    cudaStreamDestroy(ctx->prefetchStream);
    cudaEventDestroy(ctx->defEv0); cudaEventDestroy(ctx->defEv1);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" const char *skc_last_error(const skc_ctx *ctx) {
    return ctx ? ctx->err.c_str() : g_createError.c_str();
}

extern "C" int skc_synchronize(skc_ctx *ctx) {
    cudaSetDevice(ctx->device);
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

extern "C" uint64_t skc_launch_count(const skc_ctx *ctx) { return ctx->launches; }
// LayeredSkin's "texture albedo" (materials/layeredskin.cpp:183,247), an image map: nLevels levels of its MIPMap<RGBSpectrum> (sizes w[], h[],
// texels rgb = the levels one after the other, [h][w][3] each, as ImageTexture holds them: scale * (Pow(rgb, gamma) + shift)), the EWA
// filter's weight table (128 entries, core/mipmap.h:186-193) and the UVMapping2D.  The irradiance pass scales a point's E by
// Pow(albedo(u, v), mix) (lookup with zero differentials = bilinear at level 0), Li a hit's subsurface term by Pow(albedo, 1 - mix)
// with the EWA lookup under the camera ray's differentials.  nLevels == 0 removes it.
static int uploadTexture(skc_ctx *ctx, const char *what, float **store, AlbedoTexDev *dev, uint32_t nLevels, const int *w, const int *h, const float *texels, int channels,
                         const float *weightLut128, float su, float sv, float du, float dv) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (*store) { cudaFree(*store); *store = nullptr; }
    *dev = AlbedoTexDev();
    if (!nLevels) return 0;
    if (nLevels > SKC_TEX_MAX_LEVELS || !w || !h || !texels || !weightLut128) return skc_fail(ctx, "%s: bad arguments (at most %d levels)", what, SKC_TEX_MAX_LEVELS);
    AlbedoTexDev t;
    size_t total = 0;
    for (uint32_t l = 0; l < nLevels; ++l) { t.w[l] = w[l]; t.h[l] = h[l]; t.off[l] = (uint32_t)total; total += (size_t)3 * w[l] * h[l]; }
    std::vector<float> expanded;
    if (channels == 1) {          // a float texture: the same filter code on three equal channels
        expanded.resize(total);
        for (size_t k = 0; k < total / 3; ++k) expanded[3 * k] = expanded[3 * k + 1] = expanded[3 * k + 2] = texels[k];
        texels = expanded.data();
    }
    float refl[7 * 32];
    for (int k = 0; k < 7; ++k) for (int i = 0; i < 32; ++i) refl[k * 32 + i] = i < 30 ? kRGBRefl2Spect30[k][i] : 0.f;
    const size_t texFloats = (total + 3) & ~(size_t)3;
    SKC_CUDA(cudaMalloc((void **)store, (texFloats + 128 + 7 * 32) * sizeof(float)));
    SKC_CUDA(cudaMemcpyAsync(*store, texels, total * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    SKC_CUDA(cudaMemcpyAsync(*store + texFloats, weightLut128, 128 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    SKC_CUDA(cudaMemcpyAsync(*store + texFloats + 128, refl, sizeof refl, cudaMemcpyHostToDevice, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    t.rgb = *store; t.lut = *store + texFloats; t.refl = t.lut + 128;
    t.nLevels = (int)nLevels; t.su = su; t.sv = sv; t.du = du; t.dv = dv;
    *dev = t;
    return 0;
}

extern "C" int skc_upload_albedo_texture(skc_ctx *ctx, uint32_t nLevels, const int *w, const int *h, const float *rgb, const float *weightLut128,
                                         float su, float sv, float du, float dv) {
    if (!ctx) return 1;
    return uploadTexture(ctx, "skc_upload_albedo_texture", &ctx->albedoTexStore, &ctx->albedoTex, nLevels, w, h, rgb, 3, weightLut128, su, sv, du, dv);
}

// The LayeredSkin material's displacement texture ("texture bumpmap", a float imagemap: materials/layeredskin.cpp:139,299-304): MIP pyramid
// of one float per texel, finest first, the same filter table and UV mapping as the albedo texture.  It bends the normals of the tessellated
// irradiance points (trianglemesh.cpp:236-242) and of every skin hit's BSDF (core/material.cpp:47-104).  nLevels == 0 removes it.
extern "C" int skc_upload_bump_texture(skc_ctx *ctx, uint32_t nLevels, const int *w, const int *h, const float *displacement, const float *weightLut128,
                                       float su, float sv, float du, float dv) {
    if (!ctx) return 1;
    return uploadTexture(ctx, "skc_upload_bump_texture", &ctx->bumpTexStore, &ctx->bumpTex, nLevels, w, h, displacement, 1, weightLut128, su, sv, du, dv);
}

// The radiance map and sampling distribution of the scene's infinite light (light record with radius -4).  w, h: texels; rgb
// [h][w][3]; condFunc [h][w], condCdf [h][w + 1], condInt [h]; margFunc [h], margCdf [h + 1].  w == 0 removes it.
extern "C" int skc_upload_envmap(skc_ctx *ctx, uint32_t w, uint32_t h, const float *rgb, const float *condFunc, const float *condCdf,
                                 const float *condInt, const float *margFunc, const float *margCdf, float margInt) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->envStore) { cudaFree(ctx->envStore); ctx->envStore = nullptr; }
    ctx->env = EnvMapDev();
    if (!w || !h) return 0;
    if (!rgb || !condFunc || !condCdf || !condInt || !margFunc || !margCdf) return skc_fail(ctx, "skc_upload_envmap: null array");
    float illum[7 * 32];
    for (int k = 0; k < 7; ++k) for (int i = 0; i < 32; ++i) illum[k * 32 + i] = i < 30 ? kRGBIllum2Spect30[k][i] : 0.f;
    const size_t n = (size_t)w * h, sizes[7] = {3 * n, n, (size_t)h * (w + 1), h, h, (size_t)h + 1, 7 * 32};
    const float *src[7] = {rgb, condFunc, condCdf, condInt, margFunc, margCdf, illum};
    size_t total = 0, off[7];
    for (int k = 0; k < 7; ++k) { off[k] = total; total += (sizes[k] + 3) & ~(size_t)3; }
    SKC_CUDA(cudaMalloc((void **)&ctx->envStore, total * sizeof(float)));
    for (int k = 0; k < 7; ++k) SKC_CUDA(cudaMemcpyAsync(ctx->envStore + off[k], src[k], sizes[k] * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    EnvMapDev &e = ctx->env;
    e.rgb = ctx->envStore + off[0]; e.condFunc = ctx->envStore + off[1]; e.condCdf = ctx->envStore + off[2]; e.condInt = ctx->envStore + off[3];
    e.margFunc = ctx->envStore + off[4]; e.margCdf = ctx->envStore + off[5]; e.illum = ctx->envStore + off[6]; e.margInt = margInt; e.w = (int)w; e.h = (int)h;
    return 0;
}

// LayeredSkin's "color albedo" (materials/layeredskin.cpp:183,247): the irradiance pass scales E by Pow(albedo, mix) and Li the
// subsurface term by Pow(albedo, 1 - mix) (multipolesubsurface.cpp:139-140,287).  For a constant albedo both are 30 numbers,
// made by the caller with the host's powf so that they are the reference's own values.  NULL, NULL = Spectrum(1).
extern "C" int skc_set_albedo_factors(skc_ctx *ctx, const float *powMix30, const float *powOneMinusMix30) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!powMix30 || !powOneMinusMix30) {
        if (ctx->albedoFactors) { cudaStreamSynchronize(ctx->stream); cudaFree(ctx->albedoFactors); ctx->albedoFactors = nullptr; }
        return 0;
    }
    float h[64];
    for (int i = 0; i < 32; ++i) { h[i] = i < SKC_NS ? powMix30[i] : 1.f; h[32 + i] = i < SKC_NS ? powOneMinusMix30[i] : 1.f; }
    if (!ctx->albedoFactors) SKC_CUDA(cudaMalloc((void **)&ctx->albedoFactors, sizeof h));
    SKC_CUDA(cudaMemcpyAsync(ctx->albedoFactors, h, sizeof h, cudaMemcpyHostToDevice, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

extern "C" int skc_sm_count(const skc_ctx *ctx) { return ctx ? ctx->smCount : 0; }
extern "C" float skc_stage_ms(const skc_ctx *ctx, int stage) {
    return (stage >= 0 && stage < SKC_NSTAGES) ? ctx->stageMs[stage] : -1.f;
}

// whole-step timing on the context's stream (bench.py): CUDA events, not wall clock
// (the events belong to the context: a second context on another GPU of the same process must not record events of a foreign device)
extern "C" int skc_timer_begin(skc_ctx *ctx) {
    cudaSetDevice(ctx->device);
    if (!ctx->timer0) { SKC_CUDA(cudaEventCreate(&ctx->timer0)); SKC_CUDA(cudaEventCreate(&ctx->timer1)); }
    SKC_CUDA(cudaEventRecord(ctx->timer0, ctx->stream));
    return 0;
}
extern "C" float skc_timer_end_ms(skc_ctx *ctx) {
    cudaSetDevice(ctx->device);
    if (!ctx->timer0) return -1.f;
    if (cudaEventRecord(ctx->timer1, ctx->stream) != cudaSuccess || cudaEventSynchronize(ctx->timer1) != cudaSuccess) return -1.f;
    float ms = -1.f;
    cudaEventElapsedTime(&ms, ctx->timer0, ctx->timer1);
    return ms;
}
extern "C" int skc_set_point_count(skc_ctx *ctx, uint64_t n) {
    if (n > ctx->ipCap) return skc_fail(ctx, "skc_set_point_count: %llu exceeds capacity", (unsigned long long)n);
    ctx->ipCount = n;
    return 0;
}

// Rows first, first+stride, ... (count of them) of the device irradiance-point set: {p, area} and E[30] (IrradiancePoint::p / area / E,
// integrators/irradiancepoint.h:36-45).  For parity checks of sets too large to copy out whole.  Either output may be NULL.
extern "C" int skc_get_irradiance_rows(skc_ctx *ctx, uint64_t first, uint64_t count, uint64_t stride, float *posArea4, float *E30) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!count) return 0;
    if (!stride || first + (count - 1) * stride >= ctx->ipCount) return skc_fail(ctx, "skc_get_irradiance_rows: range beyond the %llu device points", (unsigned long long)ctx->ipCount);
    if (posArea4) SKC_CUDA(cudaMemcpy2DAsync(posArea4, sizeof(float4), ctx->ipPosArea + first, stride * sizeof(float4), sizeof(float4), count, cudaMemcpyDeviceToHost, ctx->stream));
    if (E30) SKC_CUDA(cudaMemcpy2DAsync(E30, SKC_NS * sizeof(float), ctx->ipE + first * SKC_EPAD, stride * SKC_EPAD * sizeof(float), SKC_NS * sizeof(float), count, cudaMemcpyDeviceToHost, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// The counterpart of skc_get_irradiance_rows: overwrite E (and, if given, {p, area}) of `count` device points starting at `first`.
// For parity checks that need one or two rows replaced by the reference's (a shadow-ray visibility flip in a set of millions).
extern "C" int skc_set_irradiance_rows(skc_ctx *ctx, uint64_t first, uint64_t count, const float *posArea4, const float *E30) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!count) return 0;
    if (first + count > ctx->ipCount) return skc_fail(ctx, "skc_set_irradiance_rows: range beyond the %llu device points", (unsigned long long)ctx->ipCount);
    if (posArea4) SKC_CUDA(cudaMemcpyAsync(ctx->ipPosArea + first, posArea4, count * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
    if (E30) SKC_CUDA(cudaMemcpy2DAsync(ctx->ipE + first * SKC_EPAD, SKC_EPAD * sizeof(float), E30, SKC_NS * sizeof(float), SKC_NS * sizeof(float), count, cudaMemcpyHostToDevice, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// Test hook: the device's sinf / cosf (skc_common.cuh: glibc's algorithm restated) on n host floats.
__global__ void k_test_sincosf(const float *__restrict__ x, uint64_t n, float *__restrict__ s, float *__restrict__ c) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { s[i] = skc_sinf(x[i]); c[i] = skc_cosf(x[i]); }
}
__global__ void k_test_libmf(int fn, const float *__restrict__ x, uint64_t n, float *__restrict__ out) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = fn == 7 ? skc_powf(x[i], x[n - 1 - i]) : fn == 6 ? skc_atan2f(x[i], x[n - 1 - i]) : fn == 5 ? skc_acosf(x[i]) : fn == 2 ? skc_logf(x[i]) : fn == 3 ? skc_expf(x[i]) : fn == 4 ? skc_atanf(x[i]) : fn == 1 ? skc_cosf(x[i]) : skc_sinf(x[i]);
}
// fn: 0 sinf, 1 cosf, 2 logf, 3 expf, 4 atanf, 5 acosf, 6 atan2f(x[i], x[n - 1 - i]), 7 powf(x[i], x[n - 1 - i])
extern "C" int skc_test_libmf(skc_ctx *ctx, int fn, const float *x, uint64_t n, float *out) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!n) return 0;
    float *d = (float *)skc_scratch(ctx, 0, 2 * n * sizeof(float));
    if (!d) return 1;
    SKC_CUDA(cudaMemcpyAsync(d, x, n * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    k_test_libmf<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(fn, d, n, d + n);
    SKC_LAUNCHED("k_test_libmf");
    SKC_CUDA(cudaMemcpyAsync(out, d + n, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}
extern "C" int skc_test_sincosf(skc_ctx *ctx, const float *x, uint64_t n, float *sinOut, float *cosOut) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!n) return 0;
    float *d = (float *)skc_scratch(ctx, 0, 3 * n * sizeof(float));
    if (!d) return 1;
    SKC_CUDA(cudaMemcpyAsync(d, x, n * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    k_test_sincosf<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(d, n, d + n, d + 2 * n);
    SKC_LAUNCHED("k_test_sincosf");
    SKC_CUDA(cudaMemcpyAsync(sinOut, d + n, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    SKC_CUDA(cudaMemcpyAsync(cosOut, d + 2 * n, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

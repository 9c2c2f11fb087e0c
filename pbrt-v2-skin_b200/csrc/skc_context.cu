// skc_context.cu -- context lifetime, error reporting, scratch memory, stage timing.
#include <cstdlib>
#include <cstring>
#include <stdarg.h>
#include "skc_common.cuh"

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
    if (const char *k = getenv("SKC_MO_WARPS")) ctx->moWarps = atoi(k);
    if (const char *k = getenv("SKC_MO_TABLE")) ctx->moTable = strcmp(k, "exact") != 0;
    if (const char *k = getenv("SKC_MO_TOL")) { float t = (float)atof(k); if (t > 0.f) ctx->mocTol = t; }
    if (const char *k = getenv("SKC_MO_SORT")) ctx->moSort = atoi(k);
    if (const char *k = getenv("SKC_MO_SPLIT")) ctx->moSplit = atoi(k) == 1 ? 1 : atoi(k) == 2 ? 2 : atoi(k) == 4 ? 4 : 0;
    return ctx;
}

extern "C" void skc_destroy(skc_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    void *ptrs[] = {ctx->meshIdx, ctx->bvhNodes, ctx->bvhPrims, ctx->primInfo, ctx->meshP, ctx->meshN, ctx->meshUV, ctx->spheres,
                    ctx->profPairs, ctx->profRcp, ctx->rhoHd, ctx->mocRows, ctx->mocHeads, ctx->mocLam, ctx->mocHeadOff, ctx->lights, ctx->ipPosArea, ctx->ipE,
                    ctx->otGeo, ctx->otEt, ctx->resPts, ctx->resRays, ctx->otPointIdx, ctx->otLeafPosArea, ctx->otLeafE};
    for (void *p : ptrs) if (p) cudaFree(p);
    skc_frame_release(ctx);
    for (int i = 0; i < 20; ++i) if (ctx->scratch[i]) cudaFree(ctx->scratch[i]);
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

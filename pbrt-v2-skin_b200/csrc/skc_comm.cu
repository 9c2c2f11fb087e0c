// skc_comm.cu -- the two exchange steps of the multi-GPU path, as NCCL calls behind the C ABI.
//
// The reference is a single process (core/parallel.cpp:722-748); nothing in it corresponds to this file.  The path shards in
// two places: irradiance points (each rank computes a contiguous range; every rank needs all of them for its replica of the
// octree -> ncclAllGather, in place, over NVLink) and camera tiles (each rank renders its render tasks into its own film; the
// reconstruction filter spreads samples across tile borders -> ncclReduce(sum) of the films to the root, not a gather).
// There is no arithmetic between compute and exchange that could be fused into one kernel: both are pure data movement, and
// the big one (128 B/point irradiance rows, 3.15 GB at 24.6 M points) is overlapped with the octree topology pass, which needs
// the 16 B/point positions only -- positions travel first on the main stream, rows on the copy stream.
//
// One communicator per context.  Bootstrap either way: one process per GPU (skc_comm_unique_id on rank 0, the 128 bytes
// handed to the other ranks by whatever the launcher offers, skc_comm_create everywhere) or one process driving several
// GPUs (skc_comm_create_local, one host thread per context afterwards).
#include <nccl.h>
#include "skc_internal.h"

struct skc_comm {
    skc_ctx *ctx = nullptr;
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1;
    cudaEvent_t rowsReady = nullptr, rowsDone = nullptr;
    float *scalar = nullptr;          // device word for the max-reduce of timings
};

static int ncclFail(skc_ctx *ctx, ncclResult_t r, const char *what) {
    if (r == ncclSuccess) return 0;
    return skc_fail(ctx, "NCCL error %s at %s", ncclGetErrorString(r), what);
}
#define SKC_NCCL(call) do { if (ncclFail(c->ctx, (call), #call)) return 1; } while (0)

extern "C" int skc_comm_unique_id(uint8_t *id128) {
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    if (ncclGetUniqueId(&id) != ncclSuccess) return 1;
    memcpy(id128, &id, sizeof id);
    return 0;
}

static skc_comm *finishComm(skc_ctx *ctx, ncclComm_t comm, int rank, int world) {
    skc_comm *c = new skc_comm();
    c->ctx = ctx; c->comm = comm; c->rank = rank; c->world = world;
    cudaSetDevice(ctx->device);
    if (cudaEventCreateWithFlags(&c->rowsReady, cudaEventDisableTiming) != cudaSuccess || cudaEventCreateWithFlags(&c->rowsDone, cudaEventDisableTiming) != cudaSuccess ||
        cudaMalloc((void **)&c->scalar, 256) != cudaSuccess) {
        skc_fail(ctx, "skc_comm: event / scratch creation failed");
        ncclCommDestroy(comm);
        delete c;
        return nullptr;
    }
    return c;
}

extern "C" skc_comm *skc_comm_create(skc_ctx *ctx, int rank, int world, const uint8_t *id128) {
    if (!ctx || world < 1 || rank < 0 || rank >= world || !id128) { skc_fail(ctx, "skc_comm_create: bad rank %d of %d", rank, world); return nullptr; }
    cudaSetDevice(ctx->device);
    ncclUniqueId id;
    memcpy(&id, id128, sizeof id);
    ncclComm_t comm = nullptr;
    ncclResult_t r = ncclCommInitRank(&comm, world, id, rank);
    if (r != ncclSuccess) { skc_fail(ctx, "skc_comm_create: ncclCommInitRank: %s", ncclGetErrorString(r)); return nullptr; }
    return finishComm(ctx, comm, rank, world);
}

extern "C" int skc_comm_create_local(skc_ctx **ctxs, int n, skc_comm **comms) {
    if (!ctxs || !comms || n < 1 || n > 64) return 1;
    int devs[64]; ncclComm_t raw[64];
    for (int i = 0; i < n; ++i) { if (!ctxs[i]) return 1; devs[i] = ctxs[i]->device; comms[i] = nullptr; }
    ncclResult_t r = ncclCommInitAll(raw, n, devs);
    if (r != ncclSuccess) return skc_fail(ctxs[0], "skc_comm_create_local: ncclCommInitAll: %s", ncclGetErrorString(r));
    for (int i = 0; i < n; ++i) {
        comms[i] = finishComm(ctxs[i], raw[i], i, n);
        if (!comms[i]) return 1;
    }
    return 0;
}

extern "C" void skc_comm_destroy(skc_comm *c) {
    if (!c) return;
    cudaSetDevice(c->ctx->device);
    cudaStreamSynchronize(c->ctx->stream); cudaStreamSynchronize(c->ctx->copyStream);
    ncclCommDestroy(c->comm);
    cudaEventDestroy(c->rowsReady); cudaEventDestroy(c->rowsDone);
    cudaFree(c->scalar);
    delete c;
}
extern "C" int skc_comm_rank(const skc_comm *c) { return c ? c->rank : 0; }
extern "C" int skc_comm_world(const skc_comm *c) { return c ? c->world : 1; }

// Every rank has computed the points [rank * shard, (rank + 1) * shard) of its context's irradiance arrays (the last shard may
// be padded; skc_set_point_count afterwards gives the true total).  Positions ({p, area}, 16 B/point): in place, on the main
// stream, so that the octree topology pass that follows sees all of them.
extern "C" int skc_comm_allgather_positions(skc_comm *c, uint64_t shard) {
    if (!c) return 1;
    cudaSetDevice(c->ctx->device);
    if (skc_ensure_points(c->ctx, shard * c->world)) return 1;
    float4 *base = c->ctx->ipPosArea;
    SKC_NCCL(ncclAllGather(base + (size_t)c->rank * shard, base, shard * 4, ncclFloat, c->comm, c->ctx->stream));
    return 0;
}
// Irradiance rows (128 B/point): in place, on the copy stream, behind everything the main stream has enqueued so far; returns at
// once.  skc_comm_allgather_rows_wait makes the main stream wait for them (on the device; the host does not block).
extern "C" int skc_comm_allgather_rows_begin(skc_comm *c, uint64_t shard) {
    if (!c) return 1;
    skc_ctx *ctx = c->ctx;
    cudaSetDevice(ctx->device);
    if (skc_ensure_points(ctx, shard * c->world)) return 1;
    SKC_CUDA(cudaEventRecord(c->rowsReady, ctx->stream));
    SKC_CUDA(cudaStreamWaitEvent(ctx->copyStream, c->rowsReady, 0));
    float *base = ctx->ipE;
    SKC_NCCL(ncclAllGather(base + (size_t)c->rank * shard * SKC_EPAD, base, shard * SKC_EPAD, ncclFloat, c->comm, ctx->copyStream));
    SKC_CUDA(cudaEventRecord(c->rowsDone, ctx->copyStream));
    return 0;
}
extern "C" int skc_comm_allgather_rows_wait(skc_comm *c) {
    if (!c) return 1;
    skc_ctx *ctx = c->ctx;
    cudaSetDevice(ctx->device);
    SKC_CUDA(cudaStreamWaitEvent(ctx->stream, c->rowsDone, 0));
    return 0;
}

// ImageFilm pixels {Lxyz[3], weightSum} of every rank summed into the root's film (film/image.cpp:100-137 adds filtered samples
// into neighbouring pixels across tile borders: a sum, not a gather).  On the main stream, after the rank's skc_frame_render.
extern "C" int skc_comm_reduce_film(skc_comm *c, int root) {
    if (!c) return 1;
    skc_ctx *ctx = c->ctx;
    cudaSetDevice(ctx->device);
    uint64_t floats = 0;
    float *film = (float *)skc_frame_film_floats(ctx, &floats);
    if (!film) return skc_fail(ctx, "skc_comm_reduce_film: no frame set up (skc_frame_setup)");
    SKC_NCCL(ncclReduce(film, film, floats, ncclFloat, ncclSum, root, c->comm, ctx->stream));
    return 0;
}

// max over ranks of a host value (device timings are reported as the max over ranks); blocks until the result is back
extern "C" int skc_comm_max_f32(skc_comm *c, float *value) {
    if (!c || !value) return 1;
    skc_ctx *ctx = c->ctx;
    cudaSetDevice(ctx->device);
    SKC_CUDA(cudaMemcpyAsync(c->scalar, value, sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    SKC_NCCL(ncclAllReduce(c->scalar, c->scalar, 1, ncclFloat, ncclMax, c->comm, ctx->stream));
    SKC_CUDA(cudaMemcpyAsync(value, c->scalar, sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}
// all ranks' main streams have reached this point (an all-reduce of one word) and the host has seen it
extern "C" int skc_comm_barrier(skc_comm *c) {
    float one = 1.f;
    return skc_comm_max_f32(c, &one);
}

// a block of host bytes from every rank to every rank (small set-up data: material tables computed rank by rank)
extern "C" int skc_comm_allgather_host(skc_comm *c, void *buf, uint64_t bytesPerRank) {
    if (!c || !buf) return 1;
    skc_ctx *ctx = c->ctx;
    cudaSetDevice(ctx->device);
    const uint64_t total = bytesPerRank * c->world;
    char *d = (char *)skc_scratch(ctx, 20, total);
    if (!d) return 1;
    SKC_CUDA(cudaMemcpyAsync(d + (size_t)c->rank * bytesPerRank, (char *)buf + (size_t)c->rank * bytesPerRank, bytesPerRank, cudaMemcpyHostToDevice, ctx->stream));
    SKC_NCCL(ncclAllGather(d + (size_t)c->rank * bytesPerRank, d, bytesPerRank, ncclChar, c->comm, ctx->stream));
    SKC_CUDA(cudaMemcpyAsync(buf, d, total, cudaMemcpyDeviceToHost, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// skc_octree.cu -- SubsurfaceOctreeNode::Insert + InitHierarchy on the device.
//
// Reference: integrators/diffusionutil.h:94-132 (Insert), :133-173 (InitHierarchy),
// core/octree.h:87-97 (octreeChildBound), integrators/multipolesubsurface.cpp:218-237
// (bounds = union of points, insertion in point order).
//
// The reference inserts points one by one; the resulting CELL DECOMPOSITION does not depend
// on the order (a cell is split iff it holds more than 8 points; child = 4*(x>mid) +
// 2*(y>mid) + (z>mid) with mid = .5f*min + .5f*max recomputed top-down), and inside a leaf
// the points sit in ascending insertion (= point index) order.  So the tree is rebuilt
// breadth-first: every level partitions the points of each over-full cell into its eight
// children with warp-aggregated atomics, children of one cell are allocated contiguously
// (flat breadth-first node array), and each leaf's <= 8 points are finally sorted by index.
// InitHierarchy then runs level by level from the deepest level up, one thread per node,
// with the reference's exact fp32 operation order (the accept/refine decision of Mo depends
// on node p and sumArea bit for bit, SURVEY.md section 7 hard part #1).
//
// Node record (3 x float4 = 48 B):  {p.xyz, sumArea} {min.xyz, max.x} {max.y, max.z, w0, w1}
//   w0 = first child (interior) or first point slot (leaf)
//   w1 = child mask | SKC_OT_LEAF | SKC_OT_BLACK ; leaf: point count in bits 0-7
#include <cstdlib>
#include <cstring>
#include <vector>
#include <algorithm>
#include "skc_internal.h"
#include "skc_cie.h"

#define OT_DONE 0xffffffffu
#define OT_MAX_LEVELS 64

__constant__ float c_cieY[32];

// ---- bounds -----------------------------------------------------------------------------
__global__ void k_bounds_partial(const float4 *__restrict__ pa, uint64_t n, float *__restrict__ partial) {
    __shared__ float s[6][256];
    float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        float4 v = __ldg(&pa[i]);
        mn[0] = fminf(mn[0], v.x); mn[1] = fminf(mn[1], v.y); mn[2] = fminf(mn[2], v.z);
        mx[0] = fmaxf(mx[0], v.x); mx[1] = fmaxf(mx[1], v.y); mx[2] = fmaxf(mx[2], v.z);
    }
    for (int k = 0; k < 3; ++k) { s[k][threadIdx.x] = mn[k]; s[3 + k][threadIdx.x] = mx[k]; }
    __syncthreads();
    for (int st = 128; st > 0; st >>= 1) {
        if (threadIdx.x < st)
            for (int k = 0; k < 3; ++k) {
                s[k][threadIdx.x] = fminf(s[k][threadIdx.x], s[k][threadIdx.x + st]);
                s[3 + k][threadIdx.x] = fmaxf(s[3 + k][threadIdx.x], s[3 + k][threadIdx.x + st]);
            }
        __syncthreads();
    }
    if (threadIdx.x < 6) partial[blockIdx.x * 6 + threadIdx.x] = s[threadIdx.x][0];
}

// ---- level kernels ------------------------------------------------------------------------
struct OtBuild {
    float4 *nodes;          // 3 per node
    uint32_t *nodeStart, *nodeCount;
    uint32_t *cnt;          // per level-node 8 counters -> exclusive prefix
    uint32_t *counters;     // [0] total nodes allocated, [1] overflow flag, [2] active nodes next level
    uint32_t cap;
};

__global__ void k_init_root(OtBuild b, uint32_t n, float4 b0, float4 b1) {
    b.nodes[1] = b0;
    b.nodes[2] = make_float4(b1.x, b1.y, __uint_as_float(0u), __uint_as_float(n <= 8 ? (SKC_OT_LEAF | n) : 0u));
    b.nodes[0] = make_float4(0.f, 0.f, 0.f, 0.f);
    b.nodeStart[0] = 0; b.nodeCount[0] = n;
    b.counters[0] = 1; b.counters[1] = 0; b.counters[2] = n > 8 ? 1 : 0;
}

__global__ void k_fill_u32(uint32_t *p, uint64_t n, uint32_t v, int iota) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = iota ? (uint32_t)i : v;
}

// K1: digit of every point that sits in an over-full cell + its rank inside (cell, digit)
__global__ void __launch_bounds__(256) k_level_count(OtBuild b, uint32_t levelStart, const float4 *__restrict__ posArea,
        const uint32_t *__restrict__ idx, const uint32_t *__restrict__ slotNode, uint32_t n, uint32_t *__restrict__ digitRank) {
    uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t node = slot < n ? slotNode[slot] : OT_DONE;
    bool active = node != OT_DONE;
    uint32_t key = OT_DONE, digit = 0;
    if (active) {
        float4 b0 = __ldg(&b.nodes[3 * (size_t)node + 1]), b1 = __ldg(&b.nodes[3 * (size_t)node + 2]);
        float4 p = __ldg(&posArea[idx[slot]]);
        float mx = .5f * b0.x + .5f * b0.w, my = .5f * b0.y + .5f * b1.x, mz = .5f * b0.z + .5f * b1.y;   // diffusionutil.h:96
        digit = (p.x > mx ? 4u : 0u) + (p.y > my ? 2u : 0u) + (p.z > mz ? 1u : 0u);                    // :116-117
        key = (node - levelStart) * 8u + digit;
    }
    unsigned act = __ballot_sync(0xffffffffu, active);
    if (!active) return;
    unsigned peers = __match_any_sync(act, key);
    int leader = __ffs(peers) - 1;
    int lane = threadIdx.x & 31;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(&b.cnt[key], (uint32_t)__popc(peers));
    base = __shfl_sync(peers, base, leader);
    uint32_t rank = base + __popc(peers & ((1u << lane) - 1u));
    digitRank[slot] = (digit << 29) | rank;
}

// K2: one thread per cell of the level: allocate children, write their bounds and ranges
__global__ void __launch_bounds__(128) k_level_split(OtBuild b, uint32_t levelStart, uint32_t levelCount) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= levelCount) return;
    uint32_t node = levelStart + t;
    uint32_t count = b.nodeCount[node];
    if (count <= 8) return;                 // leaf, finished at creation
    uint32_t c[8], pre[8], mask = 0, nCh = 0, run = 0;
    for (int k = 0; k < 8; ++k) { c[k] = b.cnt[t * 8 + k]; pre[k] = run; run += c[k]; if (c[k]) { mask |= 1u << k; ++nCh; } }
    uint32_t first = atomicAdd(&b.counters[0], nCh);
    float4 b0 = b.nodes[3 * (size_t)node + 1], b1 = b.nodes[3 * (size_t)node + 2];
    if (first + nCh > b.cap) { atomicExch(&b.counters[1], 1u); return; }
    b.nodes[3 * (size_t)node + 2] = make_float4(b1.x, b1.y, __uint_as_float(first), __uint_as_float(mask));
    float mnx = b0.x, mny = b0.y, mnz = b0.z, mxx = b0.w, mxy = b1.x, mxz = b1.y;
    float mx = .5f * mnx + .5f * mxx, my = .5f * mny + .5f * mxy, mz = .5f * mnz + .5f * mxz;
    uint32_t start = b.nodeStart[node], r = 0, nActive = 0;
    for (int k = 0; k < 8; ++k) {
        b.cnt[t * 8 + k] = pre[k];
        if (!c[k]) continue;
        uint32_t child = first + r++;
        // octreeChildBound (core/octree.h:87-97)
        float cminx = (k & 4) ? mx : mnx, cmaxx = (k & 4) ? mxx : mx;
        float cminy = (k & 2) ? my : mny, cmaxy = (k & 2) ? mxy : my;
        float cminz = (k & 1) ? mz : mnz, cmaxz = (k & 1) ? mxz : mz;
        b.nodes[3 * (size_t)child] = make_float4(0.f, 0.f, 0.f, 0.f);
        b.nodes[3 * (size_t)child + 1] = make_float4(cminx, cminy, cminz, cmaxx);
        uint32_t w1 = c[k] <= 8 ? (SKC_OT_LEAF | c[k]) : 0u;
        b.nodes[3 * (size_t)child + 2] = make_float4(cmaxy, cmaxz, __uint_as_float(start + pre[k]), __uint_as_float(w1));
        b.nodeStart[child] = start + pre[k];
        b.nodeCount[child] = c[k];
        if (c[k] > 8) ++nActive;
    }
    if (nActive) atomicAdd(&b.counters[2], nActive);
}

// K3: move every point of an over-full cell into its child's range
__global__ void __launch_bounds__(256) k_level_scatter(OtBuild b, uint32_t levelStart, const uint32_t *__restrict__ idxIn,
        const uint32_t *__restrict__ slotNodeIn, const uint32_t *__restrict__ digitRank, uint32_t n,
        uint32_t *__restrict__ idxOut, uint32_t *__restrict__ slotNodeOut) {
    uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= n) return;
    uint32_t node = slotNodeIn[slot];
    if (node == OT_DONE) { idxOut[slot] = idxIn[slot]; slotNodeOut[slot] = OT_DONE; return; }
    uint32_t dr = digitRank[slot], digit = dr >> 29, rank = dr & 0x1fffffffu;
    float4 b1 = __ldg(&b.nodes[3 * (size_t)node + 2]);
    uint32_t first = __float_as_uint(b1.z), mask = __float_as_uint(b1.w) & 0xffu;
    uint32_t child = first + __popc(mask & ((1u << digit) - 1u));
    uint32_t dst = b.nodeStart[node] + b.cnt[(node - levelStart) * 8u + digit] + rank;
    idxOut[dst] = idxIn[slot];
    slotNodeOut[dst] = b.nodeCount[child] > 8 ? child : OT_DONE;
}

// leaf points in ascending point index (= the reference's insertion order)
__global__ void k_sort_leaves(OtBuild b, uint32_t nNodes, uint32_t *__restrict__ idx) {
    uint32_t node = blockIdx.x * blockDim.x + threadIdx.x;
    if (node >= nNodes) return;
    float4 b1 = b.nodes[3 * (size_t)node + 2];
    uint32_t w1 = __float_as_uint(b1.w);
    if (!(w1 & SKC_OT_LEAF)) return;
    uint32_t cnt = w1 & 0xffu, start = __float_as_uint(b1.z);
    uint32_t v[8];
    for (uint32_t k = 0; k < 8; ++k) v[k] = k < cnt ? idx[start + k] : 0xffffffffu;
    for (int i = 1; i < 8; ++i) {
        uint32_t x = v[i]; int j = i - 1;
        while (j >= 0 && v[j] > x) { v[j + 1] = v[j]; --j; }
        v[j + 1] = x;
    }
    for (uint32_t k = 0; k < cnt; ++k) idx[start + k] = v[k];
}

// gather the point records into leaf order: 8 threads per 128 B row (one float4 each), 4 rows per warp
// {p, area} into leaf order (16 B per point).  The 128-byte E rows stay where the irradiance pass -- or the all-gather -- put
// them: InitHierarchy and Mo read them through the same slot -> point index (a whole line per row either way), which saves a
// 6.3 GB copy per build (1.4 of 10 ms) and 3.15 GB of HBM.
__global__ void __launch_bounds__(256) k_gather_points(const uint32_t *__restrict__ idx, uint64_t n, const float4 *__restrict__ posArea,
                                float4 *__restrict__ posAreaOut) {
    uint64_t slot = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= n) return;
    posAreaOut[slot] = __ldg(&posArea[__ldg(&idx[slot])]);
}

// SampledSpectrum::y() (core/spectrum.h:387-393)
__device__ __forceinline__ float spectrumY(const float *c) {
    float yy = 0.f;
#pragma unroll
    for (int i = 0; i < SKC_NS; ++i) yy += c_cieY[i] * c[i];
    return yy * 300.f / (SKC_CIE_Y_INTEGRAL * 30.f);
}

// InitHierarchy (diffusionutil.h:133-173) for the nodes [levelStart, levelStart+levelCount)
__global__ void __launch_bounds__(128) k_init_level(float4 *__restrict__ nodes, float *__restrict__ Et, uint32_t levelStart,
        uint32_t levelCount, const float4 *__restrict__ leafPosArea, const uint32_t *__restrict__ pointIdx, const float *__restrict__ E) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= levelCount) return;
    uint32_t node = levelStart + t;
    float4 b1 = nodes[3 * (size_t)node + 2];
    uint32_t w0 = __float_as_uint(b1.z), w1 = __float_as_uint(b1.w);
    float et[32];
#pragma unroll
    for (int c = 0; c < 32; ++c) et[c] = 0.f;
    float px = 0.f, py = 0.f, pz = 0.f, sumWt = 0.f, sumArea = 0.f;
    if (w1 & SKC_OT_LEAF) {
        uint32_t cnt = w1 & 0xffu;
        for (uint32_t k = 0; k < cnt; ++k) {
            float4 pa = __ldg(&leafPosArea[w0 + k]);
            const float4 *row = reinterpret_cast<const float4 *>(E + (uint64_t)__ldg(&pointIdx[w0 + k]) * SKC_EPAD);
            float ipEt[32];
#pragma unroll
            for (int q = 0; q < 8; ++q) { float4 v = __ldg(row + q); ipEt[4 * q] = v.x * pa.w; ipEt[4 * q + 1] = v.y * pa.w; ipEt[4 * q + 2] = v.z * pa.w; ipEt[4 * q + 3] = v.w * pa.w; }
            float wt = spectrumY(ipEt);
#pragma unroll
            for (int c = 0; c < SKC_NS; ++c) et[c] += ipEt[c];
            px += wt * pa.x; py += wt * pa.y; pz += wt * pa.z;
            sumWt += wt;
            sumArea += pa.w;
        }
    } else {
        uint32_t nCh = __popc(w1 & 0xffu);
        for (uint32_t k = 0; k < nCh; ++k) {
            uint32_t child = w0 + k;
            float4 g = nodes[3 * (size_t)child];
            const float4 *row = reinterpret_cast<const float4 *>(Et + (uint64_t)child * SKC_EPAD);
            float cEt[32];
#pragma unroll
            for (int q = 0; q < 8; ++q) { float4 v = row[q]; cEt[4 * q] = v.x; cEt[4 * q + 1] = v.y; cEt[4 * q + 2] = v.z; cEt[4 * q + 3] = v.w; }
            float wt = spectrumY(cEt);
#pragma unroll
            for (int c = 0; c < SKC_NS; ++c) et[c] += cEt[c];
            px += wt * g.x; py += wt * g.y; pz += wt * g.z;
            sumWt += wt;
            sumArea += g.w;
        }
    }
    if (sumWt > 0.f) { float inv = 1.f / sumWt; px *= inv; py *= inv; pz *= inv; }
    bool black = true;
#pragma unroll
    for (int c = 0; c < SKC_NS; ++c) if (et[c] != 0.f) black = false;
    float4 *orow = reinterpret_cast<float4 *>(Et + (uint64_t)node * SKC_EPAD);
#pragma unroll
    for (int q = 0; q < 8; ++q)
        orow[q] = make_float4(et[4 * q], et[4 * q + 1], et[4 * q + 2], et[4 * q + 3]);
    nodes[3 * (size_t)node] = make_float4(px, py, pz, sumArea);
    nodes[3 * (size_t)node + 2] = make_float4(b1.x, b1.y, b1.z, __uint_as_float(black ? (w1 | SKC_OT_BLACK) : (w1 & ~SKC_OT_BLACK)));
}

// ---- host driver ----------------------------------------------------------------------------
static int ensureOctreeStorage(skc_ctx *ctx, uint64_t n, uint32_t cap) {
    if (ctx->otCap < cap) {
        if (ctx->otGeo) cudaFree(ctx->otGeo);
        if (ctx->otEt) cudaFree(ctx->otEt);
        ctx->otGeo = nullptr; ctx->otEt = nullptr; ctx->otCap = 0;
        SKC_CUDA(cudaMalloc((void **)&ctx->otGeo, (size_t)cap * 3 * sizeof(float4)));
        SKC_CUDA(cudaMalloc((void **)&ctx->otEt, (size_t)cap * SKC_EPAD * sizeof(float)));
        ctx->otCap = cap;
    }
    if (ctx->otPoints < n || !ctx->otLeafPosArea) {
        if (ctx->otLeafPosArea) cudaFree(ctx->otLeafPosArea);
        if (ctx->otPointIdx) cudaFree(ctx->otPointIdx);
        ctx->otLeafPosArea = nullptr; ctx->otPointIdx = nullptr;
        uint64_t m = n ? n : 1;
        SKC_CUDA(cudaMalloc((void **)&ctx->otLeafPosArea, m * sizeof(float4)));
        SKC_CUDA(cudaMalloc((void **)&ctx->otPointIdx, m * sizeof(uint32_t)));
    }
    ctx->otPoints = n;
    return 0;
}

// Topology: bounds, breadth-first splitting, leaf order.  Reads only the point POSITIONS, so a multi-GPU caller can run it
// while the irradiance rows (8x the bytes) are still being gathered from the other ranks.
static int topologyWithCapacity(skc_ctx *ctx, uint32_t n, uint32_t cap, uint32_t *needOut);

extern "C" int skc_build_octree_topology(skc_ctx *ctx) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    ctx->otTopologyReady = false;
    uint64_t n64 = ctx->ipCount;
    if (n64 > 0x1fffffffull) return skc_fail(ctx, "skc_build_octree: more than 2^29 irradiance points");
    uint32_t n = (uint32_t)n64;
    {
        float y[32] = {0};
        for (int i = 0; i < SKC_NS; ++i) y[i] = SKC_CIE_Y30[i];
        SKC_CUDA(cudaMemcpyToSymbolAsync(c_cieY, y, sizeof y, 0, cudaMemcpyHostToDevice, ctx->stream));
    }
    // n + n/2 nodes hold every surface-like point set; tight clusters of 9+ near-coincident points grow chains of
    // single-child interior nodes (the reference's Insert has no depth limit) and can need more: grow and retry
    uint32_t cap = std::max(ctx->otCap, n + n / 2 + 4096);
    if (const char *e = getenv("SKC_OCTREE_CAP")) cap = std::max(ctx->otCap, (uint32_t)std::max(16, atoi(e)));     // tests: force the growth path
    for (int attempt = 0; attempt < 6; ++attempt) {
        uint32_t need = 0;
        int rc = topologyWithCapacity(ctx, n, cap, &need);
        if (rc != 3) return rc;
        uint64_t grown = std::max<uint64_t>((uint64_t)need + need / 8 + 4096, (uint64_t)cap * 2);
        if (grown > 0xfffffff0ull) break;
        cap = (uint32_t)grown;
    }
    return skc_fail(ctx, "skc_build_octree: the tree needs more than %u nodes", cap);
}

// 0 = done, 1 = error, 3 = node capacity exceeded (*needOut = nodes needed where known, else 0)
static int topologyWithCapacity(skc_ctx *ctx, uint32_t n, uint32_t cap, uint32_t *needOut) {
    if (ensureOctreeStorage(ctx, n, cap)) return 1;
    skc_stage_begin(ctx);
    // --- bounds (multipolesubsurface.cpp:223-225): union of the point positions
    float hb[6] = {INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY};
    if (n) {
        const int nb = 296;
        float *dPart = (float *)skc_scratch(ctx, 8, nb * 6 * sizeof(float));
        if (!dPart) return 1;
        k_bounds_partial<<<nb, 256, 0, ctx->stream>>>(ctx->ipPosArea, n, dPart);
        SKC_LAUNCHED("k_bounds_partial");
        std::vector<float> hp(nb * 6);
        SKC_CUDA(cudaMemcpyAsync(hp.data(), dPart, nb * 6 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
        SKC_CUDA(cudaStreamSynchronize(ctx->stream));
        for (int b = 0; b < nb; ++b)
            for (int k = 0; k < 3; ++k) { hb[k] = fminf(hb[k], hp[b * 6 + k]); hb[3 + k] = fmaxf(hb[3 + k], hp[b * 6 + 3 + k]); }
    }
    for (int k = 0; k < 6; ++k) ctx->otBoundsHost[k] = hb[k];
    // --- sort-based topology (skc_octree_sort.cu); trees deeper than its 21-level path keys use the level passes below
    if (n > 8 && !getenv("SKC_OCTREE_LEVELWISE")) {
        std::vector<uint32_t> ls;
        uint32_t totalSorted = 0;
        int rc = skc_octree_topology_sorted(ctx, n, hb, cap, ls, &totalSorted);
        if (rc == 1) return 1;
        if (rc == 3) { *needOut = totalSorted; return 3; }
        if (rc == 0) {
            OtBuild bs; memset(&bs, 0, sizeof bs);
            bs.nodes = ctx->otGeo;
            k_sort_leaves<<<(totalSorted + 255) / 256, 256, 0, ctx->stream>>>(bs, totalSorted, ctx->otPointIdx);
            SKC_LAUNCHED("k_sort_leaves");
            skc_stage_end(ctx, 2);
            ctx->otNodes = totalSorted;
            ctx->otLevels = (uint32_t)ls.size() - 1;
            ctx->otLevelStart = ls;
            ctx->otTopologyReady = true;
            return 0;
        }
    }
    // --- scratch: ping-pong permutation + slot->node maps, digit/rank, per-node ranges, counters
    size_t words = (size_t)(n ? n : 1);
    uint32_t *buf = (uint32_t *)skc_scratch(ctx, 9, (5 * words + 2 * (size_t)cap + 16) * sizeof(uint32_t));
    uint32_t *cnt = (uint32_t *)skc_scratch(ctx, 10, ((size_t)cap * 8 + 8) * sizeof(uint32_t));
    if (!buf || !cnt) return 1;
    uint32_t *idxA = buf, *idxB = buf + words, *snA = buf + 2 * words, *snB = buf + 3 * words, *dr = buf + 4 * words;
    OtBuild b;
    b.nodes = ctx->otGeo; b.nodeStart = buf + 5 * words; b.nodeCount = b.nodeStart + cap; b.counters = b.nodeCount + cap;
    b.cnt = cnt; b.cap = cap;
    k_init_root<<<1, 1, 0, ctx->stream>>>(b, n, make_float4(hb[0], hb[1], hb[2], hb[3]), make_float4(hb[4], hb[5], 0.f, 0.f));
    SKC_LAUNCHED("k_init_root");
    unsigned gridN = (unsigned)((words + 255) / 256);
    k_fill_u32<<<gridN, 256, 0, ctx->stream>>>(idxA, n, 0, 1);
    SKC_LAUNCHED("k_fill_u32");
    k_fill_u32<<<gridN, 256, 0, ctx->stream>>>(snA, n, n > 8 ? 0u : OT_DONE, 0);
    SKC_LAUNCHED("k_fill_u32");
    std::vector<uint32_t> levelStart;   // node index where each level begins
    levelStart.push_back(0);
    uint32_t total = 1, active = n > 8 ? 1 : 0;
    uint32_t curStart = 0, curCount = 1;
    while (active) {
        if (levelStart.size() > OT_MAX_LEVELS)
            return skc_fail(ctx, "skc_build_octree: tree deeper than %d levels (9 or more coincident points; the reference's Insert recurses without bound here)", OT_MAX_LEVELS);
        SKC_CUDA(cudaMemsetAsync(cnt, 0, (size_t)curCount * 8 * sizeof(uint32_t), ctx->stream));
        SKC_CUDA(cudaMemsetAsync(b.counters + 2, 0, sizeof(uint32_t), ctx->stream));
        k_level_count<<<gridN, 256, 0, ctx->stream>>>(b, curStart, ctx->ipPosArea, idxA, snA, n, dr);
        SKC_LAUNCHED("k_level_count");
        k_level_split<<<(curCount + 127) / 128, 128, 0, ctx->stream>>>(b, curStart, curCount);
        SKC_LAUNCHED("k_level_split");
        k_level_scatter<<<gridN, 256, 0, ctx->stream>>>(b, curStart, idxA, snA, dr, n, idxB, snB);
        SKC_LAUNCHED("k_level_scatter");
        uint32_t hc[3];
        SKC_CUDA(cudaMemcpyAsync(hc, b.counters, sizeof hc, cudaMemcpyDeviceToHost, ctx->stream));
        SKC_CUDA(cudaStreamSynchronize(ctx->stream));
        if (hc[1]) { *needOut = 0; return 3; }
        std::swap(idxA, idxB); std::swap(snA, snB);
        curStart = total; curCount = hc[0] - total; total = hc[0]; active = hc[2];
        levelStart.push_back(curStart);
    }
    levelStart.push_back(total);   // sentinel; the last real level may be empty
    if (n) {
        k_sort_leaves<<<(total + 255) / 256, 256, 0, ctx->stream>>>(b, total, idxA);
        SKC_LAUNCHED("k_sort_leaves");
        SKC_CUDA(cudaMemcpyAsync(ctx->otPointIdx, idxA, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctx->stream));
    }
    skc_stage_end(ctx, 2);
    ctx->otNodes = total;
    ctx->otLevels = (uint32_t)levelStart.size() - 1;
    ctx->otLevelStart = levelStart;
    ctx->otTopologyReady = true;
    return 0;
}

// Finalize: gather the point records into leaf order and run InitHierarchy bottom-up (needs the irradiance rows).
extern "C" int skc_build_octree_finalize(skc_ctx *ctx, uint32_t *nNodesOut, float *root34, float *bounds6) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!ctx->otTopologyReady) return skc_fail(ctx, "skc_build_octree_finalize: no topology (skc_build_octree_topology)");
    uint32_t n = (uint32_t)ctx->otPoints, total = ctx->otNodes;
    const std::vector<uint32_t> &levelStart = ctx->otLevelStart;
    const float *hb = ctx->otBoundsHost;
    float prev = ctx->stageMs[2];
    skc_stage_begin(ctx);
    if (n) {
        k_gather_points<<<(unsigned)(((uint64_t)n + 255) / 256), 256, 0, ctx->stream>>>(ctx->otPointIdx, n, ctx->ipPosArea, ctx->otLeafPosArea);
        SKC_LAUNCHED("k_gather_points");
    }
    for (int lv = (int)levelStart.size() - 2; lv >= 0; --lv) {
        uint32_t s = levelStart[lv], c = levelStart[lv + 1] - s;
        if (!c) continue;
        k_init_level<<<(c + 127) / 128, 128, 0, ctx->stream>>>(ctx->otGeo, ctx->otEt, s, c, ctx->otLeafPosArea, ctx->otPointIdx, ctx->ipE);
        SKC_LAUNCHED("k_init_level");
    }
    skc_stage_end(ctx, 2);
    ctx->stageMs[2] += prev;          // stage 2 = topology + finalize of one build
    if (nNodesOut) *nNodesOut = total;
    if (bounds6) for (int k = 0; k < 6; ++k) bounds6[k] = hb[k];
    if (root34) {
        float4 g;
        SKC_CUDA(cudaMemcpyAsync(&g, ctx->otGeo, sizeof g, cudaMemcpyDeviceToHost, ctx->stream));
        float e[32];
        SKC_CUDA(cudaMemcpyAsync(e, ctx->otEt, sizeof e, cudaMemcpyDeviceToHost, ctx->stream));
        SKC_CUDA(cudaStreamSynchronize(ctx->stream));
        root34[0] = g.x; root34[1] = g.y; root34[2] = g.z; root34[3] = g.w;
        for (int c = 0; c < SKC_NS; ++c) root34[4 + c] = e[c];
    }
    return 0;
}

extern "C" int skc_build_octree(skc_ctx *ctx, uint32_t *nNodesOut, float *root34, float *bounds6) {
    if (skc_build_octree_topology(ctx)) return 1;
    return skc_build_octree_finalize(ctx, nNodesOut, root34, bounds6);
}

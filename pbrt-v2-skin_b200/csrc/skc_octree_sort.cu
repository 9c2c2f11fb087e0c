// skc_octree_sort.cu -- octree topology from one sort instead of one pass per level.
//
// Reference: integrators/diffusionutil.h:94-132 (Insert: a cell is split when it receives its 9th point; child =
// 4*(x > mid) + 2*(y > mid) + (z > mid)), core/octree.h:87-97 (octreeChildBound).
//
// The path of a point through the reference's tree depends only on its position and the root bounds: at every level
// the child digit comes from `p > .5f*min + .5f*max` with the bounds narrowed exactly as octreeChildBound does.  So
//   k_path_keys    writes the 21-level path (63 bits, 3 per level) of every point with that very fp32 arithmetic
//   cub radix sort orders (key, point index): every cell of every level is now a contiguous run
//   k_leaf_depth   per point: the shallowest level whose cell holds <= 8 points (window of 8 neighbours each side),
//                  i.e. its leaf; and the first level at which its path leaves the previous point's
//   k_level_counts / k_level_scan   how many nodes each 256-point block starts at each level -> breadth-first
//                  numbering: level by level, inside a level in key order = child-digit order, children contiguous
//   k_emit_nodes   writes the node records (bounds by the same descent, first child / point slot, child mask by
//                  atomicOr into the parent, leaf counts)
// which yields the node array the level-by-level builder in skc_octree.cu produces (same cells, same child order),
// in ~1/2 of its time.  Trees deeper than 21 levels (more than 8 points sharing a 63-bit path) fall back to it.
#include <vector>
#include <cub/device/device_radix_sort.cuh>
#include "skc_internal.h"

#define OTS_LEVELS 21
#define OTS_FAST_LEVELS 16      // levels the first, shorter sort covers
#define OTS_BLOCK 256

struct OtsBounds { float mnx, mny, mnz, mxx, mxy, mxz; };

// one level down: digit of p in the cell and the child's bounds (diffusionutil.h:96,116-117, octree.h:87-97)
__device__ __forceinline__ uint32_t otsDescend(OtsBounds &b, float px, float py, float pz) {
    float mx = .5f * b.mnx + .5f * b.mxx, my = .5f * b.mny + .5f * b.mxy, mz = .5f * b.mnz + .5f * b.mxz;
    uint32_t d = (px > mx ? 4u : 0u) + (py > my ? 2u : 0u) + (pz > mz ? 1u : 0u);
    if (d & 4) b.mnx = mx; else b.mxx = mx;
    if (d & 2) b.mny = my; else b.mxy = my;
    if (d & 1) b.mnz = mz; else b.mxz = mz;
    return d;
}
__device__ __forceinline__ void otsChild(OtsBounds &b, uint32_t d) {
    float mx = .5f * b.mnx + .5f * b.mxx, my = .5f * b.mny + .5f * b.mxy, mz = .5f * b.mnz + .5f * b.mxz;
    if (d & 4) b.mnx = mx; else b.mxx = mx;
    if (d & 2) b.mny = my; else b.mxy = my;
    if (d & 1) b.mnz = mz; else b.mxz = mz;
}

__global__ void __launch_bounds__(256) k_path_keys(const float4 *__restrict__ posArea, uint32_t n, OtsBounds root,
                                                   unsigned long long *__restrict__ keys, uint32_t *__restrict__ idx) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 p = __ldg(&posArea[i]);
    OtsBounds b = root;
    unsigned long long key = 0;
#pragma unroll 1
    for (int l = 0; l < OTS_LEVELS; ++l) key = (key << 3) | otsDescend(b, p.x, p.y, p.z);
    keys[i] = key;
    idx[i] = i;
}

// number of leading levels two paths share (21 when equal)
__device__ __forceinline__ int otsCommonLevels(unsigned long long a, unsigned long long b) {
    unsigned long long x = a ^ b;
    if (!x) return OTS_LEVELS;
    return (__clzll((long long)x) - 1) / 3;
}

// info[i] = depth | lo << 8 | leafCount << 16 ; depth = level of the leaf holding sorted point i, lo = first level at
// which i's path differs from i-1's (0 for i = 0), leafCount = points in that leaf (valid where lo <= depth)
// sortedLevels: the keys are sorted on their first sortedLevels levels only (see skc_octree_topology_sorted); cells of deeper
// levels are then not contiguous, so a leaf deeper than that cannot be recognised here and is reported through `overflow`.
__global__ void __launch_bounds__(256) k_leaf_depth(const unsigned long long *__restrict__ keys, uint32_t n,
                                                    uint32_t *__restrict__ info, uint32_t *__restrict__ overflow, int sortedLevels) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned long long k = keys[i];
    int cl[8], cr[8];
#pragma unroll
    for (int t = 0; t < 8; ++t) {
        cl[t] = i >= (uint32_t)(t + 1) ? otsCommonLevels(k, keys[i - t - 1]) : -1;
        cr[t] = i + t + 1 < n ? otsCommonLevels(k, keys[i + t + 1]) : -1;
    }
    // The cell of level l around point i holds 1 + L(l) + R(l) points, L / R = neighbours on either side that share >= l levels
    // (cl, cr are non-increasing in t).  It holds more than 8 iff for some a in 0..8 the a-th left and the (8-a)-th right
    // neighbour both share >= l levels, i.e. iff l <= max_a min(cl[a-1], cr[7-a]); the leaf's depth is one more than that.
    int deepestCrowded = -1;
#pragma unroll
    for (int a = 0; a <= 8; ++a) {
        const int x = a == 0 ? 1000 : cl[a - 1], y = a == 8 ? 1000 : cr[7 - a];
        deepestCrowded = max(deepestCrowded, min(x, y));
    }
    int depth = deepestCrowded + 1, count = 1;
    if (depth > sortedLevels) { atomicExch(overflow, 1u); depth = OTS_LEVELS; count = 8; }
    else {
#pragma unroll
        for (int t = 0; t < 8; ++t) count += (cl[t] >= depth) + (cr[t] >= depth);
    }
    int lo = i == 0 ? 0 : cl[0] + 1;
    info[i] = (uint32_t)depth | ((uint32_t)lo << 8) | ((uint32_t)count << 16);
}

// blockCount[l * nBlocks + block] = nodes of level l started inside the block
__global__ void __launch_bounds__(OTS_BLOCK) k_level_counts(const uint32_t *__restrict__ info, uint32_t n, uint32_t nBlocks,
                                                            uint32_t *__restrict__ blockCount) {
    __shared__ uint32_t cnt[OTS_LEVELS + 1];
    uint32_t i = blockIdx.x * OTS_BLOCK + threadIdx.x;
    int depth = -1, lo = 0;
    if (i < n) { uint32_t v = info[i]; depth = v & 0xff; lo = (v >> 8) & 0xff; }
    if (threadIdx.x <= OTS_LEVELS) cnt[threadIdx.x] = 0;
    __syncthreads();
    for (int l = 0; l <= OTS_LEVELS; ++l) {
        unsigned m = __ballot_sync(0xffffffffu, lo <= l && l <= depth);
        if ((threadIdx.x & 31) == 0 && m) atomicAdd(&cnt[l], (uint32_t)__popc(m));
    }
    __syncthreads();
    if (threadIdx.x <= OTS_LEVELS) blockCount[(size_t)threadIdx.x * nBlocks + blockIdx.x] = cnt[threadIdx.x];
}

// one CTA per level: exclusive scan of the block counts in place, total into levelCount[l]
__global__ void __launch_bounds__(1024) k_level_scan(uint32_t *__restrict__ blockCount, uint32_t nBlocks, uint32_t *__restrict__ levelCount) {
    __shared__ uint32_t part[1024];
    uint32_t *row = blockCount + (size_t)blockIdx.x * nBlocks;
    uint32_t per = (nBlocks + 1023) / 1024, a = threadIdx.x * per, b = min(a + per, nBlocks);
    uint32_t sum = 0;
    for (uint32_t k = a; k < b; ++k) sum += row[k];
    part[threadIdx.x] = sum;
    __syncthreads();
    for (int st = 1; st < 1024; st <<= 1) {
        uint32_t v = threadIdx.x >= (uint32_t)st ? part[threadIdx.x - st] : 0u;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    uint32_t run = part[threadIdx.x] - sum;
    for (uint32_t k = a; k < b; ++k) { uint32_t c = row[k]; row[k] = run; run += c; }
    if (threadIdx.x == 1023) levelCount[blockIdx.x] = part[1023];
}

struct OtsLevels { uint32_t start[OTS_LEVELS + 2]; };

__global__ void __launch_bounds__(OTS_BLOCK) k_emit_nodes(const unsigned long long *__restrict__ keys, const uint32_t *__restrict__ info,
                                                          uint32_t n, uint32_t nBlocks, const uint32_t *__restrict__ blockBase, OtsLevels lv,
                                                          OtsBounds root, float4 *__restrict__ nodes, int maxLevel) {
    // warpTot[l][w] = nodes of level l started by warp w of this block: all levels' ballots first, ONE barrier, then every thread
    // ranks itself per level from the table (the first version had two barriers per level)
    __shared__ uint32_t warpTot[OTS_LEVELS + 1][OTS_BLOCK / 32];
    uint32_t i = blockIdx.x * OTS_BLOCK + threadIdx.x;
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int depth = -1, lo = 0; uint32_t count = 0;
    unsigned long long key = 0;
    if (i < n) { uint32_t v = info[i]; depth = v & 0xff; lo = (v >> 8) & 0xff; count = v >> 16; key = keys[i]; }
    const unsigned below = (1u << lane) - 1u;
    for (int l = 0; l <= maxLevel; ++l) {
        unsigned m = __ballot_sync(0xffffffffu, lo <= l && l <= depth);
        if (lane == 0) warpTot[l][warp] = __popc(m);
    }
    __syncthreads();
    OtsBounds b = root;
    uint32_t containing = 0;      // node of the previous level that holds this point
    uint32_t prevOwn = 0xffffffffu;  // node this thread started at the previous level (if interior: awaits its first child)
    const int warpDepth = __reduce_max_sync(0xffffffffu, depth);
    for (int l = 0; l <= warpDepth; ++l) {
        bool alive = l <= depth;
        uint32_t digit = 0;
        if (l > 0 && alive) { digit = (uint32_t)(key >> (3 * (OTS_LEVELS - l))) & 7u; otsChild(b, digit); }
        bool start = alive && lo <= l;
        unsigned m = __ballot_sync(0xffffffffu, start);
        uint32_t before = 0;
        for (int w = 0; w < warp; ++w) before += warpTot[l][w];
        uint32_t excl = before + __popc(m & below);
        uint32_t base = lv.start[l] + __ldg(&blockBase[(size_t)l * nBlocks + blockIdx.x]);
        if (start) {
            uint32_t id = base + excl;
            float4 *nd = nodes + 3 * (size_t)id;
            nd[1] = make_float4(b.mnx, b.mny, b.mnz, b.mxx);
            float *w = reinterpret_cast<float *>(nd + 2);
            w[0] = b.mxy; w[1] = b.mxz;
            if (l == depth) { w[2] = __uint_as_float(i); atomicOr(reinterpret_cast<unsigned int *>(w + 3), SKC_OT_LEAF | count); }
            if (l > 0) {
                atomicOr(reinterpret_cast<unsigned int *>(nodes + 3 * (size_t)containing + 2) + 3, 1u << digit);
                if (prevOwn != 0xffffffffu) reinterpret_cast<float *>(nodes + 3 * (size_t)prevOwn + 2)[2] = __uint_as_float(id);   // first child
            }
            prevOwn = l < depth ? id : 0xffffffffu;
            containing = id;
        } else {
            if (alive) containing = base + excl - 1;   // the latest node of this level started at or before this point
            prevOwn = 0xffffffffu;
        }
    }
}

// Returns 0 and fills levelStart (first node of each level + sentinel) and total; 2 = tree deeper than OTS_LEVELS
// (caller falls back to the level-by-level builder); 3 = the tree has *totalOut > cap nodes (caller grows otGeo and retries); 1 = error.  ctx->otGeo must hold `cap` nodes, ctx->otPointIdx n ids.
int skc_octree_topology_sorted(skc_ctx *ctx, uint32_t n, const float *hb, uint32_t cap, std::vector<uint32_t> &levelStart, uint32_t *totalOut) {
    uint32_t nBlocks = (n + OTS_BLOCK - 1) / OTS_BLOCK;
    size_t tempBytes = 0;
    cub::DoubleBuffer<unsigned long long> dk(nullptr, nullptr);
    cub::DoubleBuffer<uint32_t> dv(nullptr, nullptr);
    cub::DeviceRadixSort::SortPairs(nullptr, tempBytes, dk, dv, (int)n, 0, 3 * OTS_LEVELS, ctx->stream);
    size_t keyBytes = ((size_t)n * 8 + 255) / 256 * 256, idxBytes = ((size_t)n * 4 + 255) / 256 * 256;
    size_t cntBytes = ((size_t)(OTS_LEVELS + 1) * nBlocks * 4 + 255) / 256 * 256;
    char *buf = (char *)skc_scratch(ctx, 9, 2 * keyBytes + 2 * idxBytes + idxBytes + cntBytes + tempBytes + 1024);
    if (!buf) return 1;
    unsigned long long *k0 = (unsigned long long *)buf, *k1 = (unsigned long long *)(buf + keyBytes);
    uint32_t *v0 = (uint32_t *)(buf + 2 * keyBytes), *v1 = (uint32_t *)(buf + 2 * keyBytes + idxBytes);
    uint32_t *info = (uint32_t *)(buf + 2 * keyBytes + 2 * idxBytes);
    uint32_t *blockCount = (uint32_t *)(buf + 2 * keyBytes + 3 * idxBytes);
    void *temp = buf + 2 * keyBytes + 3 * idxBytes + cntBytes;
    uint32_t *small = (uint32_t *)((char *)temp + tempBytes);      // [0] overflow flag, [1..] level counts
    small = (uint32_t *)(((uintptr_t)small + 255) / 256 * 256);
    OtsBounds root = {hb[0], hb[1], hb[2], hb[3], hb[4], hb[5]};
    unsigned grid = (n + 255) / 256;
    SKC_CUDA(cudaMemsetAsync(small, 0, (OTS_LEVELS + 4) * sizeof(uint32_t), ctx->stream));
    // First try: sort on the first OTS_FAST_LEVELS levels only (48 of the 63 key bits: 6 radix passes instead of 8).  Every
    // cell down to that level is a contiguous run, which is all k_leaf_depth needs when no leaf lies deeper -- the usual case
    // (skin.scene: 15 levels); a deeper tree raises the overflow flag and the keys are sorted again in full.
    const unsigned long long *keys = nullptr;
    const uint32_t *idx = nullptr;
    uint32_t h[OTS_LEVELS + 2];
    for (int attempt = 0; attempt < 2; ++attempt) {
        const int sortedLevels = attempt == 0 ? OTS_FAST_LEVELS : OTS_LEVELS;
        SKC_CUDA(cudaMemsetAsync(small, 0, (OTS_LEVELS + 4) * sizeof(uint32_t), ctx->stream));
        k_path_keys<<<grid, 256, 0, ctx->stream>>>(ctx->ipPosArea, n, root, k0, v0);
        SKC_LAUNCHED("k_path_keys");
        dk = cub::DoubleBuffer<unsigned long long>(k0, k1);
        dv = cub::DoubleBuffer<uint32_t>(v0, v1);
        SKC_CUDA(cub::DeviceRadixSort::SortPairs(temp, tempBytes, dk, dv, (int)n, 3 * (OTS_LEVELS - sortedLevels), 3 * OTS_LEVELS, ctx->stream));
        ctx->launches += 8;
        keys = dk.Current();
        idx = dv.Current();
        k_leaf_depth<<<grid, 256, 0, ctx->stream>>>(keys, n, info, small, sortedLevels);
        SKC_LAUNCHED("k_leaf_depth");
        k_level_counts<<<nBlocks, OTS_BLOCK, 0, ctx->stream>>>(info, n, nBlocks, blockCount);
        SKC_LAUNCHED("k_level_counts");
        k_level_scan<<<OTS_LEVELS + 1, 1024, 0, ctx->stream>>>(blockCount, nBlocks, small + 1);
        SKC_LAUNCHED("k_level_scan");
        SKC_CUDA(cudaMemcpyAsync(h, small, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
        SKC_CUDA(cudaStreamSynchronize(ctx->stream));
        if (!h[0]) break;
        if (attempt == 1) return 2;
    }
    OtsLevels lv;
    levelStart.clear();
    uint32_t total = 0;
    for (int l = 0; l <= OTS_LEVELS; ++l) {
        lv.start[l] = total;
        if (h[1 + l] || l == 0) levelStart.push_back(total);
        total += h[1 + l];
    }
    lv.start[OTS_LEVELS + 1] = total;
    levelStart.push_back(total);
    if (total > cap) { *totalOut = total; return 3; }      // the exact node count is known before anything is emitted: the caller grows the storage and retries
    SKC_CUDA(cudaMemsetAsync(ctx->otGeo, 0, (size_t)total * 3 * sizeof(float4), ctx->stream));
    k_emit_nodes<<<nBlocks, OTS_BLOCK, 0, ctx->stream>>>(keys, info, n, nBlocks, blockCount, lv, root, ctx->otGeo, (int)levelStart.size() - 2);
    SKC_LAUNCHED("k_emit_nodes");
    SKC_CUDA(cudaMemcpyAsync(ctx->otPointIdx, idx, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctx->stream));
    *totalOut = total;
    return 0;
}

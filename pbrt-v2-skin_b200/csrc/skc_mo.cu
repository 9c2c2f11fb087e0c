// skc_mo.cu -- SubsurfaceOctreeNode::Mo and the subsurface term of Li on the device.
//
// Reference: integrators/diffusionutil.h:175-210 (Mo), core/geometry.h:449-453 (Inside,
// inclusive), core/multipole.cpp:60-83,103-113 (sampleProfile: Rd table lerp, one table and
// one rcpDsqSpacing per wavelength), :458-463 (rho_hd), integrators/multipolesubsurface.cpp
// :253-304 (Li).
//
// One warp per shading point.  Traversal is warp-cooperative: up to 32 pending entries
// (octree nodes, or single irradiance points of an opened leaf) are popped from a per-warp
// shared-memory stack and classified one per lane with the reference's own tests
//     black -> drop;  dw = sumArea / d2;  (dw < maxError && !Inside) -> cluster contribution
//     else leaf -> its points;  else -> its children
// and the ballots of the classification drive (a) a shuffle-compacted push of the children
// and (b) the spectral accumulation, where the warp turns round: lane = wavelength, so a
// contribution costs one coalesced 120 B row (Et or E) and one 8 B table pair per lane.
//
// Exact-zero culling (not in the reference, results identical): Rd is 0 for every
// wavelength once d2 * min(rcpDsqSpacing) >= len-1, so contributions beyond that range are
// skipped, and so are subtrees whose box lies entirely beyond it.
#include <algorithm>
#include <vector>
#include <cub/cub.cuh>
#include "skc_internal.h"
#include "skc_mo_table.h"
#include "skc_texture.cuh"

// Launch shape: a CTA of MO_WARPS warps works through a CONTIGUOUS run of MO_WARPS*MO_HITS_PER_WARP
// shading points (neighbouring camera samples walk the same nodes, rows and table sectors, so
// they must share an SM's L1 -- with hits dealt round-robin over the grid the L1 hit rate was
// 2.7 %).  Stack bound: below MO_HIGH entries the warp pops up to 32 entries per round
// (<= MO_HIGH-1 + 7*32 afterwards), above it one entry (depth-first, +7 per tree level).
// compact profile table (skc_mo_table.h): abscissae, cells per binade, shared-memory floats for the table heads
#define MOC_K 256
#define MOC_M 16
#define MOC_HEAD_CAP 0
#define MO_WARPS 8
#define MO_HITS_PER_WARP 16
#define MO_HIGH 64
#define MO_POINT 0x80000000u

struct MoTables {
    const float2 *pairs;     // [30][len] {data[i], data[i+1]}
    const float *rcp;        // [32] rcpDsqSpacing per wavelength
    uint32_t len;
    float rcpMin;            // smallest rcpDsqSpacing (longest table in d2)
    float lenM1;             // (float)(len-1)
};

struct MoTree {
    const float4 *nodes;     // 3 per node
    const float *Et;         // [nodes][32]
    const float4 *leafPosArea;   // {p, area} in leaf order
    const uint32_t *pointIdx;    // leaf slot -> point: the E rows are read in place, in point order
    const float *E;              // [points][32]
};
__device__ __forceinline__ const float *leafRow(const MoTree &tr, uint32_t slot) { return tr.E + (size_t)__ldg(&tr.pointIdx[slot]) * SKC_EPAD; }

// sampleProfile (core/multipole.cpp:60-73).  d2*rcp is an fp32 product there too (the
// promotion to double happens after the multiply); the compare against size-1 and the
// fractional part are exact in fp32 for len < 2^24.
__device__ __forceinline__ float sampleRd(const MoTables &T, int lambda, float rcp, float d2) {
    float f = d2 * rcp;
    if (f >= T.lenM1) return 0.f;
    uint32_t seg = (uint32_t)f;
    float t = f - (float)seg;
    float2 v = __ldg(&T.pairs[(size_t)lambda * T.len + seg]);
    return (1.f - t) * v.x + t * v.y;
}

template <bool COUNT, int MO_STACK>
__device__ __forceinline__ void moOne(const MoTree &tr, const MoTables &T, float px, float py, float pz,
                                      float maxError, uint32_t *stack, int lane, float &acc, uint32_t *wc) {
    const float rcp = __ldg(&T.rcp[lane]);
    const int lam = lane < SKC_NS ? lane : 0;
    int sp = 0;
    if (lane == 0) stack[0] = 0u;
    sp = 1;
    __syncwarp();
    while (sp > 0) {
        int take = sp >= MO_HIGH ? 1 : min(sp, 32);
        sp -= take;
        uint32_t entry = lane < take ? stack[sp + lane] : 0u;
        __syncwarp();
        // ---- classify: one entry per lane
        bool contrib = false;      // adds Rd(d2) * row * scale
        uint32_t rowIdx = 0; float d2 = 0.f, scale = 1.f;
        uint32_t pushFirst = 0, pushCount = 0;   // entries this lane pushes
        bool pushPoints = false;
        if (lane < take) {
            if (entry & MO_POINT) {
                uint32_t slot = entry & ~MO_POINT;
                float4 pa = __ldg(&tr.leafPosArea[slot]);
                float dx = px - pa.x, dy = py - pa.y, dz = pz - pa.z;
                d2 = dx * dx + dy * dy + dz * dz;
                if (d2 * T.rcpMin < T.lenM1) { contrib = true; rowIdx = slot | MO_POINT; scale = pa.w; }
            } else {
                const float4 *nd = tr.nodes + 3 * (size_t)entry;
                float4 g = __ldg(nd), b0 = __ldg(nd + 1), b1 = __ldg(nd + 2);
                uint32_t w0 = __float_as_uint(b1.z), w1 = __float_as_uint(b1.w);
                if (!(w1 & SKC_OT_BLACK)) {
                    float dx = px - g.x, dy = py - g.y, dz = pz - g.z;
                    d2 = dx * dx + dy * dy + dz * dz;
                    float dw = g.w / d2;
                    bool inside = px >= b0.x && px <= b0.w && py >= b0.y && py <= b1.x && pz >= b0.z && pz <= b1.y;
                    if (dw < maxError && !inside) {
                        if (d2 * T.rcpMin < T.lenM1) { contrib = true; rowIdx = entry; scale = 1.f; }
                    } else {
                        // distance from pt to the cell: beyond the table range nothing below contributes
                        float ex = fmaxf(fmaxf(b0.x - px, px - b0.w), 0.f);
                        float ey = fmaxf(fmaxf(b0.y - py, py - b1.x), 0.f);
                        float ez = fmaxf(fmaxf(b0.z - pz, pz - b1.y), 0.f);
                        float db2 = ex * ex + ey * ey + ez * ez;
                        if ((db2 * 0.999f) * T.rcpMin < T.lenM1) {
                            if (w1 & SKC_OT_LEAF) { pushFirst = w0; pushCount = w1 & 0xffu; pushPoints = true; }
                            else { pushFirst = w0; pushCount = __popc(w1 & 0xffu); }
                        }
                    }
                }
            }
        }
        // ---- push children / leaf points (exclusive scan of pushCount over lanes)
        uint32_t incl = pushCount;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { uint32_t v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
        uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
        uint32_t off = sp + incl - pushCount;
        for (uint32_t k = 0; k < pushCount; ++k) stack[off + k] = (pushFirst + k) | (pushPoints ? MO_POINT : 0u);
        sp += (int)total;
        // ---- spectral accumulation: lane = wavelength
        unsigned cm = __ballot_sync(0xffffffffu, contrib);
        if (COUNT) {
            unsigned pm = __ballot_sync(0xffffffffu, lane < take && (entry & MO_POINT));
            unsigned cp = __ballot_sync(0xffffffffu, contrib && (rowIdx & MO_POINT));
            wc[0] += take - __popc(pm); wc[1] += __popc(pm); wc[2] += __popc(cm) - __popc(cp); wc[3] += __popc(cp);
        }
        while (cm) {
            int src = __ffs(cm) - 1; cm &= cm - 1;
            uint32_t r = __shfl_sync(0xffffffffu, rowIdx, src);
            float dd = __shfl_sync(0xffffffffu, d2, src);
            float sc = __shfl_sync(0xffffffffu, scale, src);
            const float *row = (r & MO_POINT) ? leafRow(tr, r & ~MO_POINT) : tr.Et + (size_t)r * SKC_EPAD;
            float e = __ldg(&row[lane]);
            float rd = sampleRd(T, lam, rcp, dd);
            acc += (rd * e) * sc;          // diffusionutil.h:186 / :197
        }
        __syncwarp();
    }
}

template <bool COUNT, int MO_STACK>
__global__ void __launch_bounds__(MO_WARPS * 32) k_mo(MoTree tr, MoTables T, const float *__restrict__ p3, uint64_t strideFloats,
                                                     const int32_t *__restrict__ enable, uint64_t enableStrideInts,
                                                     uint64_t n, float maxError, float *__restrict__ Mo, unsigned long long *cnt) {
    __shared__ uint32_t stacks[MO_WARPS][MO_STACK];
    int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t wc[4] = {0, 0, 0, 0};
    uint64_t base = (uint64_t)blockIdx.x * (MO_WARPS * MO_HITS_PER_WARP);
    for (int it = 0; it < MO_HITS_PER_WARP; ++it) {
        uint64_t h = base + (uint64_t)it * MO_WARPS + warp;
        if (h >= n) break;
        float acc = 0.f;
        bool on = enable ? (__ldg(&enable[h * enableStrideInts]) != 0) : true;
        if (on) {
            const float *pp = p3 + h * strideFloats;
            float px = __ldg(pp), py = __ldg(pp + 1), pz = __ldg(pp + 2);
            moOne<COUNT, MO_STACK>(tr, T, px, py, pz, maxError, stacks[warp], lane, acc, wc);
            if (COUNT) wc[0] += 0;
        }
        Mo[h * SKC_EPAD + lane] = lane < SKC_NS ? acc : 0.f;
    }
    if (COUNT && lane == 0 && cnt) {
        // [0] nodes classified, [1] leaf points tested, [2] cluster contributions, [3] point contributions
        for (int k = 0; k < 4; ++k) atomicAdd(&cnt[k], (unsigned long long)wc[k]);
    }
}

// ---------------------------------------------------------------------------------------------
// Packet variant: one warp walks the octree ONCE for 32 neighbouring shading points (lane = hit).
// A stack entry is (node, mask of lanes that still have to evaluate it); every lane applies the reference's
// tests to the node with its own point, accepting lanes add the cluster, the others descend.  Neighbouring
// camera samples accept the same far clusters at almost the same d2, so the warp's 32 table gathers of one
// wavelength fall into one or two 128 B lines and the Et / E row is one broadcast load -- the warp-per-hit
// kernel above spends 30 separate lines per contribution (one per wavelength) and is bound by L1 wavefronts.
// Each lane keeps its 30 wavelength sums in registers; the result is transposed through shared memory.
#define MOP_WARPS 4
#ifndef MOP_CTAS4
#define MOP_CTAS4 8          // resident CTAs per SM of the four-warps-per-packet instance (12 at 40 registers spills: 19.1 vs 13.6 ms on a 1/8 frame)
#endif
#ifndef MOP_CTAS1
#define MOP_CTAS1 7          // resident CTAs per SM of the one-warp-per-packet instance (72 registers)
#endif
#define MOP_STACK 512          // 1 + 7 * levels entries at most (levels <= 64)

// Optional fused Li epilogue of the packet kernel (multipolesubsurface.cpp:279-287): the subsurface radiance of each
// hit is written straight from the Mo sums -- into device memory, or into mapped pinned host memory, in which case the
// 120 B/sample result crosses PCIe while the kernel is still running instead of in a copy after it.
struct LiEpilogue { const skc_hit *hits; const float *rhoHd; const float *albedo; float mix; float *L; const float *albedoPow = nullptr; const float4 *hitAlbedo = nullptr; const float *refl = nullptr; };   // albedoPow: Pow(albedo, 1 - mix) of a constant albedo, made on the host with libm's powf

// NQ float4 groups of the 32-float row starting at group q0: 8 = the whole spectrum, 4 = half of it (a packet shared
// by two warps, see k_mo_packet)
template <bool COUNT, int NQ>
__device__ __forceinline__ void mopContribute(const MoTables &T, const float *sRcp, const float4 *__restrict__ row4, int q0, float d2, float scale,
                                              bool active, float (&acc)[4 * NQ], uint32_t &gathers) {
#pragma unroll
    for (int qq = 0; qq < NQ; ++qq) {
        float4 e4 = __ldg(row4 + q0 + qq);
        float e[4] = {e4.x, e4.y, e4.z, e4.w};
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int w = 4 * (NQ == 8 ? qq : q0 + qq) + r;
            if (w < SKC_NS) {
                float f = d2 * sRcp[w];
                if (active && f < T.lenM1) {
                    if (COUNT) ++gathers;
                    uint32_t seg = (uint32_t)f;
                    float t = f - (float)seg;
                    float2 v = __ldg(&T.pairs[(size_t)w * T.len + seg]);
                    float rd = (1.f - t) * v.x + t * v.y;
                    acc[4 * qq + r] += (rd * e[r]) * scale;
                }
            }
        }
    }
}

// Li subsurface term of one (hit, wavelength) (multipolesubsurface.cpp:279-287)
__device__ __forceinline__ float liSssValue(const LiEpilogue &li, uint64_t hh, int wl, float mo) {
    const float *hp = reinterpret_cast<const float *>(li.hits + hh);
    if (!__float_as_int(__ldg(hp + 13)) || wl >= SKC_NS) return 0.f;
    f3 nn = mk3(__ldg(hp + 7), __ldg(hp + 8), __ldg(hp + 9)), wo = mk3(__ldg(hp + 10), __ldg(hp + 11), __ldg(hp + 12));
    float costheta = fminf(fabsf(dot3(wo, nn)), 1.f);
    float fid = costheta * (float)(SKC_RHO_LEN - 1);
    int id = min(max((int)fid, 0), SKC_RHO_LEN - 2);
    float lerp = fid - (float)id;
    float rho = (1.f - lerp) * __ldg(&li.rhoHd[id * SKC_EPAD + wl]) + lerp * __ldg(&li.rhoHd[(id + 1) * SKC_EPAD + wl]);
    float Ft = 1.f - rho;
    float a = 1.f;
    if (li.albedo) a = (float)pow((double)__ldg(&li.albedo[hh * SKC_NS + wl]), (double)(1.f - li.mix));
    else if (li.hitAlbedo) { const float4 c = __ldg(&li.hitAlbedo[hh]); a = skc_powf(texReflectance(li.refl, c.x, c.y, c.z, wl), 1.f - li.mix); }      // "texture albedo" (k_hit_textures)
    else if (li.albedoPow) a = __ldg(&li.albedoPow[wl]);
    return fmaxf(((0.31830988618379067154f * Ft) * mo) * a, 0.f);       // Spectrum::Clamp(0)
}

// SPLIT = 1: one warp per packet sums all 30 wavelengths.  SPLIT = 2: two warps share a packet, each walks the octree
// for the same 32 hit points (identical decisions, no communication) and sums one half of the spectrum (wavelengths 0-15 /
// 16-29).  The walk's tests are a few per cent of the work and get done twice; in return a packet finishes in half the
// time -- the heaviest packets of a frame otherwise run 15-25 ms (each warp owns 1/28 of its SM's gather pipe) and
// set the length of every launch that has few packets per warp (multi-GPU shards) -- and 16 accumulators instead of
// 30 leave room for more resident warps.
template <bool COUNT, int SPLIT, bool PERM>
__global__ void __launch_bounds__(MOP_WARPS * 32, SPLIT == 1 ? MOP_CTAS1 : (SPLIT == 2 ? 8 : MOP_CTAS4)) k_mo_packet(MoTree tr, MoTables T, const float *__restrict__ p3, uint64_t strideFloats,
                                                             const int32_t *__restrict__ enable, uint64_t enableStrideInts,
                                                             uint64_t n, float maxError, float *__restrict__ Mo, unsigned long long *cnt, LiEpilogue li,
                                                             uint32_t *__restrict__ nextPacket, const uint32_t *__restrict__ perm) {
    __shared__ uint2 stacks[MOP_WARPS][MOP_STACK];
    __shared__ float sRcp[32];
    int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x < 32) sRcp[threadIdx.x] = __ldg(&T.rcp[threadIdx.x]);
    __syncthreads();
    uint2 *stack = stacks[warp];
    uint32_t wc[4] = {0, 0, 0, 0};
    uint32_t gathers = 0;          // this lane's table reads (COUNT only)
    const uint32_t nPackets = (uint32_t)((n + 31) / 32), nTickets = nPackets * SPLIT;
    constexpr int NQ = 8 / SPLIT;
    // persistent warps: packets are handed out in order from a global counter, so the per-packet cost (none for sky,
    // ~0.1 ms for a packet on the skin) balances across the grid and no wave of CTAs leaves a tail
    for (;;) {
    uint32_t packet = 0;
    if (lane == 0) packet = atomicAdd(nextPacket, 1u);
    packet = __shfl_sync(0xffffffffu, packet, 0);
    if (packet >= nTickets) break;
    const int part = SPLIT == 1 ? 0 : (int)(packet % SPLIT), q0 = part * NQ;
    packet /= SPLIT;
    // perm (optional): packet position -> index of the hit in the caller's arrays (hits sorted along a space-filling curve, moLaunch)
    uint64_t h0 = (uint64_t)packet * 32, h = h0 + lane;
    const uint32_t src = (PERM && h < n) ? __ldg(&perm[h]) : (uint32_t)h;      // n < 2^32 (checked at launch)
    bool on = h < n && (enable ? (__ldg(&enable[(uint64_t)src * enableStrideInts]) != 0) : true);
    float px = 0.f, py = 0.f, pz = 0.f;
    if (on) { const float *pp = p3 + (uint64_t)src * strideFloats; px = __ldg(pp); py = __ldg(pp + 1); pz = __ldg(pp + 2); }
    float acc[4 * NQ];
#pragma unroll
    for (int w = 0; w < 4 * NQ; ++w) acc[w] = 0.f;
    unsigned all = __ballot_sync(0xffffffffu, on);
    int sp = 0;
    if (all) { if (lane == 0) stack[0] = make_uint2(0u, all); sp = 1; }
    __syncwarp();
    while (sp > 0) {
        --sp;
        uint2 ent = stack[sp];
        __syncwarp();
        bool mine = (ent.y >> lane) & 1u;
        const float4 *nd = tr.nodes + 3 * (size_t)ent.x;
        float4 g = __ldg(nd), b0 = __ldg(nd + 1), b1 = __ldg(nd + 2);
        uint32_t w0 = __float_as_uint(b1.z), w1 = __float_as_uint(b1.w);
        if (COUNT && part == 0) wc[0] += __popc(ent.y);
        if (w1 & SKC_OT_BLACK) continue;
        float dx = px - g.x, dy = py - g.y, dz = pz - g.z;
        float d2 = dx * dx + dy * dy + dz * dz;
        float dw = g.w / d2;
        bool inside = px >= b0.x && px <= b0.w && py >= b0.y && py <= b1.x && pz >= b0.z && pz <= b1.y;
        bool accept = mine && dw < maxError && !inside;
        bool descend = mine && !accept;
        bool addC = accept && d2 * T.rcpMin < T.lenM1;
        if (descend) {      // beyond the table range nothing below contributes
            float ex = fmaxf(fmaxf(b0.x - px, px - b0.w), 0.f);
            float ey = fmaxf(fmaxf(b0.y - py, py - b1.x), 0.f);
            float ez = fmaxf(fmaxf(b0.z - pz, pz - b1.y), 0.f);
            float db2 = ex * ex + ey * ey + ez * ez;
            descend = (db2 * 0.999f) * T.rcpMin < T.lenM1;
        }
        unsigned am = __ballot_sync(0xffffffffu, addC), dm = __ballot_sync(0xffffffffu, descend);
        if (am) {
            if (COUNT && part == 0) wc[2] += __popc(am);
            mopContribute<COUNT, NQ>(T, sRcp, reinterpret_cast<const float4 *>(tr.Et + (size_t)ent.x * SKC_EPAD), q0, d2, 1.f, addC, acc, gathers);
        }
        if (dm) {
            if (w1 & SKC_OT_LEAF) {
                uint32_t cntP = w1 & 0xffu;
                if (COUNT && part == 0) wc[1] += cntP * __popc(dm);
                for (uint32_t k = 0; k < cntP; ++k) {
                    uint32_t slot = w0 + k;
                    float4 pa = __ldg(&tr.leafPosArea[slot]);
                    float qx = px - pa.x, qy = py - pa.y, qz = pz - pa.z;
                    float q2 = qx * qx + qy * qy + qz * qz;
                    bool c = descend && q2 * T.rcpMin < T.lenM1;
                    unsigned pm = __ballot_sync(0xffffffffu, c);
                    if (pm) {
                        if (COUNT && part == 0) wc[3] += __popc(pm);
                        mopContribute<COUNT, NQ>(T, sRcp, reinterpret_cast<const float4 *>(leafRow(tr, slot)), q0, q2, pa.w, c, acc, gathers);
                    }
                }
            } else {
                int nc = __popc(w1 & 0xffu);
                if (lane < nc) stack[sp + lane] = make_uint2(w0 + lane, dm);
                sp += nc;
                __syncwarp();
            }
        }
    }
    // transpose through the (now idle) stack memory: lane = hit -> lane = wavelength rows
    float *tile = reinterpret_cast<float *>(stack);           // 32 x 33 floats = 4224 B > 4096: use two halves
    __syncwarp();
    if constexpr (SPLIT == 1) {
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        // hits [16*half, 16*half+16): lanes of that half write their 30 sums, then all lanes stream rows out
        if ((lane >> 4) == half) {
#pragma unroll
            for (int w = 0; w < SKC_NS; ++w) tile[(lane & 15) * 33 + w] = acc[w];
            tile[(lane & 15) * 33 + 30] = 0.f; tile[(lane & 15) * 33 + 31] = 0.f;
        }
        __syncwarp();
        for (int r = 0; r < 16; ++r) {
            uint64_t hh = h0 + 16 * half + r;
            if (hh >= n) break;
            if (PERM) hh = (uint64_t)__shfl_sync(0xffffffffu, src, 16 * half + r);
            float mo = tile[r * 33 + lane];
            if (Mo) Mo[hh * SKC_EPAD + lane] = mo;
            if (li.L) {
                float out = liSssValue(li, hh, lane, mo);
                if (lane < SKC_NS) li.L[hh * SKC_NS + lane] = out;
            }
        }
        __syncwarp();
    }
    } else {
        // this warp's 4*NQ wavelengths of all 32 hits: 32 x (4*NQ+1) floats, then 32/(4*NQ) hits per pass
#pragma unroll
        for (int w = 0; w < 4 * NQ; ++w) tile[lane * (4 * NQ + 1) + w] = acc[w];
        __syncwarp();
        constexpr int NW = 4 * NQ;                     // wavelengths of this warp; 32 / NW hits per pass
        const int wl = 4 * q0 + (lane % NW);
        for (int r = 0; r < 32; r += 32 / NW) {
            int rr = r + lane / NW;
            uint64_t hh = PERM ? (uint64_t)__shfl_sync(0xffffffffu, src, rr) : h0 + rr;
            if (h0 + rr < n) {
                float mo = tile[rr * (NW + 1) + (lane % NW)];
                if (Mo) Mo[hh * SKC_EPAD + wl] = mo;
                if (li.L) {
                    float out = liSssValue(li, hh, wl, mo);
                    if (wl < SKC_NS) li.L[hh * SKC_NS + wl] = out;
                }
            }
        }
        __syncwarp();
    }
    }   // packet loop
    if (COUNT && cnt) {
        if (lane == 0) for (int k = 0; k < 4; ++k) atomicAdd(&cnt[k], (unsigned long long)wc[k]);
        atomicAdd(&cnt[4], (unsigned long long)gathers);
    }
}


// ---------------------------------------------------------------------------------------------
// k_mo_lambda: the packet walk of k_mo_packet (lane = hit for the node tests), but the contributions are summed with
// lane = WAVELENGTH and the profile values come from shared memory (skc_mo_table.h): every accepting lane leaves a
// record {d2, scale, abscissa cell, the four Lagrange weights} in shared memory, then the warp goes through the
// accepting hits one by one -- the Et / E row is one coalesced 128 B load for the whole loop, a profile value is four
// conflict-free LDS (row[k..k+3][lane]) and three FMAs, and the hit's sum is a read-modify-write of its own row of a
// per-warp 32 x 32 accumulator tile in shared memory.  Outside a wavelength's verified window the lane does the
// reference's exact lookup instead: from the table head held in shared memory, or from the global table (the far,
// noisy part of the tables: ~1 in 8 lookups).  No transposition at the end: lane = wavelength is the output layout.
#define MOL_STACK 256          // (node, mask) entries: 1 + 7 * levels, levels <= 36
#define MOL_WARP_BYTES (32 * 32 * 4 + 32 * 32 + MOL_STACK * 8)
#define MOL_MAX_WARPS 28

struct MoCompactDev {
    const float *rows, *heads;
    const float4 *lam;            // {rcp, fLo, fHi, headLen}
    const uint32_t *headOff;
    uint32_t rowFloats, headFloats;   // headFloats padded to a multiple of 4
    MoCompactParams P;
};

// G records at once: the records first, then every lane's global-table reads of the group (predicated, independent -- the
// L2 round trip of the exact tail lookups is what a one-at-a-time loop waits for), then the shared-memory work, then the sums.
template <bool COUNT, bool HEADS, int G>
__device__ __forceinline__ void molGroup(const float4 *rec, float e, float *acc, const float *sRows, const float *sHeads,
                                         const float2 *myPairs, int lane, const float4 &lp, uint32_t hoff, float lenM1L, uint32_t (&gc)[3]) {
    float4 a[G], w[G];    // {d2, scale, cell row offset, accumulator row offset} and the four Lagrange weights: two 16-byte broadcast reads per record
#pragma unroll
    for (int g = 0; g < G; ++g) { a[g] = rec[2 * g]; w[g] = rec[2 * g + 1]; }
    float f[G]; bool inC[G], ex[G], gl[G]; float2 v[G];
#pragma unroll
    for (int g = 0; g < G; ++g) {
        f[g] = a[g].x * lp.x;
        inC[g] = f[g] >= lp.y && f[g] < lp.z;                       // the verified window of this wavelength (fHi <= len-2)
        ex[g] = !inC[g] && f[g] < lenM1L;
        gl[g] = HEADS ? (ex[g] && f[g] >= lp.w) : ex[g];
        v[g] = make_float2(0.f, 0.f);
        if (gl[g]) v[g] = __ldg(myPairs + (uint32_t)f[g]);
    }
    if (HEADS)
#pragma unroll
    for (int g = 0; g < G; ++g)
        if (ex[g] && !gl[g]) { const uint32_t seg = (uint32_t)f[g]; v[g].x = sHeads[hoff + seg]; v[g].y = sHeads[hoff + seg + 1]; }
#pragma unroll
    for (int g = 0; g < G; ++g) {
        const float *r = sRows + __float_as_int(a[g].z) + lane;
        const float ww[4] = {w[g].x, w[g].y, w[g].z, w[g].w};
        const float rdC = mocCubic(ww, r[0], r[32], r[64], r[96]);
        const float t = f[g] - (float)(uint32_t)f[g];
        const float rdE = (1.f - t) * v[g].x + t * v[g].y;          // sampleProfile's own arithmetic (core/multipole.cpp:60-73)
        const float rd = inC[g] ? rdC : rdE;                        // rdE is 0 when neither path applies
        if (COUNT) { gc[0] += (inC[g] || ex[g]); gc[1] += (ex[g] && !gl[g]); gc[2] += gl[g]; }
        float *ap = acc + __float_as_int(a[g].w) + lane;
        *ap += (rd * e) * a[g].y;
    }
}

// an accepting lane's record: everything about (hit, cluster) that does not depend on the wavelength, computed once by the hit's lane
__device__ __forceinline__ void molRecord(float4 *rec, int slot, const MoCompactParams &P, int kMax32, float d2, float scale, int lane) {
    int k; float t;
    mocCell(P, d2, k, t);
    float w[4];
    mocWeights(t, w);
    rec[2 * slot] = make_float4(d2, scale, __int_as_float(min(k * 32, kMax32)), __int_as_float(lane * 32));
    rec[2 * slot + 1] = make_float4(w[0], w[1], w[2], w[3]);
}

// sums `e * Rd(d2) * scale` into the accumulator rows of the cnt hits whose records lie densely in rec[0 .. cnt)
template <bool COUNT, bool HEADS>
__device__ __forceinline__ void molContribute(int cnt, float e, const float4 *rec, float *acc, const float *sRows, const float *sHeads,
                                              const MoTables &T, int lane, const float4 &lp, uint32_t hoff, float lenM1L, uint32_t (&gc)[3]) {
    const float2 *myPairs = T.pairs + (size_t)lane * T.len;
    int i = 0;
    for (; i + 4 <= cnt; i += 4) molGroup<COUNT, HEADS, 4>(rec + 2 * i, e, acc, sRows, sHeads, myPairs, lane, lp, hoff, lenM1L, gc);
    if (cnt & 2) { molGroup<COUNT, HEADS, 2>(rec + 2 * i, e, acc, sRows, sHeads, myPairs, lane, lp, hoff, lenM1L, gc); i += 2; }
    if (cnt & 1) molGroup<COUNT, HEADS, 1>(rec + 2 * i, e, acc, sRows, sHeads, myPairs, lane, lp, hoff, lenM1L, gc);
}

template <bool COUNT, bool PERM, bool HEADS, int MAXW>
__global__ void __launch_bounds__(MAXW * 32, 1) k_mo_lambda(MoTree tr, MoTables T, MoCompactDev C, const float *__restrict__ p3, uint64_t strideFloats,
                                                                     const int32_t *__restrict__ enable, uint64_t enableStrideInts,
                                                                     uint64_t n, float maxError, float *__restrict__ Mo, unsigned long long *cnt, LiEpilogue li,
                                                                     uint32_t *__restrict__ nextPacket, const uint32_t *__restrict__ perm, uint32_t packetHits) {
    extern __shared__ float4 molSmem[];
    float *sRows = reinterpret_cast<float *>(molSmem);
    float *sHeads = sRows + C.rowFloats;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    char *wb = reinterpret_cast<char *>(sHeads + C.headFloats) + (size_t)warp * MOL_WARP_BYTES;
    float *acc = reinterpret_cast<float *>(wb);                          // [32 hits][32 wavelengths]
    float4 *rec = reinterpret_cast<float4 *>(wb + 32 * 32 * 4);          // [32 records] of two float4
    uint2 *stack = reinterpret_cast<uint2 *>(wb + 32 * 32 * 4 + 32 * 32);
    {   // the compact rows and the table heads -> shared memory (float4 copies; both sizes are multiples of 4 floats)
        const float4 *src = reinterpret_cast<const float4 *>(C.rows);
        for (uint32_t i = threadIdx.x; i < C.rowFloats / 4; i += blockDim.x) molSmem[i] = __ldg(src + i);
        const float4 *hs = reinterpret_cast<const float4 *>(C.heads);
        float4 *hd = reinterpret_cast<float4 *>(sHeads);
        for (uint32_t i = threadIdx.x; i < C.headFloats / 4; i += blockDim.x) hd[i] = __ldg(hs + i);
    }
    const float4 lp = __ldg(&C.lam[lane]);
    const uint32_t hoff = __ldg(&C.headOff[lane]);
    const float lenM1L = lane < SKC_NS ? T.lenM1 : -1.f;
    const int kMax32 = (C.P.K - 2) * 32;
    const unsigned lanesBelow = (1u << lane) - 1u;
    __syncthreads();
    uint32_t wc[4] = {0, 0, 0, 0};
    uint32_t hist[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // contributions by size of their accept group (<8, 8-15, 16-27, 28-32): clusters, then leaf points (COUNT only)
    uint32_t grp[2] = {0, 0};          // accept events of this warp: (cluster, packet) and (leaf point, packet) groups (COUNT only)
    uint32_t gc[3] = {0, 0, 0};        // this lane's profile lookups: all in range / from the shared-memory heads / from the global tables
    // packetHits = 32, or 16 / 8 for short launches (shards of a multi-GPU frame): the contributions -- most of the work -- are
    // summed one accepting hit at a time whatever the packet size, so smaller packets cost only node-test lanes and cut the
    // run time of the heaviest packet, which is what a launch with a few packets per warp waits for
    const uint32_t nPackets = (uint32_t)((n + packetHits - 1) / packetHits);
    for (;;) {
        uint32_t packet = 0;
        if (lane == 0) packet = atomicAdd(nextPacket, 1u);
        packet = __shfl_sync(0xffffffffu, packet, 0);
        if (packet >= nPackets) break;
        uint64_t h0 = (uint64_t)packet * packetHits, h = h0 + lane;
        const bool inPacket = (uint32_t)lane < packetHits && h < n;
        const uint32_t src = (PERM && inPacket) ? __ldg(&perm[h]) : (uint32_t)h;
        bool on = inPacket && (enable ? (__ldg(&enable[(uint64_t)src * enableStrideInts]) != 0) : true);
        float px = 0.f, py = 0.f, pz = 0.f;
        if (on) { const float *pp = p3 + (uint64_t)src * strideFloats; px = __ldg(pp); py = __ldg(pp + 1); pz = __ldg(pp + 2); }
#pragma unroll
        for (int r = 0; r < 32; ++r) acc[r * 32 + lane] = 0.f;
        unsigned all = __ballot_sync(0xffffffffu, on);
        int sp = 0;
        if (all) { if (lane == 0) stack[0] = make_uint2(0u, all); sp = 1; }
        __syncwarp();
        while (sp > 0) {
            --sp;
            uint2 ent = stack[sp];
            __syncwarp();
            bool mine = (ent.y >> lane) & 1u;
            const float4 *nd = tr.nodes + 3 * (size_t)ent.x;
            float4 g = __ldg(nd), b0 = __ldg(nd + 1), b1 = __ldg(nd + 2);
            uint32_t w0 = __float_as_uint(b1.z), w1 = __float_as_uint(b1.w);
            if (COUNT) wc[0] += __popc(ent.y);
            if (w1 & SKC_OT_BLACK) continue;
            float dx = px - g.x, dy = py - g.y, dz = pz - g.z;
            float d2 = dx * dx + dy * dy + dz * dz;
            float dw = g.w / d2;
            bool inside = px >= b0.x && px <= b0.w && py >= b0.y && py <= b1.x && pz >= b0.z && pz <= b1.y;
            bool accept = mine && dw < maxError && !inside;
            bool descend = mine && !accept;
            bool addC = accept && d2 * T.rcpMin < T.lenM1;
            if (descend) {      // beyond the table range nothing below contributes
                float ex = fmaxf(fmaxf(b0.x - px, px - b0.w), 0.f);
                float ey = fmaxf(fmaxf(b0.y - py, py - b1.x), 0.f);
                float ez = fmaxf(fmaxf(b0.z - pz, pz - b1.y), 0.f);
                float db2 = ex * ex + ey * ey + ez * ez;
                descend = (db2 * 0.999f) * T.rcpMin < T.lenM1;
            }
            unsigned am = __ballot_sync(0xffffffffu, addC), dm = __ballot_sync(0xffffffffu, descend);
            if (am) {
                if (COUNT) { wc[2] += __popc(am); ++grp[0]; const int A = __popc(am); hist[A >= 28 ? 3 : A >= 16 ? 2 : A >= 8 ? 1 : 0] += A; }
                const float e = __ldg(tr.Et + (size_t)ent.x * SKC_EPAD + lane);
                if (addC) molRecord(rec, __popc(am & lanesBelow), C.P, kMax32, d2, 1.f, lane);     // records densely, in hit order
                __syncwarp();
                molContribute<COUNT, HEADS>(__popc(am), e, rec, acc, sRows, sHeads, T, lane, lp, hoff, lenM1L, gc);
                __syncwarp();
            }
            if (dm) {
                if (w1 & SKC_OT_LEAF) {
                    uint32_t cntP = w1 & 0xffu;
                    if (COUNT) wc[1] += cntP * __popc(dm);
                    const uint32_t myPoint = (uint32_t)lane < cntP ? __ldg(&tr.pointIdx[w0 + lane]) : 0u;      // the leaf's <= 8 rows, looked up once
                    for (uint32_t kp = 0; kp < cntP; ++kp) {
                        uint32_t slot = w0 + kp;
                        float4 pa = __ldg(&tr.leafPosArea[slot]);
                        float qx = px - pa.x, qy = py - pa.y, qz = pz - pa.z;
                        float q2 = qx * qx + qy * qy + qz * qz;
                        bool c = descend && q2 * T.rcpMin < T.lenM1;
                        unsigned pm = __ballot_sync(0xffffffffu, c);
                        if (pm) {
                            if (COUNT) { wc[3] += __popc(pm); ++grp[1]; const int A = __popc(pm); hist[4 + (A >= 28 ? 3 : A >= 16 ? 2 : A >= 8 ? 1 : 0)] += A; }
                            const float e = __ldg(tr.E + (size_t)__shfl_sync(0xffffffffu, myPoint, (int)kp) * SKC_EPAD + lane);
                            if (c) molRecord(rec, __popc(pm & lanesBelow), C.P, kMax32, q2, pa.w, lane);
                            __syncwarp();
                            molContribute<COUNT, HEADS>(__popc(pm), e, rec, acc, sRows, sHeads, T, lane, lp, hoff, lenM1L, gc);
                            __syncwarp();
                        }
                    }
                } else {
                    int nc = __popc(w1 & 0xffu);
                    if (lane < nc) stack[sp + lane] = make_uint2(w0 + lane, dm);
                    sp += nc;
                    __syncwarp();
                }
            }
        }
        // lane = wavelength already: stream the 32 rows out
        for (int r = 0; r < (int)packetHits; ++r) {
            uint64_t hh = h0 + r;
            if (hh >= n) break;
            if (PERM) hh = (uint64_t)__shfl_sync(0xffffffffu, src, r);
            float mo = lane < SKC_NS ? acc[r * 32 + lane] : 0.f;
            if (Mo) Mo[hh * SKC_EPAD + lane] = mo;
            if (li.L) {
                float out = liSssValue(li, hh, lane, mo);
                if (lane < SKC_NS) li.L[hh * SKC_NS + lane] = out;
            }
        }
        __syncwarp();
    }
    if (COUNT && cnt) {
        if (lane == 0) for (int k = 0; k < 4; ++k) atomicAdd(&cnt[k], (unsigned long long)wc[k]);
        atomicAdd(&cnt[4], (unsigned long long)gc[0]);
        atomicAdd(&cnt[5], (unsigned long long)gc[1]);
        atomicAdd(&cnt[6], (unsigned long long)gc[2]);
        if (lane == 0) { atomicAdd(&cnt[7], (unsigned long long)grp[0]); atomicAdd(&cnt[8], (unsigned long long)grp[1]); for (int k = 0; k < 8; ++k) atomicAdd(&cnt[9 + k], (unsigned long long)hist[k]); }
    }
}

// Li subsurface term (multipolesubsurface.cpp:279-287); one warp per hit, lane = wavelength
__global__ void k_li_sss(const skc_hit *__restrict__ hits, uint64_t n, const float *__restrict__ Mo, const float *__restrict__ rhoHd,
                         const float *__restrict__ albedo, float mix, float *__restrict__ L, const float *__restrict__ albedoPow) {
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t h = t >> 5; int lane = t & 31;
    if (h >= n) return;
    const float *hp = reinterpret_cast<const float *>(hits + h);
    int hasB = __float_as_int(__ldg(hp + 13));
    float out = 0.f;
    if (hasB && lane < SKC_NS) {
        f3 nn = mk3(__ldg(hp + 7), __ldg(hp + 8), __ldg(hp + 9)), wo = mk3(__ldg(hp + 10), __ldg(hp + 11), __ldg(hp + 12));
        float costheta = fminf(fabsf(dot3(wo, nn)), 1.f);
        float fid = costheta * (float)(SKC_RHO_LEN - 1);
        int id = min(max((int)fid, 0), SKC_RHO_LEN - 2);
        float lerp = fid - (float)id;
        float rho = (1.f - lerp) * __ldg(&rhoHd[id * SKC_EPAD + lane]) + lerp * __ldg(&rhoHd[(id + 1) * SKC_EPAD + lane]);
        float Ft = 1.f - rho;
        float a = 1.f;
        if (albedo) a = (float)pow((double)__ldg(&albedo[h * SKC_NS + lane]), (double)(1.f - mix));
        else if (albedoPow) a = __ldg(&albedoPow[lane]);
        float v = ((0.31830988618379067154f * Ft) * Mo[h * SKC_EPAD + lane]) * a;
        out = fmaxf(v, 0.f);       // Spectrum::Clamp(0)
    }
    if (lane < SKC_NS) L[h * SKC_NS + lane] = out;
}

__global__ void k_unpad(const float *__restrict__ in, uint64_t n, float *__restrict__ out) {
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t h = t >> 5; int lane = t & 31;
    if (h < n && lane < SKC_NS) out[h * SKC_NS + lane] = in[h * SKC_EPAD + lane];
}

// ---- host side ------------------------------------------------------------------------------
extern "C" int skc_upload_profile(skc_ctx *ctx, const float *table, const float *rcpDsqSpacing, uint32_t len, const float *rhoHd) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (table) {
        if (len < 2 || len >= (1u << 24)) return skc_fail(ctx, "skc_upload_profile: table length %u out of range", len);
        std::vector<float2> pairs((size_t)SKC_NS * len);
        for (int l = 0; l < SKC_NS; ++l)
            for (uint32_t i = 0; i < len; ++i)
                pairs[(size_t)l * len + i] = make_float2(table[(size_t)l * len + i], i + 1 < len ? table[(size_t)l * len + i + 1] : 0.f);
        float rcp[32]; float rmin = INFINITY;
        for (int l = 0; l < 32; ++l) { rcp[l] = l < SKC_NS ? rcpDsqSpacing[l] : rcpDsqSpacing[SKC_NS - 1]; if (l < SKC_NS) rmin = fminf(rmin, rcp[l]); }
        if (!ctx->profPairs || !ctx->profRcp || ctx->profLen != len) {      // a table of the same length goes into the arrays that are there
            if (ctx->profPairs) cudaFree(ctx->profPairs);
            if (ctx->profRcp) cudaFree(ctx->profRcp);
            ctx->profPairs = nullptr; ctx->profRcp = nullptr; ctx->profLen = 0;
            SKC_CUDA(cudaMalloc((void **)&ctx->profPairs, pairs.size() * sizeof(float2)));
            SKC_CUDA(cudaMalloc((void **)&ctx->profRcp, sizeof rcp));
        }
        SKC_CUDA(cudaMemcpyAsync(ctx->profPairs, pairs.data(), pairs.size() * sizeof(float2), cudaMemcpyHostToDevice, ctx->stream));
        SKC_CUDA(cudaMemcpyAsync(ctx->profRcp, rcp, sizeof rcp, cudaMemcpyHostToDevice, ctx->stream));
        SKC_CUDA(cudaStreamSynchronize(ctx->stream));
        ctx->profLen = len; ctx->profMaxD2 = rmin;   // rmin kept here: smallest rcpDsqSpacing
        // the compact shared-memory form (skc_mo_table.h); tables too short to be worth it stay on the exact kernel
        for (void *q : {(void *)ctx->mocRows, (void *)ctx->mocHeads, (void *)ctx->mocLam, (void *)ctx->mocHeadOff}) if (q) cudaFree(q);
        ctx->mocRows = nullptr; ctx->mocHeads = nullptr; ctx->mocLam = nullptr; ctx->mocHeadOff = nullptr; ctx->mocRowCount = 0;
        MoCompactTable C;
        if (ctx->moTable && len >= 4096 && mocBuild(table, rcpDsqSpacing, len, ctx->mocTol, MOC_K, MOC_M, getenv("SKC_MO_HEADCAP") ? (uint32_t)atoi(getenv("SKC_MO_HEADCAP")) : MOC_HEAD_CAP, C)) {
            float4 lam[32]; uint32_t hoff[32];
            for (int l = 0; l < 32; ++l) {
                lam[l] = l < SKC_NS ? make_float4(rcpDsqSpacing[l], C.fLo[l], C.fHi[l], (float)C.headLen[l]) : make_float4(0.f, 0.f, 0.f, 0.f);
                hoff[l] = C.headOff[l];
            }
            while (C.heads.size() % 4) C.heads.push_back(0.f);
            SKC_CUDA(cudaMalloc((void **)&ctx->mocRows, C.rows.size() * sizeof(float)));
            SKC_CUDA(cudaMalloc((void **)&ctx->mocHeads, C.heads.size() * sizeof(float)));
            SKC_CUDA(cudaMalloc((void **)&ctx->mocLam, sizeof lam));
            SKC_CUDA(cudaMalloc((void **)&ctx->mocHeadOff, sizeof hoff));
            SKC_CUDA(cudaMemcpyAsync(ctx->mocRows, C.rows.data(), C.rows.size() * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
            SKC_CUDA(cudaMemcpyAsync(ctx->mocHeads, C.heads.data(), C.heads.size() * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
            SKC_CUDA(cudaMemcpyAsync(ctx->mocLam, lam, sizeof lam, cudaMemcpyHostToDevice, ctx->stream));
            SKC_CUDA(cudaMemcpyAsync(ctx->mocHeadOff, hoff, sizeof hoff, cudaMemcpyHostToDevice, ctx->stream));
            SKC_CUDA(cudaStreamSynchronize(ctx->stream));
            ctx->mocRowCount = (uint32_t)C.P.K + 3; ctx->mocHeadFloats = (uint32_t)C.heads.size();
            ctx->mocInvA = C.P.invA; ctx->mocM = C.P.cellsPerBinade;
            ctx->mocCoverage = (float)C.coverage;
            float me = 0.f; for (int l = 0; l < SKC_NS; ++l) me = fmaxf(me, C.maxErr[l]);
            ctx->mocMaxErr = me;
        }
    }
    if (rhoHd) {
        std::vector<float> pad((size_t)SKC_RHO_LEN * SKC_EPAD, 0.f);
        for (int i = 0; i < SKC_RHO_LEN; ++i)
            for (int l = 0; l < SKC_NS; ++l) pad[(size_t)i * SKC_EPAD + l] = rhoHd[(size_t)i * SKC_NS + l];
        if (ctx->rhoHd) cudaFree(ctx->rhoHd);
        ctx->rhoHd = nullptr;
        SKC_CUDA(cudaMalloc((void **)&ctx->rhoHd, pad.size() * sizeof(float)));
        SKC_CUDA(cudaMemcpyAsync(ctx->rhoHd, pad.data(), pad.size() * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
        SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return 0;
}

// Hit points -> 30-bit Morton keys over the octree's bounds (10 bits per axis); hits that carry no BSSRDF sort last.
__global__ void __launch_bounds__(256) k_mo_sort_keys(const float *__restrict__ p3, uint64_t strideFloats, const int32_t *__restrict__ enable,
                                                      uint64_t enableStrideInts, uint32_t n, float3 bmin, float3 scale,
                                                      uint32_t *__restrict__ keys, uint32_t *__restrict__ vals) {
    uint32_t i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    uint32_t key = 0x40000000u;
    if (!enable || __ldg(&enable[(uint64_t)i * enableStrideInts]) != 0) {
        const float *pp = p3 + (uint64_t)i * strideFloats;
        auto cell = [](float v) { return (uint32_t)fminf(fmaxf(v, 0.f), 1023.f); };
        auto spread = [](uint32_t x) { x = (x | (x << 16)) & 0x030000FFu; x = (x | (x << 8)) & 0x0300F00Fu; x = (x | (x << 4)) & 0x030C30C3u; return (x | (x << 2)) & 0x09249249u; };
        key = spread(cell((__ldg(pp) - bmin.x) * scale.x)) | (spread(cell((__ldg(pp + 1) - bmin.y) * scale.y)) << 1) | (spread(cell((__ldg(pp + 2) - bmin.z) * scale.z)) << 2);
    }
    keys[i] = key; vals[i] = i;
}

// The packet kernel walks the octree once for 32 hit points, so it wants neighbouring points in a packet.  The frame path
// renders its samples along a Z-curve over the film; callers of skc_mo / skc_li_sss / skc_camera_pass* own the order of
// their points, so those are sorted here along a Morton curve through the octree's bounds (one 32-bit radix sort) and the
// kernel reads and writes through the permutation -- results land at the caller's indices, bit-identical to the unsorted run.
// *failed: the sort was wanted and could not run (scratch allocation or launch error; skc_last_error says which) -- the caller fails
// instead of silently running the kernel on unsorted points at a fraction of its speed
static const uint32_t *moSortHits(skc_ctx *ctx, const float *p3Dev, uint64_t strideFloats, const int32_t *enableDev, uint64_t enableStride, uint64_t n, bool *failed) {
    *failed = false;
    if (n < 64 || n >= (1ull << 31)) return nullptr;
    size_t tempBytes = 0;
    cub::DoubleBuffer<uint32_t> dk(nullptr, nullptr), dv(nullptr, nullptr);
    cub::DeviceRadixSort::SortPairs(nullptr, tempBytes, dk, dv, (int)n, 0, 31, ctx->stream);
    size_t arr = ((size_t)n * 4 + 255) & ~(size_t)255;
    char *buf = (char *)skc_scratch(ctx, 16, 4 * arr + tempBytes);
    if (!buf) { *failed = true; return nullptr; }
    dk = cub::DoubleBuffer<uint32_t>((uint32_t *)buf, (uint32_t *)(buf + arr));
    dv = cub::DoubleBuffer<uint32_t>((uint32_t *)(buf + 2 * arr), (uint32_t *)(buf + 3 * arr));
    const float *b = ctx->otBoundsHost;
    float3 bmin = make_float3(b[0], b[1], b[2]);
    auto sc = [](float lo, float hi) { return hi > lo ? 1024.f / (hi - lo) : 0.f; };
    float3 scale = make_float3(sc(b[0], b[3]), sc(b[1], b[4]), sc(b[2], b[5]));
    k_mo_sort_keys<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(p3Dev, strideFloats, enableDev, enableStride, (uint32_t)n, bmin, scale, dk.Current(), dv.Current());
    ctx->launches += 2;
    if (cudaGetLastError() != cudaSuccess || cub::DeviceRadixSort::SortPairs(buf + 4 * arr, tempBytes, dk, dv, (int)n, 0, 31, ctx->stream) != cudaSuccess) {
        skc_fail(ctx, "skc_mo: sorting the hit points failed: %s", cudaGetErrorString(cudaGetLastError()));
        *failed = true; return nullptr;
    }
    return dv.Current();
}

// k_mo_lambda or k_mo_packet?  The compact table pays when most lookups fall inside its verified windows.  Where the windows are
// short -- profiles whose near range is sampled too coarsely by the reference's own table to be smooth at 1e-5, e.g. LayeredSkin
// at its defaults on a unit sphere: coverage 0.11, 82 % of the lookups exact -- lane = wavelength is the wrong layout for the
// remaining gathers (72 vs 47 ms there), while skin.scene has coverage 0.62 and runs 42 vs 60 ms.
#define MOC_MIN_COVERAGE 0.35
static bool molWanted(const skc_ctx *ctx) {
    if (!ctx->moTable || !ctx->mocRows || 7 * (int)ctx->otLevels + 1 > MOL_STACK) return false;
    return ctx->moTable == 2 || ctx->mocCoverage >= MOC_MIN_COVERAGE;
}

static int moLaunch(skc_ctx *ctx, const float *p3Dev, uint64_t strideFloats, const int32_t *enableDev, uint64_t enableStride,
                    uint64_t n, float maxError, float *MoDev, unsigned long long *cntDev = nullptr, const LiEpilogue *li = nullptr,
                    bool callerOrder = true) {
    if (!ctx->otGeo || !ctx->otNodes) return skc_fail(ctx, "skc_mo: no octree built (skc_build_octree)");
    if (!ctx->profPairs) return skc_fail(ctx, "skc_mo: no profile table uploaded (skc_upload_profile)");
    if (!n) return 0;
    MoTree tr; tr.nodes = ctx->otGeo; tr.Et = ctx->otEt; tr.leafPosArea = ctx->otLeafPosArea; tr.pointIdx = ctx->otPointIdx; tr.E = ctx->ipE;
    MoTables T; T.pairs = ctx->profPairs; T.rcp = ctx->profRcp; T.len = ctx->profLen; T.rcpMin = ctx->profMaxD2; T.lenM1 = (float)(ctx->profLen - 1);
    if (ctx->moPacket && molWanted(ctx)) {
        if (n >= (1ull << 32)) return skc_fail(ctx, "skc_mo: more than 2^32 hit points in one call");
        MoCompactDev C; C.rows = ctx->mocRows; C.heads = ctx->mocHeads; C.lam = ctx->mocLam; C.headOff = ctx->mocHeadOff;
        C.rowFloats = ctx->mocRowCount * 32; C.headFloats = ctx->mocHeadFloats;
        C.P.invA = ctx->mocInvA; C.P.cellsPerBinade = ctx->mocM; C.P.K = (int)ctx->mocRowCount - 3;
        const size_t tableBytes = ((size_t)C.rowFloats + C.headFloats) * sizeof(float);
        int warps = (int)std::min<size_t>(MOL_MAX_WARPS, ((size_t)ctx->smemOptin - 1024 - tableBytes) / MOL_WARP_BYTES);
        if (ctx->moWarps > 0) warps = std::min(warps, ctx->moWarps);
        // long launches (from 16 packets of 32 per resident warp: a whole frame) run 2 % faster with 23 warps per CTA than with the 27
        // that fit -- skin.scene 40.6 -> 39.8 ms alone, 82.6 -> 82.0 ms in the frame, three runs each -- while a rank's share of a
        // frame is faster with 27 (1/2: 21.7 vs 22.1 ms, 1/8: 7.07 vs 7.34 ms; tools/mo_shard_ab.py with SKC_MO_WARPS)
        else if (warps > 23 && (n + 31) / 32 >= 16ull * (uint64_t)ctx->smCount * (uint64_t)warps) warps = 23;
        if (warps < 4) return skc_fail(ctx, "skc_mo: the compact profile table (%zu B) leaves no room for the traversal state in shared memory", tableBytes);
        const size_t smem = tableBytes + (size_t)warps * MOL_WARP_BYTES;
        // packets per resident warp decide the packet size (see the kernel): 32 hits from 8 packets per warp up, else 16
        // (tools/mo_shard_ab.py, a rank's share of the skin.scene frame: 1/4 of it 13.1 -> 11.6 ms, 1/8 8.3 -> 7.1 ms; packets of
        // 8 are no better than 16, and on the whole frame 16 costs 3 %)
        const uint64_t residentWarps = (uint64_t)ctx->smCount * warps;
        uint32_t packetHits = 32;
        if (ctx->moPacketHits > 0) packetHits = (uint32_t)ctx->moPacketHits;
        else if ((n + 31) / 32 < 8 * residentWarps) packetHits = 16;
        const uint64_t nPk = (n + packetHits - 1) / packetHits;
        const unsigned grid = (unsigned)std::min<uint64_t>((nPk + warps - 1) / warps, (uint64_t)ctx->smCount);
        LiEpilogue ep = li ? *li : LiEpilogue{nullptr, nullptr, nullptr, 0.f, nullptr};
        uint32_t *next = (uint32_t *)skc_scratch(ctx, 6, 256);
        if (!next) return 1;
        SKC_CUDA(cudaMemsetAsync(next, 0, sizeof(uint32_t), ctx->stream));
        bool sortFailed = false;
        const uint32_t *perm = (ctx->moSort > 0 || (ctx->moSort < 0 && callerOrder)) ? moSortHits(ctx, p3Dev, strideFloats, enableDev, enableStride, n, &sortFailed) : nullptr;
        if (sortFailed) return 1;
#define MOL_GO2(CNT, P, H, W) do { \
            SKC_CUDA(cudaFuncSetAttribute(k_mo_lambda<CNT, P, H, W>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
            k_mo_lambda<CNT, P, H, W><<<grid, warps * 32, smem, ctx->stream>>>(tr, T, C, p3Dev, strideFloats, enableDev, enableStride, n, maxError, MoDev, cntDev, ep, next, perm, packetHits); } while (0)
#define MOL_GO(CNT, P) do { if (C.headFloats <= 4) { if (warps > 20) MOL_GO2(CNT, P, false, 28); else MOL_GO2(CNT, P, false, 20); } else { if (warps > 20) MOL_GO2(CNT, P, true, 28); else MOL_GO2(CNT, P, true, 20); } } while (0)
        if (cntDev) { if (perm) MOL_GO(true, true); else MOL_GO(true, false); } else { if (perm) MOL_GO(false, true); else MOL_GO(false, false); }
#undef MOL_GO2
#undef MOL_GO
        SKC_LAUNCHED("k_mo_lambda");
        return 0;
    }
    if (ctx->moPacket) {
        if (n >= (1ull << 32)) return skc_fail(ctx, "skc_mo: more than 2^32 hit points in one call");
        if (7 * (int)ctx->otLevels + 1 > MOP_STACK) return skc_fail(ctx, "skc_mo: octree with %u levels exceeds the packet traversal stack", ctx->otLevels);
        uint64_t pb = (n + MOP_WARPS * 32 - 1) / (MOP_WARPS * 32);
        if (pb > (uint64_t)ctx->smCount * MOP_CTAS1) pb = (uint64_t)ctx->smCount * MOP_CTAS1;       // persistent: at most the resident CTAs
        LiEpilogue ep = li ? *li : LiEpilogue{nullptr, nullptr, nullptr, 0.f, nullptr};
        uint32_t *next = (uint32_t *)skc_scratch(ctx, 6, 256);
        if (!next) return 1;
        SKC_CUDA(cudaMemsetAsync(next, 0, sizeof(uint32_t), ctx->stream));
        bool sortFailed = false;
        const uint32_t *perm = (ctx->moSort > 0 || (ctx->moSort < 0 && callerOrder)) ? moSortHits(ctx, p3Dev, strideFloats, enableDev, enableStride, n, &sortFailed) : nullptr;
        if (sortFailed) return 1;
        // one warp per packet when there are plenty of packets per resident warp; two warps per packet (half the spectrum each)
        // when a launch is short (multi-GPU shards), where the heaviest packets' run time is what the launch waits for.
        // Measured on skin.scene: 28 packets/warp 60.4 vs 61.9 ms, 14 packets/warp 38.1 vs 35.7 ms, 3.5 packets/warp 23.0 vs 17.5 ms
        // ... and four warps per packet (a quarter of the spectrum each) below 5 packets per resident warp: 1/8 frame 15.5 -> 13.6 ms, 1/4 frame 22.7 vs 22.5 ms
        const uint64_t nPk = (n + 31) / 32, resident = (uint64_t)ctx->smCount * 8 * MOP_WARPS;
        int split = ctx->moSplit ? ctx->moSplit : (nPk >= 20ull * ctx->smCount * 7 * MOP_WARPS ? 1 : (nPk < 5 * resident ? 4 : 2));
        if (split >= 2) pb = std::min<uint64_t>((uint64_t)split * ((n + MOP_WARPS * 32 - 1) / (MOP_WARPS * 32)), (uint64_t)ctx->smCount * (split == 4 ? MOP_CTAS4 : 8));
#define MOP_GO(C, S, P) k_mo_packet<C, S, P><<<(unsigned)pb, MOP_WARPS * 32, 0, ctx->stream>>>(tr, T, p3Dev, strideFloats, enableDev, enableStride, n, maxError, MoDev, cntDev, ep, next, perm)
        if (split == 4) { if (cntDev) { if (perm) MOP_GO(true, 4, true); else MOP_GO(true, 4, false); } else { if (perm) MOP_GO(false, 4, true); else MOP_GO(false, 4, false); } }
        else if (split == 2) { if (cntDev) { if (perm) MOP_GO(true, 2, true); else MOP_GO(true, 2, false); } else { if (perm) MOP_GO(false, 2, true); else MOP_GO(false, 2, false); } }
        else            { if (cntDev) { if (perm) MOP_GO(true, 1, true); else MOP_GO(true, 1, false); } else { if (perm) MOP_GO(false, 1, true); else MOP_GO(false, 1, false); } }
#undef MOP_GO
        SKC_LAUNCHED("k_mo_packet");
        return 0;
    }
    if (li) return skc_fail(ctx, "skc_mo: the fused Li epilogue needs the packet kernel");
    uint64_t per = MO_WARPS * MO_HITS_PER_WARP;
    uint64_t blocks = (n + per - 1) / per;
    // stack need: MO_HIGH-1 + 7*32 + 7*levels
    bool deep = 63 + 224 + 7 * (int)ctx->otLevels > 512;
    if (63 + 224 + 7 * (int)ctx->otLevels > 1024) return skc_fail(ctx, "skc_mo: octree with %u levels exceeds the traversal stack", ctx->otLevels);
    if (cntDev) {
        if (deep) k_mo<true, 1024><<<(unsigned)blocks, MO_WARPS * 32, 0, ctx->stream>>>(tr, T, p3Dev, strideFloats, enableDev, enableStride, n, maxError, MoDev, cntDev);
        else k_mo<true, 512><<<(unsigned)blocks, MO_WARPS * 32, 0, ctx->stream>>>(tr, T, p3Dev, strideFloats, enableDev, enableStride, n, maxError, MoDev, cntDev);
    } else {
        if (deep) k_mo<false, 1024><<<(unsigned)blocks, MO_WARPS * 32, 0, ctx->stream>>>(tr, T, p3Dev, strideFloats, enableDev, enableStride, n, maxError, MoDev, nullptr);
        else k_mo<false, 512><<<(unsigned)blocks, MO_WARPS * 32, 0, ctx->stream>>>(tr, T, p3Dev, strideFloats, enableDev, enableStride, n, maxError, MoDev, nullptr);
    }
    SKC_LAUNCHED("k_mo");
    return 0;
}

int skc_mo_dev(skc_ctx *ctx, const float *p3Dev, uint64_t strideFloats, uint64_t n, float maxError, float *MoDev) {
    return moLaunch(ctx, p3Dev, strideFloats, nullptr, 0, n, maxError, MoDev);
}

extern "C" int skc_mo(skc_ctx *ctx, const float *p3, uint64_t n, float maxError, float *Mo30) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!n) return 0;
    float *dP = (float *)skc_scratch(ctx, 0, n * 3 * sizeof(float));
    float *dMo = (float *)skc_scratch(ctx, 1, n * SKC_EPAD * sizeof(float));
    float *dOut = (float *)skc_scratch(ctx, 2, n * SKC_NS * sizeof(float));
    if (!dP || !dMo || !dOut) return 1;
    SKC_CUDA(cudaMemcpyAsync(dP, p3, n * 3 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    skc_stage_begin(ctx);
    if (moLaunch(ctx, dP, 3, nullptr, 0, n, maxError, dMo)) return 1;
    skc_stage_end(ctx, 4);
    uint64_t threads = n * 32;
    k_unpad<<<(unsigned)((threads + 255) / 256), 256, 0, ctx->stream>>>(dMo, n, dOut);
    SKC_LAUNCHED("k_unpad");
    SKC_CUDA(cudaMemcpyAsync(Mo30, dOut, n * SKC_NS * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// hits on the device -> Mo (padded, scratch 1) -> L (scratch 2); shared by li_sss and camera_pass
static int liDev(skc_ctx *ctx, const skc_hit *dHits, uint64_t n, float maxError, float mix, const float *albedo30,
                 float *Mo30, float *L30, unsigned long long *cntDev = nullptr, bool computeL = false) {
    float *dMo = (float *)skc_scratch(ctx, 4, n * SKC_EPAD * sizeof(float));
    float *dL = (float *)skc_scratch(ctx, 2, n * SKC_NS * sizeof(float));
    if (!dMo || !dL) return 1;
    if (!ctx->rhoHd) return skc_fail(ctx, "skc_li_sss: no rho table uploaded (skc_upload_profile)");
    float *dAlb = nullptr;
    if (albedo30) {
        dAlb = (float *)skc_scratch(ctx, 3, n * SKC_NS * sizeof(float));
        if (!dAlb) return 1;
        SKC_CUDA(cudaMemcpyAsync(dAlb, albedo30, n * SKC_NS * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    }
    skc_stage_begin(ctx);
    // p at float offset 4, hasBssrdf at int offset 13 of the 14-word hit record
    if (moLaunch(ctx, reinterpret_cast<const float *>(dHits) + 4, 14, reinterpret_cast<const int32_t *>(dHits) + 13, 14, n, maxError, dMo, cntDev)) return 1;
    skc_stage_end(ctx, 4);
    uint64_t threads = n * 32;
    if (L30 || computeL) {
        k_li_sss<<<(unsigned)((threads + 255) / 256), 256, 0, ctx->stream>>>(dHits, n, dMo, ctx->rhoHd, dAlb, mix, dL, ctx->albedoFactors ? ctx->albedoFactors + 32 : nullptr);
        SKC_LAUNCHED("k_li_sss");
        if (L30) SKC_CUDA(cudaMemcpyAsync(L30, dL, n * SKC_NS * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    }
    if (Mo30) {
        float *dOut = (float *)skc_scratch(ctx, 5, n * SKC_NS * sizeof(float));
        if (!dOut) return 1;
        k_unpad<<<(unsigned)((threads + 255) / 256), 256, 0, ctx->stream>>>(dMo, n, dOut);
        SKC_LAUNCHED("k_unpad");
        SKC_CUDA(cudaMemcpyAsync(Mo30, dOut, n * SKC_NS * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    }
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// (the frame path: its samples already follow a Z-curve over the film)
int skc_li_fused_dev(skc_ctx *ctx, const skc_hit *dHits, uint64_t n, float maxError, float mix, float *dL) {
    if (!ctx->rhoHd) return skc_fail(ctx, "skc_li: no rho table uploaded (skc_upload_profile)");
    if (!ctx->moPacket) return skc_fail(ctx, "skc_li: the fused Li epilogue needs the packet kernel");
    LiEpilogue ep{dHits, ctx->rhoHd, nullptr, mix, dL, ctx->albedoFactors ? ctx->albedoFactors + 32 : nullptr, ctx->hitAlbedo, ctx->albedoTex.refl};
    skc_stage_begin(ctx);
    if (moLaunch(ctx, reinterpret_cast<const float *>(dHits) + 4, 14, reinterpret_cast<const int32_t *>(dHits) + 13, 14, n, maxError, nullptr, nullptr, &ep, false)) return 1;
    skc_stage_end(ctx, 4);
    return 0;
}

extern "C" int skc_li_sss(skc_ctx *ctx, const skc_hit *hits, uint64_t n, float maxError, float mix,
                          const float *albedo30, float *Mo30, float *L30) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!n) return 0;
    skc_hit *dHits = (skc_hit *)skc_scratch(ctx, 1, n * sizeof(skc_hit));
    if (!dHits) return 1;
    SKC_CUDA(cudaMemcpyAsync(dHits, hits, n * sizeof(skc_hit), cudaMemcpyHostToDevice, ctx->stream));
    return liDev(ctx, dHits, n, maxError, mix, albedo30, Mo30, L30);
}

// Host-buffer camera pass: rays up, closest hit, Mo with the fused Li epilogue.  When L30 is pinned host memory
// (skc_alloc_pinned / cudaHostAlloc) the kernel writes the radiance rows straight into it; otherwise they are staged
// in device memory and copied afterwards.  The hit records travel back on the copy stream while Mo runs.
extern "C" int skc_camera_pass(skc_ctx *ctx, const skc_ray *rays, uint64_t n, float maxError, float mix,
                               skc_hit *hitsOut, float *L30) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!n) return 0;
    if (!ctx->rhoHd) return skc_fail(ctx, "skc_camera_pass: no rho table uploaded (skc_upload_profile)");
    skc_ray *dRays = (skc_ray *)skc_scratch(ctx, 0, n * sizeof(skc_ray));
    skc_hit *dHits = (skc_hit *)skc_scratch(ctx, 1, n * sizeof(skc_hit));
    if (!dRays || !dHits) return 1;
    SKC_CUDA(cudaMemcpyAsync(dRays, rays, n * sizeof(skc_ray), cudaMemcpyHostToDevice, ctx->stream));
    skc_stage_begin(ctx);
    if (skc_trace_closest_dev(ctx, dRays, n, dHits, nullptr)) return 1;
    skc_stage_end(ctx, 3);
    if (hitsOut) {
        cudaEvent_t traced = skc_event(ctx, 0);
        if (!traced) return 1;
        SKC_CUDA(cudaEventRecord(traced, ctx->stream));
        SKC_CUDA(cudaStreamWaitEvent(ctx->copyStream, traced, 0));
        SKC_CUDA(cudaMemcpyAsync(hitsOut, dHits, n * sizeof(skc_hit), cudaMemcpyDeviceToHost, ctx->copyStream));
    }
    if (!ctx->moPacket || !L30) {
        int rc = liDev(ctx, dHits, n, maxError, mix, nullptr, nullptr, L30);
        cudaError_t e = cudaStreamSynchronize(ctx->copyStream);
        if (rc) return rc;
        SKC_CUDA(e);
        return 0;
    }
    float *target = nullptr, *dL = nullptr;
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, L30) == cudaSuccess && attr.type == cudaMemoryTypeHost && attr.devicePointer) target = (float *)attr.devicePointer;
    else {
        cudaGetLastError();
        dL = (float *)skc_scratch(ctx, 2, n * SKC_NS * sizeof(float));
        if (!dL) return 1;
        target = dL;
    }
    LiEpilogue ep{dHits, ctx->rhoHd, nullptr, mix, target, ctx->albedoFactors ? ctx->albedoFactors + 32 : nullptr, ctx->hitAlbedo, ctx->albedoTex.refl};
    skc_stage_begin(ctx);
    if (moLaunch(ctx, reinterpret_cast<const float *>(dHits) + 4, 14, reinterpret_cast<const int32_t *>(dHits) + 13, 14, n, maxError, nullptr, nullptr, &ep)) return 1;
    skc_stage_end(ctx, 4);
    if (dL) SKC_CUDA(cudaMemcpyAsync(L30, dL, n * SKC_NS * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->copyStream));
    return 0;
}

// ---- resident camera rays -----------------------------------------------------------------
extern "C" int skc_set_camera_rays(skc_ctx *ctx, const skc_ray *rays, uint64_t n) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (ctx->resRays) { cudaFree(ctx->resRays); ctx->resRays = nullptr; }
    ctx->resRaysCount = 0;
    if (!n) return 0;
    SKC_CUDA(cudaMalloc((void **)&ctx->resRays, n * sizeof(skc_ray)));
    SKC_CUDA(cudaMemcpyAsync(ctx->resRays, rays, n * sizeof(skc_ray), cudaMemcpyHostToDevice, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->resRaysCount = n;
    return 0;
}

// Scene::Intersect / IntersectP alone over the resident rays (BVH microbenchmark of SURVEY.md 8d): anyHit = 0 closest hit
// (hitsOut: host, count skc_hit records, or NULL), 1 = any hit (hitsOut: host, count bytes, or NULL).  counters2
// (optional) = {slab tests, primitive tests}.  Device time lands in stage "closest" / "any".
extern "C" int skc_trace_resident(skc_ctx *ctx, uint64_t first, uint64_t count, int anyHit, void *hitsOut, uint64_t *counters2) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!ctx->resRays) return skc_fail(ctx, "skc_trace_resident: no camera rays set");
    if (first + count > ctx->resRaysCount) return skc_fail(ctx, "skc_trace_resident: bad range");
    if (!count) return 0;
    size_t outBytes = anyHit ? count : count * sizeof(skc_hit);
    void *dOut = skc_scratch(ctx, 1, outBytes);
    if (!dOut) return 1;
    unsigned long long *dCnt = nullptr;
    if (counters2) {
        dCnt = (unsigned long long *)skc_scratch(ctx, 7, 8 * sizeof(unsigned long long));
        if (!dCnt) return 1;
        SKC_CUDA(cudaMemsetAsync(dCnt, 0, 8 * sizeof(unsigned long long), ctx->stream));
    }
    skc_stage_begin(ctx);
    int rc = anyHit ? skc_trace_any_dev(ctx, ctx->resRays + first, count, (uint8_t *)dOut, dCnt)
                    : skc_trace_closest_dev(ctx, ctx->resRays + first, count, (skc_hit *)dOut, dCnt);
    if (rc) return 1;
    skc_stage_end(ctx, anyHit ? 5 : 3);
    if (hitsOut) SKC_CUDA(cudaMemcpyAsync(hitsOut, dOut, outBytes, cudaMemcpyDeviceToHost, ctx->stream));
    if (counters2) {
        unsigned long long h[2];
        SKC_CUDA(cudaMemcpyAsync(h, dCnt, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
        SKC_CUDA(cudaStreamSynchronize(ctx->stream));
        counters2[0] = h[0]; counters2[1] = h[1];
    }
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// counters7 (optional): [0] slab tests [1] primitive tests (closest hit), [2] octree nodes classified,
// [3] leaf points tested, [4] cluster contributions, [5] point contributions, [6] profile-table reads (packet kernel)
extern "C" int skc_camera_pass_resident(skc_ctx *ctx, uint64_t first, uint64_t count, float maxError, float mix,
                                        float *L30, uint64_t *counters7) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!ctx->resRays) return skc_fail(ctx, "skc_camera_pass_resident: no camera rays set");
    if (first + count > ctx->resRaysCount) return skc_fail(ctx, "skc_camera_pass_resident: bad range");
    if (!count) return 0;
    skc_hit *dHits = (skc_hit *)skc_scratch(ctx, 1, count * sizeof(skc_hit));
    if (!dHits) return 1;
    unsigned long long *dCnt = nullptr;
    if (counters7) {
        dCnt = (unsigned long long *)skc_scratch(ctx, 7, 32 * sizeof(unsigned long long));
        if (!dCnt) return 1;
        SKC_CUDA(cudaMemsetAsync(dCnt, 0, 32 * sizeof(unsigned long long), ctx->stream));
    }
    skc_stage_begin(ctx);
    if (skc_trace_closest_dev(ctx, ctx->resRays + first, count, dHits, dCnt)) return 1;
    skc_stage_end(ctx, 3);
    if (liDev(ctx, dHits, count, maxError, mix, nullptr, nullptr, L30, dCnt ? dCnt + 2 : nullptr, true)) return 1;
    if (counters7) {
        unsigned long long h[32];
        SKC_CUDA(cudaMemcpyAsync(h, dCnt, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
        SKC_CUDA(cudaStreamSynchronize(ctx->stream));
        for (int k = 0; k < 7; ++k) counters7[k] = h[k];
        ctx->moLookups[0] = h[6]; ctx->moLookups[1] = h[7]; ctx->moLookups[2] = h[8]; ctx->moGroups[0] = h[9]; ctx->moGroups[1] = h[10]; for (int k = 0; k < 8; ++k) ctx->moHist[k] = h[11 + k];
    }
    return 0;
}

// Compact profile table of the Mo kernel (skc_mo_table.h): stats8 = {rows, shared-memory bytes (rows + heads), coverage (log-weighted
// share of table segments served from shared memory), largest verified relative error, and from the last counted camera pass:
// profile lookups in range, of those from the shared-memory heads, of those from the global tables, 0}.  rows = 0: exact tables only.
extern "C" int skc_mo_table_stats(skc_ctx *ctx, double *stats8) {
    if (!ctx) return 1;
    stats8[0] = ctx->mocRows ? ctx->mocRowCount : 0; stats8[1] = ctx->mocRows ? ((double)ctx->mocRowCount * 32 + ctx->mocHeadFloats) * 4 : 0;
    stats8[2] = ctx->mocRows ? ctx->mocCoverage : 0; stats8[3] = ctx->mocRows ? ctx->mocMaxErr : 0;
    stats8[4] = (double)ctx->moLookups[0]; stats8[5] = (double)ctx->moLookups[1]; stats8[6] = (double)ctx->moLookups[2]; stats8[7] = (double)ctx->moGroups[0] + (double)ctx->moGroups[1];
    return 0;
}
// contributions of the last counted camera pass by the size of their accept group: clusters {<8, 8-15, 16-27, 28-32}, leaf points {same}
extern "C" int skc_mo_group_histogram(skc_ctx *ctx, uint64_t *hist8) { if (!ctx) return 1; for (int k = 0; k < 8; ++k) hist8[k] = ctx->moHist[k]; return 0; }

// ---------------------------------------------------------------------------------------------
// skc_mo_static: SubsurfaceOctreeNode::Mo for callers that must not allocate (code under CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Same kernels as skc_mo, but every byte of scratch -- the packet ticket counter, the staged points, the padded [n][32]
// rows and the [n][30] output staging -- lives in ONE file-scope device array of fixed size, so the call makes no
// cudaMalloc / cudaFree and uses no allocating library routine (no hit sort either: the points are walked in the
// caller's order, which changes packet coherence, not results).  Every GPU that loads the library holds the array until
// exit.  All calls share it and it keeps the last call's contents: the ticket counter is cleared per call, everything
// else is fully overwritten before it is read; two calls that overlap on the GPU (two contexts or host threads on one
// device) race on it.
#define MOS_OFF_P3 256u                                                     // after the counter's 256 B
#define MOS_OFF_PAD (MOS_OFF_P3 + SKC_MO_STATIC_MAX_HITS * 12u)              // 12 B per point
#define MOS_OFF_OUT (MOS_OFF_PAD + SKC_MO_STATIC_MAX_HITS * SKC_EPAD * 4u)    // 128 B per padded row
#define MOS_BYTES (MOS_OFF_OUT + SKC_MO_STATIC_MAX_HITS * SKC_NS * 4u)        // 120 B per output row: 4 260 096 B in all
static __device__ __align__(256) unsigned char g_moStaticScratch[MOS_BYTES];

// In host code `g_moStaticScratch` is a host-side stand-in, not a device address: resolve it once per device, with that
// device current, and hand the kernels the resolved pointer.
static unsigned char *moStaticBase(skc_ctx *ctx) {
    static unsigned char *base[64] = {nullptr};
    const int d = ctx->device;
    if (d < 0 || d >= 64) { skc_fail(ctx, "skc_mo_static: device %d out of range", d); return nullptr; }
    if (!base[d]) {
        unsigned char *p = nullptr;
        if (skc_check(ctx, cudaGetSymbolAddress((void **)&p, g_moStaticScratch), "cudaGetSymbolAddress(g_moStaticScratch)")) return nullptr;
        base[d] = p;
    }
    return base[d];
}

extern "C" int skc_mo_static(skc_ctx *ctx, const float *p3, uint64_t n, float maxError, float *Mo30) {
    if (!ctx) return 1;
    if (cudaSetDevice(ctx->device) != cudaSuccess) return skc_fail(ctx, "skc_mo_static: cudaSetDevice(%d) failed", ctx->device);
    if (n > SKC_MO_STATIC_MAX_HITS)
        return skc_fail(ctx, "skc_mo_static: %llu points exceed the static scratch capacity of %u points per call (split the call, or use skc_mo)",
                        (unsigned long long)n, SKC_MO_STATIC_MAX_HITS);
    if (!n) return 0;
    if (!p3 || !Mo30) return skc_fail(ctx, "skc_mo_static: null input or output");
    if (!ctx->otGeo || !ctx->otNodes) return skc_fail(ctx, "skc_mo_static: no octree built (skc_build_octree)");
    if (!ctx->profPairs) return skc_fail(ctx, "skc_mo_static: no profile table uploaded (skc_upload_profile)");
    unsigned char *base = moStaticBase(ctx);
    if (!base) return 1;
    uint32_t *next = reinterpret_cast<uint32_t *>(base);
    float *dP = reinterpret_cast<float *>(base + MOS_OFF_P3), *dPad = reinterpret_cast<float *>(base + MOS_OFF_PAD), *dOut = reinterpret_cast<float *>(base + MOS_OFF_OUT);
    // p3 / Mo30 may be host or device memory (unified addressing picks the direction)
    SKC_CUDA(cudaMemcpyAsync(dP, p3, n * 3 * sizeof(float), cudaMemcpyDefault, ctx->stream));
    SKC_CUDA(cudaMemsetAsync(next, 0, sizeof(uint32_t), ctx->stream));
    MoTree tr; tr.nodes = ctx->otGeo; tr.Et = ctx->otEt; tr.leafPosArea = ctx->otLeafPosArea; tr.pointIdx = ctx->otPointIdx; tr.E = ctx->ipE;
    MoTables T; T.pairs = ctx->profPairs; T.rcp = ctx->profRcp; T.len = ctx->profLen; T.rcpMin = ctx->profMaxD2; T.lenM1 = (float)(ctx->profLen - 1);
    const LiEpilogue ep{nullptr, nullptr, nullptr, 0.f, nullptr};
    if (molWanted(ctx)) {
        MoCompactDev C; C.rows = ctx->mocRows; C.heads = ctx->mocHeads; C.lam = ctx->mocLam; C.headOff = ctx->mocHeadOff;
        C.rowFloats = ctx->mocRowCount * 32; C.headFloats = ctx->mocHeadFloats;
        C.P.invA = ctx->mocInvA; C.P.cellsPerBinade = ctx->mocM; C.P.K = (int)ctx->mocRowCount - 3;
        const size_t tableBytes = ((size_t)C.rowFloats + C.headFloats) * sizeof(float);
        const int warps = (int)std::min<size_t>(20, ((size_t)ctx->smemOptin - 1024 - tableBytes) / MOL_WARP_BYTES);
        if (warps < 4) return skc_fail(ctx, "skc_mo_static: the compact profile table (%zu B) leaves no room for the traversal state in shared memory", tableBytes);
        const size_t smem = tableBytes + (size_t)warps * MOL_WARP_BYTES;
        const unsigned grid = (unsigned)std::min<uint64_t>(((n + 31) / 32 + warps - 1) / warps, (uint64_t)ctx->smCount);
        if (C.headFloats <= 4) {
            SKC_CUDA(cudaFuncSetAttribute(k_mo_lambda<false, false, false, 20>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            k_mo_lambda<false, false, false, 20><<<grid, warps * 32, smem, ctx->stream>>>(tr, T, C, dP, 3, nullptr, 0, n, maxError, dPad, nullptr, ep, next, nullptr, 32u);
        } else {
            SKC_CUDA(cudaFuncSetAttribute(k_mo_lambda<false, false, true, 20>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            k_mo_lambda<false, false, true, 20><<<grid, warps * 32, smem, ctx->stream>>>(tr, T, C, dP, 3, nullptr, 0, n, maxError, dPad, nullptr, ep, next, nullptr, 32u);
        }
        SKC_LAUNCHED("k_mo_lambda (static scratch)");
    } else {
        if (7 * (int)ctx->otLevels + 1 > MOP_STACK) return skc_fail(ctx, "skc_mo_static: octree with %u levels exceeds the packet traversal stack", ctx->otLevels);
        const uint64_t pb = std::min<uint64_t>((n + MOP_WARPS * 32 - 1) / (MOP_WARPS * 32), (uint64_t)ctx->smCount * MOP_CTAS1);
        k_mo_packet<false, 1, false><<<(unsigned)pb, MOP_WARPS * 32, 0, ctx->stream>>>(tr, T, dP, 3, nullptr, 0, n, maxError, dPad, nullptr, ep, next, nullptr);
        SKC_LAUNCHED("k_mo_packet (static scratch)");
    }
    k_unpad<<<(unsigned)((n * 32 + 255) / 256), 256, 0, ctx->stream>>>(dPad, n, dOut);
    SKC_LAUNCHED("k_unpad (static scratch)");
    SKC_CUDA(cudaMemcpyAsync(Mo30, dOut, n * SKC_NS * sizeof(float), cudaMemcpyDefault, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

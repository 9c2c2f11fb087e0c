// skc_tess.cu -- surface-point generation on the device.
//
// Reference: TriangleMesh::TessellateSurfacePoints (shapes/trianglemesh.cpp:187-257), tessellator (:264-316),
// matching (:319-351); GetSurfacePointsThroughTessellation (renderers/surfacepoints.cpp:362-369).
//
// Per triangle the reference runs a DX11-style ring tessellator and emits one SurfacePoint per sub-triangle, in a
// fixed order: ring by ring from the centre outwards, per ring the three sides, per side a zipper walk (`matching`)
// between the inner and the outer edge.  The order matters downstream (a point's index selects its scramble stream and
// its place in an octree leaf), so it is reproduced exactly -- but a walk of segsInner + segsOuter steps always emits
// exactly that many points, so every (ring, side) walk knows its output offset up front:
//   k_tess_count  one thread per triangle: integer tessellation factors, points per triangle
//   exclusive scan (cub) -> first point of every triangle
//   k_tess_emit   one CTA per triangle, one thread per (ring, side) walk; each step evaluates the sub-triangle's
//                 centroid (or incentre), area, shading normal and uv with the reference's fp32 arithmetic
#include <cub/device/device_scan.cuh>
#include <algorithm>
#include "skc_internal.h"
#include "skc_texture.cuh"

struct TessBC { float b0, b1, b2; };
__host__ __device__ __forceinline__ TessBC bc3(float a, float b, float c) { TessBC r; r.b0 = a; r.b1 = b; r.b2 = c; return r; }
__host__ __device__ __forceinline__ float lerpF(float t, float v1, float v2) { return (1.f - t) * v1 + t * v2; }   // core/pbrt.h:250-252
__host__ __device__ __forceinline__ TessBC bcLerp(float v, TessBC a, TessBC b) { return bc3(lerpF(v, a.b0, b.b0), lerpF(v, a.b1, b.b1), lerpF(v, a.b2, b.b2)); }
__host__ __device__ __forceinline__ f3 bcPoint(TessBC w, f3 v0, f3 v1, f3 v2) {      // b0*v0 + b1*v1 + b2*v2
    return add3(add3(mul3(v0, w.b0), mul3(v1, w.b1)), mul3(v2, w.b2));
}
__host__ __device__ __forceinline__ TessBC bcOfBc(TessBC w, TessBC a, TessBC b, TessBC c) {
    return bc3(w.b0 * a.b0 + w.b1 * b.b0 + w.b2 * c.b0, w.b0 * a.b1 + w.b1 * b.b1 + w.b2 * c.b1, w.b0 * a.b2 + w.b1 * b.b2 + w.b2 * c.b2);
}

struct TessFactors { int e0, e1, e2, c, rings; };

// trianglemesh.cpp:201-214 (rounded factors) + tessellator's integer factors (:266-270)
__host__ __device__ __forceinline__ TessFactors tessFactors(f3 v0, f3 v1, f3 v2, float minDist) {
    float le0 = len3(sub3(v1, v2)), le1 = len3(sub3(v2, v0)), le2 = len3(sub3(v0, v1));
    float tfe0 = le0 / minDist * 0.8f, tfe1 = le1 / minDist * 0.8f, tfe2 = le2 / minDist * 0.8f;
    float tfc = floorf((tfe0 + tfe1 + tfe2) / 3.f + .5f);
    tfe0 = floorf(tfe0 + .5f); tfe1 = floorf(tfe1 + .5f); tfe2 = floorf(tfe2 + .5f);
    TessFactors t;
    auto atLeast = [](int v, int lo) { return v < lo ? lo : v; };
    t.e0 = atLeast((int)ceilf(tfe0), 1); t.e1 = atLeast((int)ceilf(tfe1), 1); t.e2 = atLeast((int)ceilf(tfe2), 1); t.c = atLeast((int)ceilf(tfc), 1);
    if (t.e0 > 1 || t.e1 > 1 || t.e2 > 1) t.c = atLeast(t.c, 2);
    t.rings = (t.c + 1) / 2;
    return t;
}
// points emitted by ring r (all three sides)
__host__ __device__ __forceinline__ uint32_t ringPoints(const TessFactors &t, int r) {
    if (r < t.rings - 1) {
        int edgeInner = t.c - (t.rings - r) * 2;
        return edgeInner >= 0 ? 3u * (uint32_t)(edgeInner + edgeInner + 2) : 1u;
    }
    int edgeInner = t.c - 2;
    return edgeInner >= 0 ? (uint32_t)(3 * edgeInner + t.e0 + t.e1 + t.e2) : 1u;
}

__global__ void __launch_bounds__(256) k_tess_count(const float *__restrict__ P, const int32_t *__restrict__ idx, uint32_t nTris, float minDist,
                                                    uint32_t *__restrict__ counts) {
    uint32_t tri = blockIdx.x * blockDim.x + threadIdx.x;
    if (tri >= nTris) return;
    f3 v0 = ld3(P, idx[3 * tri]), v1 = ld3(P, idx[3 * tri + 1]), v2 = ld3(P, idx[3 * tri + 2]);
    TessFactors t = tessFactors(v0, v1, v2, minDist);
    uint32_t n = 0;
    for (int r = 0; r < t.rings; ++r) n += ringPoints(t, r);
    counts[tri] = n;
}

struct TessEmit {
    MeshView mv; int4 vi; f3 v0, v1, v2;
    int incenter; uint32_t materialId; float rayEpsilon;
    skc_surface_point *out;
    const AlbedoTexDev *bump;           // "texture bumpmap" of the material (device only), or NULL
};

// the domain shader of TessellateSurfacePoints (trianglemesh.cpp:216-255)
template <bool BUMP> __host__ __device__ inline void tessPoint(const TessEmit &e, TessBC bv0, TessBC bv1, TessBC bv2, uint32_t slot) {
    f3 sv0 = bcPoint(bv0, e.v0, e.v1, e.v2), sv1 = bcPoint(bv1, e.v0, e.v1, e.v2), sv2 = bcPoint(bv2, e.v0, e.v1, e.v2);
    TessBC bc;
    if (!e.incenter) bc = bcOfBc(bc3(1.f / 3.f, 1.f / 3.f, 1.f / 3.f), bv0, bv1, bv2);
    else {
        float l0 = len3(sub3(sv1, sv2)), l1 = len3(sub3(sv2, sv0)), l2 = len3(sub3(sv0, sv1));
        bc = bcOfBc(bc3(l0 / (l0 + l1 + l2), l1 / (l0 + l1 + l2), l2 / (l0 + l1 + l2)), bv0, bv1, bv2);
    }
    f3 pp = bcPoint(bc, e.v0, e.v1, e.v2), nn;
    float tu, tv;
    triShadingGeometry(e.mv, e.vi, bc.b0, bc.b1, bc.b2, &nn, &tu, &tv);
#ifdef __CUDA_ARCH__
    if (BUMP) { f3 dpduBump; triBump(e.mv, e.vi, bc.b0, bc.b1, bc.b2, *e.bump, 0.f, 0.f, 0.f, 0.f, &nn, &dpduBump); }    // sp.n = dgBump.nn (trianglemesh.cpp:236-242)
#endif
    float area = .5f * len3(cross3(sub3(sv1, sv0), sub3(sv2, sv0)));
    float *o = reinterpret_cast<float *>(e.out + slot);
    o[0] = pp.x; o[1] = pp.y; o[2] = pp.z; o[3] = nn.x; o[4] = nn.y; o[5] = nn.z; o[6] = tu; o[7] = tv;
    e.out[slot].materialId = e.materialId; o[9] = area; o[10] = e.rayEpsilon;
}

// matching (trianglemesh.cpp:319-351): zipper between an inner edge of segsInner and an outer edge of segsOuter segments
template <bool BUMP> __host__ __device__ inline void tessMatching(const TessEmit &e, TessBC b0Inner, TessBC b1Inner, int segsInner, TessBC b0Outer, TessBC b1Outer, int segsOuter, uint32_t slot) {
    int innerPos = 0, outerPos = 0;
    while (innerPos < segsInner || outerPos < segsOuter) {
        TessBC bInner = segsInner ? bcLerp((float)innerPos / (float)segsInner, b0Inner, b1Inner) : b0Inner;
        TessBC bOuter = bcLerp((float)outerPos / (float)segsOuter, b0Outer, b1Outer);
        float slopeInner = (innerPos < segsInner) ? fabsf((float)(innerPos + 1) + 1.f - (float)outerPos / (float)segsOuter * (float)(segsInner + 2)) : INFINITY;
        float slopeOuter = (outerPos < segsOuter) ? fabsf((float)innerPos + 1.f - (float)(outerPos + 1) / (float)segsOuter * (float)(segsInner + 2)) : INFINITY;
        // one call site for the two cases (a warp whose lanes disagree would otherwise run the domain shader twice): the new vertex
        // is on the inner edge -> sub-triangle (new, inner, outer), else on the outer edge -> (inner, outer, new)
        const bool stepInner = slopeInner < slopeOuter;
        const TessBC bNew = bcLerp(stepInner ? (float)(innerPos + 1) / (float)segsInner : (float)(outerPos + 1) / (float)segsOuter,
                                   stepInner ? b0Inner : b0Outer, stepInner ? b1Inner : b1Outer);
        tessPoint<BUMP>(e, stepInner ? bNew : bInner, stepInner ? bInner : bOuter, stepInner ? bOuter : bNew, slot++);
        if (stepInner) innerPos++; else outerPos++;
    }
}

// One job = one (ring, side) walk of a triangle's tessellation; a ring without an inner edge is a single job (side 0).  Every
// job knows where its points go (see the header comment), so the jobs of a triangle run in any order -- threads of a CTA on
// the device, a plain loop in skc_tessellate_host.
template <bool BUMP> __host__ __device__ inline void tessJob(const TessEmit &e, const TessFactors &t, int job) {
    const TessBC b0 = bc3(1.f, 0.f, 0.f), b1 = bc3(0.f, 1.f, 0.f), b2 = bc3(0.f, 0.f, 1.f), bcC = bc3(1.f / 3.f, 1.f / 3.f, 1.f / 3.f);
    const TessBC corner[3] = {b0, b1, b2};
    int r = job / 3, side = job - 3 * r;
    uint32_t slot = 0;
    for (int q = 0; q < r; ++q) slot += ringPoints(t, q);
    bool last = r == t.rings - 1;
    int edgeInner = last ? t.c - 2 : t.c - (t.rings - r) * 2;
    if (edgeInner < 0) {
        if (side == 0) {
            if (last) tessPoint<BUMP>(e, b0, b1, b2, slot);
            else {
                float f = (float)(r + 1) / (float)t.rings;
                tessPoint<BUMP>(e, bcLerp(f, bcC, b0), bcLerp(f, bcC, b1), bcLerp(f, bcC, b2), slot);
            }
        }
        return;
    }
    // sides in the reference's order: (0,1), (1,2), (2,0); outer segment counts of the last ring: itfe2, itfe0, itfe1
    int a = side, b = (side + 1) % 3;
    float fIn = (float)r / (float)t.rings;
    TessBC inA = bcLerp(fIn, bcC, corner[a]), inB = bcLerp(fIn, bcC, corner[b]);
    TessBC outA, outB; int segsOuter, outerBefore;
    if (last) {
        outA = corner[a]; outB = corner[b];
        const int segs[3] = {t.e2, t.e0, t.e1};
        segsOuter = segs[side];
        outerBefore = side == 0 ? 0 : (side == 1 ? t.e2 : t.e2 + t.e0);
    } else {
        float fOut = (float)(r + 1) / (float)t.rings;
        outA = bcLerp(fOut, bcC, corner[a]); outB = bcLerp(fOut, bcC, corner[b]);
        segsOuter = edgeInner + 2;
        outerBefore = side * segsOuter;
    }
    slot += (uint32_t)(side * edgeInner + outerBefore);
    tessMatching<BUMP>(e, inA, inB, edgeInner, outA, outB, segsOuter, slot);
}

template <bool BUMP> __global__ void __launch_bounds__(128) k_tess_emit(MeshView mv, const int32_t *__restrict__ idx, uint32_t nTris, float minDist, int incenter,
                                                   uint32_t materialId, const uint32_t *__restrict__ offsets, skc_surface_point *__restrict__ out,
                                                   uint32_t rangeLo, uint32_t rangeHi, const __grid_constant__ AlbedoTexDev bump) {
    uint32_t tri = blockIdx.x;
    if (tri >= nTris) return;
    if (offsets[tri + 1] <= rangeLo || offsets[tri] >= rangeHi) return;      // none of this triangle's points is wanted (a shard of the set)
    TessEmit e;
    e.mv = mv; e.vi = make_int4(idx[3 * tri], idx[3 * tri + 1], idx[3 * tri + 2], 0);
    e.v0 = ld3(mv.P, e.vi.x); e.v1 = ld3(mv.P, e.vi.y); e.v2 = ld3(mv.P, e.vi.z);
    e.incenter = incenter; e.materialId = materialId; e.rayEpsilon = minDist / 10.f; e.out = out + offsets[tri];
    e.bump = BUMP ? &bump : nullptr;
    TessFactors t = tessFactors(e.v0, e.v1, e.v2, minDist);
    for (int job = threadIdx.x; job < 3 * t.rings; job += blockDim.x) tessJob<BUMP>(e, t, job);
}

// GetSurfacePointsThroughTessellation for the uploaded mesh, written into the resident surface-point set
static int tessellate(skc_ctx *ctx, float minDist, int incenter, uint32_t materialId, uint64_t first, uint64_t count, bool ranged, uint64_t *countOut) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!ctx->meshIdx || !ctx->meshTris || !ctx->meshP) return skc_fail(ctx, "skc_tessellate_resident: no mesh uploaded (skc_upload_bvh)");
    if (!(minDist > 0.f)) return skc_fail(ctx, "skc_tessellate_resident: minSampleDist must be positive");
    uint32_t nTris = ctx->meshTris;
    size_t tempBytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tempBytes, (uint32_t *)nullptr, (uint32_t *)nullptr, (int)nTris + 1, ctx->stream);
    size_t cntBytes = ((size_t)(nTris + 1) * 4 + 255) / 256 * 256;
    char *buf = (char *)skc_scratch(ctx, 11, 2 * cntBytes + tempBytes + 256);
    if (!buf) return 1;
    uint32_t *counts = (uint32_t *)buf, *offsets = (uint32_t *)(buf + cntBytes);
    void *temp = buf + 2 * cntBytes;
    skc_stage_begin(ctx);
    SKC_CUDA(cudaMemsetAsync(counts + nTris, 0, sizeof(uint32_t), ctx->stream));
    k_tess_count<<<(nTris + 255) / 256, 256, 0, ctx->stream>>>(ctx->meshP, ctx->meshIdx, nTris, minDist, counts);
    SKC_LAUNCHED("k_tess_count");
    SKC_CUDA(cub::DeviceScan::ExclusiveSum(temp, tempBytes, counts, offsets, (int)nTris + 1, ctx->stream));
    ctx->launches += 2;
    uint32_t total = 0;
    SKC_CUDA(cudaMemcpyAsync(&total, offsets + nTris, sizeof total, cudaMemcpyDeviceToHost, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    if (total > 0x7fffffffu) return skc_fail(ctx, "skc_tessellate_resident: %u surface points exceed the reference's int indexing", total);
    if (ctx->resPtsCap < total) {
        if (ctx->resPts) cudaFree(ctx->resPts);
        ctx->resPts = nullptr; ctx->resPtsCap = 0; ctx->resPtsCount = 0;
        SKC_CUDA(cudaMalloc((void **)&ctx->resPts, (size_t)(total ? total : 1) * sizeof(skc_surface_point)));
        ctx->resPtsCap = total;
    }
    if (total) {
        uint32_t lo = ranged ? (uint32_t)std::min<uint64_t>(first, total) : 0u, hi = ranged ? (uint32_t)std::min<uint64_t>(first + count, total) : total;
        if (ctx->bumpTex.rgb) k_tess_emit<true><<<nTris, 128, 0, ctx->stream>>>(skc_mesh_view(ctx), ctx->meshIdx, nTris, minDist, incenter, materialId, offsets, ctx->resPts, lo, hi, ctx->bumpTex);
        else k_tess_emit<false><<<nTris, 128, 0, ctx->stream>>>(skc_mesh_view(ctx), ctx->meshIdx, nTris, minDist, incenter, materialId, offsets, ctx->resPts, lo, hi, ctx->bumpTex);
        SKC_LAUNCHED("k_tess_emit");
    }
    skc_stage_end(ctx, 8);
    ctx->resPtsCount = total;
    if (countOut) *countOut = total;
    return 0;
}

extern "C" int skc_tessellate_resident(skc_ctx *ctx, float minDist, int incenter, uint32_t materialId, uint64_t *countOut) {
    return tessellate(ctx, minDist, incenter, materialId, 0, 0, false, countOut);
}
extern "C" int skc_tessellate_range(skc_ctx *ctx, float minDist, int incenter, uint32_t materialId, uint64_t first, uint64_t count,
                                    uint64_t *countOut) {
    return tessellate(ctx, minDist, incenter, materialId, first, count, true, countOut);
}

// copy of resident surface points [first, first+count) for the host (tests, "pointsfile" export)
extern "C" int skc_get_surface_points(skc_ctx *ctx, uint64_t first, uint64_t count, skc_surface_point *out) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (first + count > ctx->resPtsCount) return skc_fail(ctx, "skc_get_surface_points: bad range");
    if (!count) return 0;
    SKC_CUDA(cudaMemcpyAsync(out, ctx->resPts + first, count * sizeof(skc_surface_point), cudaMemcpyDeviceToHost, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// The same jobs evaluated on the HOST (no context, no GPU): the host mirror's TriangleMesh::TessellateSurfacePoints /
// CountSurfacePoints (host/skin_geom.cpp) call this, so that host and device share one tessellator -- the CPU tests that
// pin the host mirror to the reference's exported points (tests/test_host_mirror.py) pin the device kernel's job functions
// at the same time.  Not a fallback: RenderScene and bench.py generate their points with k_tess_emit.
//   P [nVerts][3] world-space points, N object-space normals or NULL, UV or NULL, idx [nTris][3], o2w32 = {ObjectToWorld.m,
//   ObjectToWorld.mInv} row-major, flip = ReverseOrientation ^ TransformSwapsHandedness.
// out == NULL: count only.  Returns the number of points (*countOut); fails (1) if cap is too small.
extern "C" int skc_tessellate_host(const float *P, const float *N, const float *UV, const int32_t *idx, uint32_t nTris, const float *o2w32, int flip,
                                   float minDist, int incenter, uint32_t materialId, skc_surface_point *out, uint64_t cap, uint64_t *countOut) {
    if (!P || !idx || !(minDist > 0.f)) return 1;
    MeshView mv;
    mv.primInfo = nullptr; mv.P = P; mv.N = N; mv.UV = UV; mv.flip = flip;
    for (int i = 0; i < 16; ++i) { mv.o2w[i] = o2w32 ? o2w32[i] : (i % 5 == 0 ? 1.f : 0.f); mv.w2o[i] = o2w32 ? o2w32[16 + i] : (i % 5 == 0 ? 1.f : 0.f); }
    uint64_t total = 0;
    for (uint32_t tri = 0; tri < nTris; ++tri) {
        TessEmit e;
        e.mv = mv; e.vi = make_int4(idx[3 * tri], idx[3 * tri + 1], idx[3 * tri + 2], 0);
        e.v0 = ld3(P, e.vi.x); e.v1 = ld3(P, e.vi.y); e.v2 = ld3(P, e.vi.z);
        e.incenter = incenter; e.materialId = materialId; e.rayEpsilon = minDist / 10.f; e.bump = nullptr;
        const TessFactors t = tessFactors(e.v0, e.v1, e.v2, minDist);
        uint64_t n = 0;
        for (int r = 0; r < t.rings; ++r) n += ringPoints(t, r);
        if (out) {
            if (total + n > cap) return 1;
            e.out = out + total;
            for (int job = 0; job < 3 * t.rings; ++job) tessJob<false>(e, t, job);
        }
        total += n;
    }
    if (countOut) *countOut = total;
    return 0;
}

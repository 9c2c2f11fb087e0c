// skc_bvh.cu -- BVHAccel on the device: upload of the flat tree and the two ray kernels.
// Reference: accelerators/bvh.cpp (LinearBVHNode :113-123, Intersect :388-439,
// IntersectP :442-488), core/primitive.cpp:180-193, shapes/trianglemesh.inl:55-136,224-299.
#include <vector>
#include <string.h>
#include "skc_light.cuh"
#include "skc_internal.h"

// ---- kernels --------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_trace_any(BvhView bv, const skc_ray *__restrict__ rays, uint64_t n,
                                                  uint8_t *__restrict__ out, unsigned long long *cnt) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 *rp = reinterpret_cast<const float4 *>(rays + i);
    float4 r0 = __ldg(rp), r1 = __ldg(rp + 1);
    DRay ray; ray.o = mk3(r0.x, r0.y, r0.z); ray.d = mk3(r0.w, r1.x, r1.y); ray.mint = r1.z; ray.maxt = r1.w;
    bool occ = cnt ? bvhOccluded<true>(bv, ray, cnt) : bvhOccludedFast(bv, ray);
    out[i] = occ ? 1 : 0;
}

__device__ inline void triShading(const MeshView &mv, int prim, const DRay &ray, float t, float b1, float b2,
                                  f3 *pOut, f3 *nOut, f3 *ngOut = nullptr, f3 *ssOut = nullptr) {
    int4 vi = __ldg(&mv.primInfo[prim]);
    float b0 = 1 - b1 - b2, tu, tv;
    *pOut = add3(ray.o, mul3(ray.d, t));
    triShadingGeometry(mv, vi, b0, b1, b2, nOut, &tu, &tv, ngOut, ssOut);
}

__global__ void __launch_bounds__(128) k_trace_closest(BvhView bv, MeshView mv, const skc_ray *__restrict__ rays,
                                                      uint64_t n, skc_hit *__restrict__ out, unsigned long long *cnt,
                                                      float4 *__restrict__ frame) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 *rp = reinterpret_cast<const float4 *>(rays + i);
    float4 r0 = __ldg(rp), r1 = __ldg(rp + 1);
    DRay ray; ray.o = mk3(r0.x, r0.y, r0.z); ray.d = mk3(r0.w, r1.x, r1.y); ray.mint = r1.z; ray.maxt = r1.w;
    float b1 = 0.f, b2 = 0.f;
    int prim = cnt ? bvhClosest<true>(bv, ray, &b1, &b2, cnt) : bvhClosest<false>(bv, ray, &b1, &b2, nullptr);
    skc_hit h;
    h.hit = prim >= 0; h.prim = prim; h.t = 0.f; h.rayEpsilon = 0.f;
    h.p[0] = h.p[1] = h.p[2] = 0.f; h.n[0] = h.n[1] = h.n[2] = 0.f;
    h.wo[0] = h.wo[1] = h.wo[2] = 0.f; h.hasBssrdf = 0;
    if (prim >= 0) {
        float t = ray.maxt;
        h.t = t;
        h.wo[0] = -ray.d.x; h.wo[1] = -ray.d.y; h.wo[2] = -ray.d.z;
        int type = __ldg(&mv.primInfo[prim]).w;
        if (type == 0) {
            f3 p, nn, ng, ss;
            triShading(mv, prim, ray, t, b1, b2, &p, &nn, frame ? &ng : nullptr, frame ? &ss : nullptr);
            if (frame) { frame[2 * i] = make_float4(ng.x, ng.y, ng.z, ss.x); frame[2 * i + 1] = make_float4(ss.y, ss.z, 0.f, 0.f); }
            h.p[0] = p.x; h.p[1] = p.y; h.p[2] = p.z; h.n[0] = nn.x; h.n[1] = nn.y; h.n[2] = nn.z;
            h.rayEpsilon = 1e-3f * t;        // trianglemesh.inl:134
            h.hasBssrdf = 1;
        } else {
            f3 p = add3(ray.o, mul3(ray.d, t));
            h.p[0] = p.x; h.p[1] = p.y; h.p[2] = p.z;
            h.rayEpsilon = 5e-4f * t;        // shapes/sphere.cpp:156
            if (frame) {   // the emitter's dg (shapes/sphere.cpp:103-150): geometric = shading frame of its matte BSDF
                const skc_light &sp = bv.spheres[__float_as_int(__ldg(&bv.prims[3 * prim + 2]).z)];
                DRay r0; r0.o = ray.o; r0.d = ray.d; r0.mint = r1.z; r0.maxt = r1.w;
                float ts; f3 nn, dpdu;
                if (sphereIntersectDg(sp, r0, &ts, &nn, &dpdu)) {
                    if (sp.reverseOrientation != 0.f) nn = mul3(nn, -1.f);
                    h.n[0] = nn.x; h.n[1] = nn.y; h.n[2] = nn.z;
                    frame[2 * i] = make_float4(nn.x, nn.y, nn.z, dpdu.x); frame[2 * i + 1] = make_float4(dpdu.y, dpdu.z, 0.f, 0.f);
                }
            }
        }
    }
    // 56 B record: write as 14 words
    uint32_t *dst = reinterpret_cast<uint32_t *>(out + i);
    const uint32_t *src = reinterpret_cast<const uint32_t *>(&h);
#pragma unroll
    for (int k = 0; k < 14; ++k) dst[k] = src[k];
}

// ---- host side --------------------------------------------------------------------------
BvhView skc_bvh_view(const skc_ctx *ctx) {
    BvhView bv; bv.nodes = ctx->bvhNodes; bv.prims = ctx->bvhPrims; bv.spheres = ctx->spheres; bv.wide = ctx->bvhWideOn ? ctx->bvhWide : nullptr;
    bv.alphaTex = ctx->alphaTex; bv.alphaUV = ctx->alphaUV; bv.alphaW = ctx->alphaW; bv.alphaH = ctx->alphaH;
    bv.alphaSu = ctx->alphaMap[0]; bv.alphaSv = ctx->alphaMap[1]; bv.alphaDu = ctx->alphaMap[2]; bv.alphaDv = ctx->alphaMap[3];
    return bv;
}
// Triangle::GetUVs (shapes/trianglemesh.h:84-97) of every BVH primitive: the mesh's uvs, or (0,0) (1,0) (1,1)
__global__ void k_alpha_uv(const int4 *__restrict__ primInfo, const float *__restrict__ UV, uint32_t nPrims, float *__restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nPrims) return;
    const int4 v = primInfo[i];
    float *o = out + 6 * (size_t)i;
    if (UV && v.w == 0) { o[0] = UV[2 * v.x]; o[1] = UV[2 * v.x + 1]; o[2] = UV[2 * v.y]; o[3] = UV[2 * v.y + 1]; o[4] = UV[2 * v.z]; o[5] = UV[2 * v.z + 1]; }
    else { o[0] = 0.f; o[1] = 0.f; o[2] = 1.f; o[3] = 0.f; o[4] = 1.f; o[5] = 1.f; }
}

// The mesh's "texture alpha" (a float imagemap: shapes/trianglemesh.cpp:411-421, textures/imagemap.cpp): level-0 texels [h][w] as
// ImageTexture<float, float> holds them (scale * (powf(y, gamma) + shift), made by the caller), and the UVMapping2D.  Call after
// skc_upload_bvh (the primitives' uvs are taken from the uploaded mesh).  w == 0 removes it.
extern "C" int skc_upload_alpha(skc_ctx *ctx, uint32_t w, uint32_t h, const float *texels, float su, float sv, float du, float dv) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->alphaTex) { cudaFree(ctx->alphaTex); ctx->alphaTex = nullptr; }
    if (ctx->alphaUV) { cudaFree(ctx->alphaUV); ctx->alphaUV = nullptr; }
    ctx->alphaW = ctx->alphaH = 0;
    if (!w || !h) return 0;
    if (!texels) return skc_fail(ctx, "skc_upload_alpha: null texels");
    if (!ctx->primInfo || !ctx->nPrims) return skc_fail(ctx, "skc_upload_alpha: upload the BVH and the mesh first (skc_upload_bvh)");
    SKC_CUDA(cudaMalloc((void **)&ctx->alphaTex, (size_t)w * h * sizeof(float)));
    SKC_CUDA(cudaMalloc((void **)&ctx->alphaUV, (size_t)ctx->nPrims * 6 * sizeof(float)));
    SKC_CUDA(cudaMemcpyAsync(ctx->alphaTex, texels, (size_t)w * h * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    k_alpha_uv<<<(ctx->nPrims + 255) / 256, 256, 0, ctx->stream>>>(ctx->primInfo, ctx->meshUV, ctx->nPrims, ctx->alphaUV);
    SKC_LAUNCHED("k_alpha_uv");
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->alphaW = (int)w; ctx->alphaH = (int)h;
    ctx->alphaMap[0] = su; ctx->alphaMap[1] = sv; ctx->alphaMap[2] = du; ctx->alphaMap[3] = dv;
    return 0;
}

MeshView skc_mesh_view(const skc_ctx *ctx) {
    MeshView mv; mv.primInfo = ctx->primInfo; mv.P = ctx->meshP; mv.N = ctx->meshN; mv.UV = ctx->meshUV;
    memcpy(mv.o2w, ctx->meshO2W, 64); memcpy(mv.w2o, ctx->meshO2W + 16, 64); mv.flip = ctx->meshFlip;
    return mv;
}

// A scene is uploaded once per frame in the host-buffer flow: the device arrays are kept and reused while they are large
// enough (cudaFree / cudaMalloc synchronise the device), the copies are stream-ordered and the caller synchronises once.
template <class T> static int upload(skc_ctx *ctx, T **dst, const T *src, size_t n) {
    size_t &cap = ctx->uploadCap[(void *)dst];
    if (!n || !src) {
        if (*dst) { cudaFree(*dst); *dst = nullptr; }
        cap = 0;
        return 0;
    }
    if (!*dst || cap < n * sizeof(T)) {
        if (*dst) { cudaFree(*dst); *dst = nullptr; }
        cap = 0;
        SKC_CUDA(cudaMalloc((void **)dst, n * sizeof(T)));
        cap = n * sizeof(T);
    }
    SKC_CUDA(cudaMemcpyAsync(*dst, src, n * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));     // pageable source: staged before the call returns
    return 0;
}

// The 4-wide form of the binary tree for any-hit rays (skc_bvh.cuh bvhOccludedWide): every wide node gathers up to four
// descendants of one binary node -- its two children, then, while fewer than four, the interior child with the largest surface
// area is replaced by its own two children -- and stores THEIR boxes (the reference's boxes, unchanged).  Returns false when a
// leaf does not fit the child word (more than 127 primitives or 2^24 primitives in all): the binary walk is used then.
static bool buildWide(const skc_bvh_node *nodes, uint32_t nNodes, uint32_t nPrims, std::vector<float4> &out) {
    if (nPrims >= (1u << 24)) return false;
    struct Item { uint32_t bin, wide; };
    std::vector<Item> todo;
    auto area = [&](uint32_t b) { const skc_bvh_node &n = nodes[b]; float dx = n.pmax[0] - n.pmin[0], dy = n.pmax[1] - n.pmin[1], dz = n.pmax[2] - n.pmin[2]; return dx * dy + dx * dz + dy * dz; };
    out.clear();
    out.resize(8);
    todo.push_back({0u, 0u});
    while (!todo.empty()) {
        const Item it = todo.back();
        todo.pop_back();
        uint32_t kids[4]; int nk = 0;
        if (nodes[it.bin].nPrimitives) kids[nk++] = it.bin;                     // a tree that is one leaf
        else {
            kids[nk++] = it.bin + 1; kids[nk++] = nodes[it.bin].offset;
            while (nk < 4) {
                int best = -1; float bestArea = -1.f;
                for (int k = 0; k < nk; ++k) if (!nodes[kids[k]].nPrimitives && area(kids[k]) > bestArea) { bestArea = area(kids[k]); best = k; }
                if (best < 0) break;
                const uint32_t b = kids[best];
                kids[best] = b + 1; kids[nk++] = nodes[b].offset;
            }
        }
        float box[6][4]; uint32_t word[4];
        for (int k = 0; k < 4; ++k) {
            if (k >= nk) { for (int a = 0; a < 3; ++a) { box[a][k] = INFINITY; box[3 + a][k] = -INFINITY; } word[k] = SKC_WIDE_EMPTY; continue; }
            const skc_bvh_node &n = nodes[kids[k]];
            for (int a = 0; a < 3; ++a) { box[a][k] = n.pmin[a]; box[3 + a][k] = n.pmax[a]; }
            if (n.nPrimitives) {
                if (n.nPrimitives > 127) return false;
                word[k] = SKC_WIDE_LEAF | ((uint32_t)n.nPrimitives << 24) | n.offset;
            } else {
                const uint32_t w = (uint32_t)(out.size() / 8);
                if (w >= SKC_WIDE_LEAF) return false;
                out.resize(out.size() + 8);
                word[k] = w;
                todo.push_back({kids[k], w});
            }
        }
        float4 *q = out.data() + 8 * (size_t)it.wide;
        for (int a = 0; a < 6; ++a) q[a] = make_float4(box[a][0], box[a][1], box[a][2], box[a][3]);
        float f[4]; memcpy(f, word, 16);
        q[6] = make_float4(f[0], f[1], f[2], f[3]);
        q[7] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    (void)nNodes;
    return true;
}

extern "C" int skc_upload_bvh(skc_ctx *ctx, const skc_bvh_node *nodes, uint32_t nNodes,
                              const skc_prim *prims, uint32_t nPrims,
                              const float *P, const float *N, const float *uv, uint32_t nVerts,
                              const int32_t *indices, uint32_t nTris, const float *meshO2W,
                              int reverseOrientation, int transformSwapsHandedness,
                              const skc_light *spheres, uint32_t nSpheres) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!nodes || !nNodes || !prims || !nPrims) return skc_fail(ctx, "skc_upload_bvh: empty tree");
    cudaStreamSynchronize(ctx->prefetchStream);     // a prefetch still walking the old tree (skc_frame_prefetch) finishes first
    std::vector<float4> dn(2 * (size_t)nNodes), dp(3 * (size_t)nPrims);
    std::vector<int4> info(nPrims);
    for (uint32_t i = 0; i < nNodes; ++i) {
        const skc_bvh_node &nd = nodes[i];
        uint32_t meta = nd.nPrimitives | ((uint32_t)nd.axis << 8);
        float fo, fm; memcpy(&fo, &nd.offset, 4); memcpy(&fm, &meta, 4);
        dn[2 * i] = make_float4(nd.pmin[0], nd.pmin[1], nd.pmin[2], nd.pmax[0]);
        dn[2 * i + 1] = make_float4(nd.pmax[1], nd.pmax[2], fo, fm);
        if (nd.nPrimitives && nd.offset + nd.nPrimitives > nPrims) return skc_fail(ctx, "skc_upload_bvh: leaf %u out of range", i);
        if (!nd.nPrimitives && nd.offset >= nNodes) return skc_fail(ctx, "skc_upload_bvh: child of node %u out of range", i);
    }
    for (uint32_t i = 0; i < nPrims; ++i) {
        const skc_prim &pr = prims[i];
        float ft, fi; int32_t ty = pr.type, ix = pr.index; memcpy(&ft, &ty, 4); memcpy(&fi, &ix, 4);
        if (pr.type == 0) {
            if (!P || !indices || (uint32_t)pr.index >= nTris) return skc_fail(ctx, "skc_upload_bvh: triangle %u out of range", i);
            int v0 = indices[3 * pr.index], v1 = indices[3 * pr.index + 1], v2 = indices[3 * pr.index + 2];
            if ((uint32_t)v0 >= nVerts || (uint32_t)v1 >= nVerts || (uint32_t)v2 >= nVerts) return skc_fail(ctx, "skc_upload_bvh: vertex index out of range");
            const float *p1 = P + 3 * v0, *p2 = P + 3 * v1, *p3 = P + 3 * v2;
            float e1[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
            float e2[3] = {p3[0] - p1[0], p3[1] - p1[1], p3[2] - p1[2]};
            dp[3 * i] = make_float4(p1[0], p1[1], p1[2], e1[0]);
            dp[3 * i + 1] = make_float4(e1[1], e1[2], e2[0], e2[1]);
            dp[3 * i + 2] = make_float4(e2[2], ft, fi, 0.f);
            info[i] = make_int4(v0, v1, v2, 0);
        } else if (pr.type == 1) {
            if ((uint32_t)pr.index >= nSpheres) return skc_fail(ctx, "skc_upload_bvh: sphere %u out of range", i);
            dp[3 * i] = make_float4(0, 0, 0, 0); dp[3 * i + 1] = make_float4(0, 0, 0, 0);
            dp[3 * i + 2] = make_float4(0.f, ft, fi, 0.f);
            info[i] = make_int4(0, 0, 0, 1);
        } else return skc_fail(ctx, "skc_upload_bvh: unsupported primitive type %d", pr.type);
    }
    {
        std::vector<float4> wide;
        if (buildWide(nodes, nNodes, nPrims, wide)) { if (upload(ctx, &ctx->bvhWide, wide.data(), wide.size())) return 1; ctx->nWideNodes = (uint32_t)(wide.size() / 8); }
        else { if (upload(ctx, &ctx->bvhWide, (const float4 *)nullptr, 0)) return 1; ctx->nWideNodes = 0; }
    }
    if (upload(ctx, &ctx->bvhNodes, dn.data(), dn.size())) return 1;
    if (upload(ctx, &ctx->bvhPrims, dp.data(), dp.size())) return 1;
    if (upload(ctx, &ctx->primInfo, info.data(), info.size())) return 1;
    if (upload(ctx, &ctx->meshP, P, 3 * (size_t)nVerts)) return 1;
    if (upload(ctx, &ctx->meshN, N, N ? 3 * (size_t)nVerts : 0)) return 1;
    if (upload(ctx, &ctx->meshUV, uv, uv ? 2 * (size_t)nVerts : 0)) return 1;
    if (upload(ctx, &ctx->meshIdx, indices, indices ? 3 * (size_t)nTris : 0)) return 1;
    ctx->meshTris = indices ? nTris : 0;
    if (upload(ctx, &ctx->spheres, spheres, nSpheres)) return 1;
    ctx->nNodes = nNodes; ctx->nPrims = nPrims; ctx->nVerts = nVerts; ctx->nSpheres = nSpheres;
    {   // longest root-to-leaf path of the depth-first array: the most entries a traversal stack can hold
        std::vector<std::pair<uint32_t, uint32_t>> st; st.push_back({0u, 1u});
        uint32_t depth = 0;
        while (!st.empty()) {
            auto [n, d] = st.back(); st.pop_back();
            depth = std::max(depth, d);
            if (n < nNodes && nodes[n].nPrimitives == 0) { st.push_back({n + 1, d + 1}); if (nodes[n].offset < nNodes) st.push_back({nodes[n].offset, d + 1}); }
        }
        ctx->bvhDepth = depth;
    }
    if (meshO2W) memcpy(ctx->meshO2W, meshO2W, 128);
    else { float id[32] = {1,0,0,0, 0,1,0,0, 0,0,1,0, 0,0,0,1, 1,0,0,0, 0,1,0,0, 0,0,1,0, 0,0,0,1}; memcpy(ctx->meshO2W, id, 128); }
    ctx->meshFlip = (reverseOrientation != 0) ^ (transformSwapsHandedness != 0);
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int skc_trace_any_dev(skc_ctx *ctx, const skc_ray *raysDev, uint64_t n, uint8_t *outDev, unsigned long long *cntDev) {
    if (!ctx->bvhNodes) return skc_fail(ctx, "skc_trace_any: no BVH uploaded");
    if (!n) return 0;
    uint64_t blocks = (n + 127) / 128;
    k_trace_any<<<(unsigned)blocks, 128, 0, ctx->stream>>>(skc_bvh_view(ctx), raysDev, n, outDev, cntDev);
    SKC_LAUNCHED("k_trace_any");
    return 0;
}

int skc_trace_closest_dev(skc_ctx *ctx, const skc_ray *raysDev, uint64_t n, skc_hit *outDev, unsigned long long *cntDev, float4 *frameDev) {
    if (!ctx->bvhNodes) return skc_fail(ctx, "skc_trace_closest: no BVH uploaded");
    if (!n) return 0;
    uint64_t blocks = (n + 127) / 128;
    k_trace_closest<<<(unsigned)blocks, 128, 0, ctx->stream>>>(skc_bvh_view(ctx), skc_mesh_view(ctx), raysDev, n, outDev, cntDev, frameDev);
    SKC_LAUNCHED("k_trace_closest");
    return 0;
}

extern "C" int skc_trace_any(skc_ctx *ctx, const skc_ray *rays, uint64_t n, uint8_t *out) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!n) return 0;
    skc_ray *d = (skc_ray *)skc_scratch(ctx, 0, n * sizeof(skc_ray));
    uint8_t *o = (uint8_t *)skc_scratch(ctx, 1, n);
    if (!d || !o) return 1;
    SKC_CUDA(cudaMemcpyAsync(d, rays, n * sizeof(skc_ray), cudaMemcpyHostToDevice, ctx->stream));
    skc_stage_begin(ctx);
    if (skc_trace_any_dev(ctx, d, n, o, nullptr)) return 1;
    skc_stage_end(ctx, 5);
    SKC_CUDA(cudaMemcpyAsync(out, o, n, cudaMemcpyDeviceToHost, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

extern "C" int skc_trace_closest(skc_ctx *ctx, const skc_ray *rays, uint64_t n, skc_hit *out) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!n) return 0;
    skc_ray *d = (skc_ray *)skc_scratch(ctx, 0, n * sizeof(skc_ray));
    skc_hit *o = (skc_hit *)skc_scratch(ctx, 1, n * sizeof(skc_hit));
    if (!d || !o) return 1;
    SKC_CUDA(cudaMemcpyAsync(d, rays, n * sizeof(skc_ray), cudaMemcpyHostToDevice, ctx->stream));
    skc_stage_begin(ctx);
    if (skc_trace_closest_dev(ctx, d, n, o, nullptr)) return 1;
    skc_stage_end(ctx, 3);
    SKC_CUDA(cudaMemcpyAsync(out, o, n * sizeof(skc_hit), cudaMemcpyDeviceToHost, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// skc_poisson.cu -- the Poisson-disk surface-point finder (renderers/surfacepoints.cpp:103-284, FindPoissonPointDistribution
// :354-359): MultipoleSubsurfaceIntegrator's alternative to the tessellator ("bool usepoissonpointfinder").
//
// The reference runs one SurfacePointTask per core; with --ncores 1 (the only deterministic configuration: the tasks race for
// the octree otherwise) task 0 draws everything from RNG(0):
//   round:  20 000 random walks from the camera position, 30 bounces each, every bounce 2 floats (UniformSampleSphere), the
//           scene's closest hit or -- outside it -- a hit on the scene's bounding sphere; from bounce 3 on, every hit on a
//           translucent primitive leaves a candidate {p, shading normal, u, v, area = pi (d/2)^2, rayEpsilon};
//   then    the candidates, in order, are accepted unless an accepted point lies within minSampleDist; the finder stops after
//           maxFails rejections in a row.
// The walks of a round are independent once every walk knows where its floats start in the stream: 2 + 2 * 30 words per walk,
// unless a ray misses the bounding sphere too and the walk ends early.  So the words of a round are made on the host, the walks
// run one per thread (k_poisson_walk: bvhClosest, the emitters' and the bounding sphere's Sphere::Intersect, glibc's sinf / cosf
// restated), every walk reports how many words it consumed, and the round is repeated with corrected offsets in the rare case
// that one ended early.  The accept / reject chain is inherently serial and cheap (a hash grid of cell size minSampleDist in
// place of the reference's octree: the same predicate "some accepted point with DistanceSquared < d^2"); it runs on the host.
#include <algorithm>
#include <unordered_map>
#include <vector>
#include "skc_light.cuh"
#include "skc_texture.cuh"
#include "skc_internal.h"

#define PSN_PATHS 20000           // surfacepoints.cpp:182
#define PSN_DEPTH 30              // :187
#define PSN_FIRST 3               // candidates from ray.depth >= 3 (:201)
#define PSN_SLOTS (PSN_DEPTH - PSN_FIRST)

namespace {

// core/rng.cpp: MT19937 (init_genrand / genrand_int32), the reference's RNG
struct HostMT {
    uint32_t mt[624]; int mti;
    explicit HostMT(uint32_t seed) { mt[0] = seed; for (mti = 1; mti < 624; ++mti) mt[mti] = 1812433253u * (mt[mti - 1] ^ (mt[mti - 1] >> 30)) + (uint32_t)mti; }
    uint32_t next() {
        if (mti >= 624) {
            int kk;
            for (kk = 0; kk < 624 - 397; ++kk) { uint32_t y = (mt[kk] & 0x80000000u) | (mt[kk + 1] & 0x7fffffffu); mt[kk] = mt[kk + 397] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u); }
            for (; kk < 623; ++kk) { uint32_t y = (mt[kk] & 0x80000000u) | (mt[kk + 1] & 0x7fffffffu); mt[kk] = mt[kk + (397 - 624)] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u); }
            uint32_t y = (mt[623] & 0x80000000u) | (mt[0] & 0x7fffffffu); mt[623] = mt[396] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
            mti = 0;
        }
        uint32_t y = mt[mti++];
        y ^= (y >> 11); y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= (y >> 18);
        return y;
    }
};

}  // namespace

// UniformSampleSphere(rng.RandomFloat(), rng.RandomFloat()) (core/montecarlo.cpp:283-290): g++ evaluates the two arguments right
// to left, so the FIRST word drawn is u2
__device__ __forceinline__ f3 psnSampleSphere(const uint32_t *__restrict__ words, uint64_t &w) {
    const float u2 = (float)(__ldg(&words[w]) & 0xffffffu) / (float)(1 << 24);
    const float u1 = (float)(__ldg(&words[w + 1]) & 0xffffffu) / (float)(1 << 24);
    w += 2;
    const float z = 1.f - 2.f * u1;
    const float r = sqrtf(fmaxf(0.f, 1.f - z * z));
    const float phi = (2.f * SKC_PI_F) * u2;
    return mk3(r * skc_cosf(phi), r * skc_sinf(phi), z);
}
__device__ __forceinline__ f3 psnFaceforward(f3 v, f3 n) { return dot3(v, n) < 0.f ? mul3(v, -1.f) : v; }

// dg.p of Sphere::Intersect (shapes/sphere.cpp:103-150): ObjectToWorld(phit), phit from the object-space ray
__device__ inline f3 psnSpherePoint(const skc_light &s, const DRay &r, float thit) {
    const f3 o = xformPoint(s.w2o, r.o), d = xformVector(s.w2o, r.d);
    f3 phit = add3(o, mul3(d, thit));
    if (phit.x == 0.f && phit.y == 0.f) phit.x = 1e-5f * s.radius;
    return xformPoint(s.o2w, phit);
}

__global__ void __launch_bounds__(128) k_poisson_walk(BvhView bv, MeshView mv, const __grid_constant__ skc_light bound, const __grid_constant__ AlbedoTexDev bump,
                                                      f3 origin, const uint32_t *__restrict__ words, const uint64_t *__restrict__ wordOffset, uint32_t nPaths,
                                                      float area, uint32_t materialId, skc_surface_point *__restrict__ cand, uint32_t *__restrict__ candMask,
                                                      uint32_t *__restrict__ used, uint32_t *__restrict__ raysTraced) {
    const uint32_t path = blockIdx.x * blockDim.x + threadIdx.x;
    if (path >= nPaths) return;
    uint64_t w = wordOffset[path];
    const uint64_t w0 = w;
    uint32_t mask = 0, nRays = 0;
    DRay ray; ray.o = origin; ray.d = psnSampleSphere(words, w); ray.mint = 0.f; ray.maxt = INFINITY;
    for (int depth = 0; depth < PSN_DEPTH; ++depth) {
        ++nRays;
        DRay walk = ray;                               // bvhClosest shrinks maxt
        float b1 = 0.f, b2 = 0.f;
        const int prim = bvhClosest<false>(bv, walk, &b1, &b2, nullptr);
        f3 p, ng;
        float rayEps;
        if (prim >= 0) {
            const float t = walk.maxt;
            if (__ldg(&mv.primInfo[prim]).w == 0) {    // a triangle of the skin mesh: isect.GetBSSRDF != NULL (materials/layeredskin.cpp:170-177)
                f3 ns, ss;
                const int4 vi = __ldg(&mv.primInfo[prim]);
                float tu, tv;
                p = add3(ray.o, mul3(ray.d, t));                          // dg.p = ray(thit), shapes/trianglemesh.inl:128
                triShadingGeometry(mv, vi, 1 - b1 - b2, b1, b2, &ns, &tu, &tv, &ng, &ss);
                rayEps = 1e-3f * t;
                ng = psnFaceforward(ng, mul3(ray.d, -1.f));             // hitGeometry.nn = Faceforward(nn, -ray.d), :196-197
                if (depth >= PSN_FIRST) {
                    if (bump.rgb) {                    // mat->GetBumpMapping().Bump(hitGeometry, dgShading, &dgs): Faceforward to the flipped geometric normal
                        f3 nb, dpduB;
                        triBump(mv, vi, 1 - b1 - b2, b1, b2, bump, 0.f, 0.f, 0.f, 0.f, &nb, &dpduB);
                        ns = psnFaceforward(nb, ng);      // triBump oriented it towards dgGeom.nn; the reference's dgGeom here is hitGeometry, whose nn was flipped above
                    }
                    float *o = reinterpret_cast<float *>(cand + (size_t)path * PSN_SLOTS + (depth - PSN_FIRST));
                    o[0] = p.x; o[1] = p.y; o[2] = p.z; o[3] = ns.x; o[4] = ns.y; o[5] = ns.z; o[6] = tu; o[7] = tv;
                    reinterpret_cast<uint32_t *>(o)[8] = materialId; o[9] = area; o[10] = rayEps;
                    mask |= 1u << (depth - PSN_FIRST);
                }
            } else {                                   // an emitter sphere (matte: no BSSRDF)
                const skc_light &sp = bv.spheres[__float_as_int(__ldg(&bv.prims[3 * prim + 2]).z)];
                float ts = t; f3 nn = mk3(0.f, 0.f, 1.f);
                sphereIntersectDg(sp, ray, &ts, &nn);
                if (sp.reverseOrientation != 0.f) nn = mul3(nn, -1.f);
                p = psnSpherePoint(sp, ray, ts);
                rayEps = 5e-4f * ts;
                ng = psnFaceforward(nn, mul3(ray.d, -1.f));
            }
        } else {                                       // outside the scene: the bounding sphere catches the ray (:190-194)
            float ts; f3 nn;
            if (!sphereIntersectDg(bound, ray, &ts, &nn)) break;
            nn = mul3(nn, -1.f);                       // ReverseOrientation = true
            p = psnSpherePoint(bound, ray, ts);
            rayEps = 5e-4f * ts;
            ng = psnFaceforward(nn, mul3(ray.d, -1.f));
        }
        f3 dir = psnSampleSphere(words, w);
        dir = psnFaceforward(dir, ng);
        ray.o = p; ray.d = dir; ray.mint = rayEps; ray.maxt = INFINITY;
    }
    candMask[path] = mask; used[path] = (uint32_t)(w - w0); raysTraced[path] = nRays;
}

// FindPoissonPointDistribution for the uploaded scene; the accepted points become the resident surface-point set (as after
// skc_tessellate_resident).  origin3 = the camera position; maxFails 2000 (200 under --quick).  stats4 (optional) = {points,
// paths traced, rays traced, rounds}.
extern "C" int skc_poisson_resident(skc_ctx *ctx, const float *origin3, float minSampleDist, uint32_t materialId, int maxFails, uint64_t *countOut, uint64_t *stats4) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!ctx->bvhNodes || !ctx->meshP) return skc_fail(ctx, "skc_poisson_resident: no scene uploaded (skc_upload_bvh)");
    if (!origin3 || !(minSampleDist > 0.f) || maxFails < 1) return skc_fail(ctx, "skc_poisson_resident: bad arguments");
    // the scene's bounding sphere (surfacepoints.cpp:124-133): BBox::BoundingSphere of the BVH's root bounds
    float4 rootNode[2];
    SKC_CUDA(cudaMemcpyAsync(rootNode, ctx->bvhNodes, sizeof rootNode, cudaMemcpyDeviceToHost, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    const float mn[3] = {rootNode[0].x, rootNode[0].y, rootNode[0].z}, mx[3] = {rootNode[0].w, rootNode[1].x, rootNode[1].y};
    float c[3];
    for (int k = 0; k < 3; ++k) c[k] = .5f * mn[k] + .5f * mx[k];
    const bool inside = c[0] >= mn[0] && c[0] <= mx[0] && c[1] >= mn[1] && c[1] <= mx[1] && c[2] >= mn[2] && c[2] <= mx[2];
    const float dx = c[0] - mx[0], dy = c[1] - mx[1], dz = c[2] - mx[2];
    const float radius = inside ? sqrtf(dx * dx + dy * dy + dz * dz) : 0.f;
    skc_light bound; memset(&bound, 0, sizeof bound);
    for (int i = 0; i < 16; ++i) bound.o2w[i] = bound.w2o[i] = (i % 5 == 0) ? 1.f : 0.f;
    // Translate(sceneCenter - Point(0,0,0)) and its inverse (core/transform.cpp:125-136)
    bound.o2w[3] = c[0]; bound.o2w[7] = c[1]; bound.o2w[11] = c[2];
    bound.w2o[3] = -c[0]; bound.w2o[7] = -c[1]; bound.w2o[11] = -c[2];
    bound.radius = radius; bound.zmin = -radius; bound.zmax = radius;       // Sphere(.., true, r, -r, r, 360)
    bound.thetaMin = acosf(-1.f); bound.thetaMax = acosf(1.f); bound.phiMax = 360.f * (SKC_PI_F / 180.f);
    bound.reverseOrientation = 1.f; bound.nSamples = 1.f;

    const uint32_t nPaths = PSN_PATHS;
    const uint32_t wordsPerPath = 2 + 2 * PSN_DEPTH;
    const size_t candBytes = (size_t)nPaths * PSN_SLOTS * sizeof(skc_surface_point);
    char *buf = (char *)skc_scratch(ctx, 0, candBytes + (size_t)nPaths * (wordsPerPath * 4 + 8 + 12) + 1024);
    if (!buf) return 1;
    skc_surface_point *dCand = (skc_surface_point *)buf;
    uint32_t *dWords = (uint32_t *)(buf + candBytes);
    uint64_t *dOffset = (uint64_t *)(dWords + (size_t)nPaths * wordsPerPath);
    uint32_t *dMask = (uint32_t *)(dOffset + nPaths), *dUsed = dMask + nPaths, *dRays = dUsed + nPaths;

    HostMT rng(0u);                                                          // RNG rng(37 * taskNum), task 0
    std::vector<uint32_t> words;                                             // the stream from the start of the current round on
    std::vector<uint64_t> offset(nPaths);
    std::vector<uint32_t> mask(nPaths), used(nPaths), rays(nPaths);
    std::vector<skc_surface_point> cand((size_t)nPaths * PSN_SLOTS), points;
    std::unordered_map<uint64_t, std::vector<uint32_t>> grid;                // cell -> indices into points
    const float d2max = minSampleDist * minSampleDist;
    const float area = (SKC_PI_F * (minSampleDist / 2.f)) * (minSampleDist / 2.f);      // M_PI * (d / 2) * (d / 2), :203
    auto cellOf = [&](const float *p, int k) { return (int64_t)floorf(p[k] / minSampleDist); };
    auto key = [](int64_t x, int64_t y, int64_t z) { return ((uint64_t)(x & 0x1fffff) << 42) | ((uint64_t)(y & 0x1fffff) << 21) | (uint64_t)(z & 0x1fffff); };
    int repeatedFails = 0;
    uint64_t totalPaths = 0, totalRays = 0, rounds = 0;
    bool done = false;
    const f3 origin = mk3(origin3[0], origin3[1], origin3[2]);
    while (!done) {
        ++rounds;
        while (words.size() < (size_t)nPaths * wordsPerPath) words.push_back(rng.next());
        for (uint32_t i = 0; i < nPaths; ++i) offset[i] = (uint64_t)i * wordsPerPath;
        uint64_t consumed = 0;
        for (int attempt = 0; ; ++attempt) {
            if (attempt > (int)nPaths) return skc_fail(ctx, "skc_poisson_resident: the walk offsets do not settle");
            SKC_CUDA(cudaMemcpyAsync(dWords, words.data(), (size_t)nPaths * wordsPerPath * 4, cudaMemcpyHostToDevice, ctx->stream));
            SKC_CUDA(cudaMemcpyAsync(dOffset, offset.data(), (size_t)nPaths * 8, cudaMemcpyHostToDevice, ctx->stream));
            k_poisson_walk<<<(nPaths + 127) / 128, 128, 0, ctx->stream>>>(skc_bvh_view(ctx), skc_mesh_view(ctx), bound, ctx->bumpTex, origin, dWords, dOffset, nPaths,
                                                                         area, materialId, dCand, dMask, dUsed, dRays);
            SKC_LAUNCHED("k_poisson_walk");
            SKC_CUDA(cudaMemcpyAsync(mask.data(), dMask, (size_t)nPaths * 4, cudaMemcpyDeviceToHost, ctx->stream));
            SKC_CUDA(cudaMemcpyAsync(used.data(), dUsed, (size_t)nPaths * 4, cudaMemcpyDeviceToHost, ctx->stream));
            SKC_CUDA(cudaMemcpyAsync(rays.data(), dRays, (size_t)nPaths * 4, cudaMemcpyDeviceToHost, ctx->stream));
            SKC_CUDA(cudaStreamSynchronize(ctx->stream));
            bool settled = true;
            uint64_t at = 0;
            for (uint32_t i = 0; i < nPaths; ++i) { if (offset[i] != at) { settled = false; offset[i] = at; } at += used[i]; }
            consumed = at;
            if (settled) break;
        }
        SKC_CUDA(cudaMemcpyAsync(cand.data(), dCand, candBytes, cudaMemcpyDeviceToHost, ctx->stream));
        SKC_CUDA(cudaStreamSynchronize(ctx->stream));
        words.erase(words.begin(), words.begin() + consumed);              // an early end leaves words for the next round
        totalPaths += nPaths;
        for (uint32_t i = 0; i < nPaths; ++i) totalRays += rays[i];
        // :228-276 for one task: candidates in order; a rejection counts towards maxFails, an acceptance resets the count
        for (uint32_t i = 0; i < nPaths && !done; ++i)
            for (int s = 0; s < PSN_SLOTS && !done; ++s) {
                if (!((mask[i] >> s) & 1u)) continue;
                const skc_surface_point &sp = cand[(size_t)i * PSN_SLOTS + s];
                const float *p = reinterpret_cast<const float *>(&sp);
                const int64_t cx = cellOf(p, 0), cy = cellOf(p, 1), cz = cellOf(p, 2);
                bool failed = false;
                for (int64_t z = cz - 1; z <= cz + 1 && !failed; ++z) for (int64_t y = cy - 1; y <= cy + 1 && !failed; ++y) for (int64_t x = cx - 1; x <= cx + 1 && !failed; ++x) {
                    auto it = grid.find(key(x, y, z));
                    if (it == grid.end()) continue;
                    for (uint32_t q : it->second) {
                        const float *pq = reinterpret_cast<const float *>(&points[q]);
                        const float ex = pq[0] - p[0], ey = pq[1] - p[1], ez = pq[2] - p[2];
                        if (ex * ex + ey * ey + ez * ez < d2max) { failed = true; break; }        // PoissonCheck, :86-98
                    }
                }
                if (failed) { if (++repeatedFails >= maxFails) done = true; }
                else { repeatedFails = 0; grid[key(cx, cy, cz)].push_back((uint32_t)points.size()); points.push_back(sp); }
            }
        if (!done && totalPaths > 50000 && points.empty()) return skc_fail(ctx, "skc_poisson_resident: there don't seem to be any objects with BSSRDFs in this scene");
    }
    if (skc_set_surface_points(ctx, points.data(), points.size())) return 1;
    if (countOut) *countOut = points.size();
    if (stats4) { stats4[0] = points.size(); stats4[1] = totalPaths; stats4[2] = totalRays; stats4[3] = rounds; }
    return 0;
}

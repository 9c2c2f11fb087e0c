// skc_frame.cu -- the rest of the frame around the subsurface term, on the device.
//
// Reference: SamplerRenderer::Render / SamplerRendererTask::Run (renderers/samplerrenderer.cpp:59-167,191-225),
// LDSampler (samplers/lowdiscrepancy.cpp:39-79) + LDPixelSample (core/montecarlo.cpp:200-252) + the scrambled
// (0,2)-sequence and Shuffle (core/montecarlo.h:183-189,278-331), Sampler::ComputeSubWindow (core/sampler.cpp:56-75),
// PerspectiveCamera::GenerateRay (cameras/perspective.cpp:79-113), MultipoleSubsurfaceIntegrator::Li
// (integrators/multipolesubsurface.cpp:253-304), UniformSampleAllLights / EstimateDirect (core/integrator.cpp:47-155),
// BSDF::f / Pdf / Sample_f (core/reflection.cpp:675-780), ImageFilm::AddSample (film/image.cpp:77-137),
// SampledSpectrum::ToXYZ / y (core/spectrum.h:374-393).
//
// Kernels:
//   k_mt_words      one CTA per SamplerRendererTask: the task's MT19937 stream RNG(taskNum), tempered, 624 words at a
//                   time.  LDPixelSample draws a fixed number K of words per pixel (scrambles + Fisher-Yates shuffles),
//                   and the task's RNG is ALSO drawn from by Li: SpecularReflect and SpecularTransmit each build a
//                   BSDFSample(rng) (3 floats, core/integrator.cpp:184,223) for every camera ray that hits something,
//                   before finding that the skin BSDF has no specular lobe.  So pixel k of a task starts at word
//                   k*K + 6*(camera rays of pixels < k that hit)
//   k_ld_chain      one warp per task walks its pixels in order: image samples of the pixel (10 words at 4 spp) ->
//                   camera rays -> any-hit -> start of the next pixel.  This is the only sequential part of the frame
//   k_ld_pixel      one thread per pixel: the pixel's image/lens/time samples and every 1-D / 2-D integrator array,
//                   generated, shuffled within and across the pixel's samples exactly as LDShuffleScrambled1D/2D do
//   k_camera_rays   one thread per camera sample
//   k_direct_terms  one thread per (camera sample, light, light sample): EstimateDirect's two MIS terms -- light
//                   sampling with a shadow ray (any-hit) and BSDF sampling with a closest-hit ray that must reach the
//                   sampled emitter -- reduced to the scalars the spectral sum needs (the BSDF is grey)
//   k_frame_sum     one warp per camera sample, lane = wavelength: L = Le + subsurface term + sum over lights of
//                   (sum over samples)/nSamples in the reference's order, radiance sanity checks, ToXYZ and
//                   ImageFilm::AddSample (atomic adds: the filter support may cross pixels)
#include <algorithm>
#include <vector>
#include "skc_light.cuh"
#include "skc_texture.cuh"
#include "skc_bsdf.cuh"
#include "skc_internal.h"

#define FR_MAX_ARRAYS 36          // 2 per light (16 lights) + extras
#define FR_MAX_VALUES 64          // nSamples * spp of one array
#define FR_MAX_LIGHTS 16

struct FrameLayout {
    uint32_t spp, n1D, n2D;
    uint32_t size1D[FR_MAX_ARRAYS], size2D[FR_MAX_ARRAYS];
    uint32_t draws;               // MT words per pixel
    uint32_t floats;              // floats per camera sample (all 1-D arrays, then all 2-D arrays)
    uint32_t nLights, slots;      // slots = sum of nSamples over lights
    uint32_t nS[FR_MAX_LIGHTS];
    uint32_t off1L[FR_MAX_LIGHTS], off2L[FR_MAX_LIGHTS], off1B[FR_MAX_LIGHTS], off2B[FR_MAX_LIGHTS];  // float offsets inside a sample's block
};

struct FrameConsts {              // device copy of the per-frame tables
    float c2w[16], r2c[16];
    float filterTable[256];
    float cie[3][32];
    float Kr[32], Kt[32];
    float xyzScale, filterXWidth, filterYWidth, invXWidth, invYWidth;
    int xres, yres;
};

struct skc_frame_state {
    skc_frame_desc desc;
    FrameLayout lay;
    std::vector<int4> win;        // emitted tasks, in emission order: {x0, x1, y0, y1}
    std::vector<uint32_t> seed, pixBase;
    uint64_t nPixels = 0, nSamples = 0;
    int4 *dWin = nullptr; uint32_t *dSeed = nullptr, *dPixBase = nullptr;
    float2 *dXY = nullptr; float *dArrays = nullptr; float4 *dFilm = nullptr;
    uint32_t *dOrder = nullptr;   // render order: pixel slot -> pixel of the sampler stream (Z-order inside each task window)
    FrameConsts *dConsts = nullptr;
    bool samplesReady = false;
    // skc_frame_prefetch: the part of the frame that does not depend on Preprocess, enqueued on prefetchStream
    bool prefetched = false;
    bool lazyArrays = false;            // skc_frame_prefetch: no integrator arrays for pixels whose camera rays all miss (k_ld_pixel)
    cudaEvent_t evP0 = nullptr, evP1 = nullptr, evDone = nullptr, evHits = nullptr;   // evHits: closest hits done (what Mo needs); evP1: direct terms done (what the spectral sum needs)
};

void skc_frame_release(skc_ctx *ctx) {
    skc_frame_state *f = ctx->frame;
    if (!f) return;
    void *ptrs[] = {f->dWin, f->dSeed, f->dPixBase, f->dXY, f->dArrays, f->dFilm, f->dConsts, f->dOrder};
    for (void *p : ptrs) if (p) cudaFree(p);
    for (cudaEvent_t e : {f->evP0, f->evP1, f->evDone, f->evHits}) if (e) cudaEventDestroy(e);
    delete f;
    ctx->frame = nullptr;
}

// ---------------------------------------------------------------------------------------------------------------
// sampler stream
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_mt_words(const uint32_t *__restrict__ seed, const uint32_t *__restrict__ pixBase,
                                                 uint32_t draws, uint32_t *__restrict__ out) {
    __shared__ uint32_t mt[MT_N];
    uint32_t t = blockIdx.x;
    uint64_t base0 = (uint64_t)pixBase[t] * draws, nWords = (uint64_t)(pixBase[t + 1] - pixBase[t]) * draws;
    int tid = threadIdx.x;
    if (tid == 0) {   // RNG::Seed, core/rng.cpp:43-55
        uint32_t s = seed[t];
        mt[0] = s;
        for (int i = 1; i < MT_N; ++i) { s = 1812433253u * (s ^ (s >> 30)) + (uint32_t)i; mt[i] = s; }
    }
    __syncthreads();
    for (uint64_t base = 0; base < nWords; base += MT_N) {
        uint32_t v = 0;     // core/rng.cpp:84-103 in three dependency-free phases
        if (tid < 227) v = mtTwist(mt[tid], mt[tid + 1], mt[tid + MT_M]);
        __syncthreads();
        if (tid < 227) mt[tid] = v;
        __syncthreads();
        if (tid < 227) v = mtTwist(mt[tid + 227], mt[tid + 228], mt[tid]);
        __syncthreads();
        if (tid < 227) mt[tid + 227] = v;
        __syncthreads();
        if (tid < 169) v = mtTwist(mt[tid + 454], mt[tid + 455], mt[tid + 227]);
        __syncthreads();
        if (tid < 169) mt[tid + 454] = v;
        __syncthreads();
        if (tid == 0) mt[623] = mtTwist(mt[623], mt[0], mt[396]);
        __syncthreads();
        for (int k = tid; k < MT_N; k += 256) {
            uint64_t w = base + k;
            if (w < nWords) {
                uint32_t y = mt[k];       // tempering, core/rng.cpp:105-110
                y ^= (y >> 11);
                y ^= (y << 7) & 0x9d2c5680u;
                y ^= (y << 15) & 0xefc60000u;
                y ^= (y >> 18);
                out[base0 + w] = y;
            }
        }
        __syncthreads();
    }
}

// Shuffle (core/montecarlo.h:183-189)
__device__ __forceinline__ void ldShuffle(float *samp, uint32_t count, uint32_t dims, const uint32_t *__restrict__ r, uint32_t &ri) {
    for (uint32_t i = 0; i < count; ++i) {
        uint32_t other = i + (r[ri++] % (count - i));
        for (uint32_t j = 0; j < dims; ++j) { float t = samp[dims * i + j]; samp[dims * i + j] = samp[dims * other + j]; samp[dims * other + j] = t; }
    }
}
// LDShuffleScrambled1D / 2D (core/montecarlo.h:312-331)
__device__ inline void ldScrambled1D(uint32_t nSamples, uint32_t nPixel, float *samples, const uint32_t *__restrict__ r, uint32_t &ri) {
    uint32_t scramble = r[ri++];
    for (uint32_t i = 0; i < nSamples * nPixel; ++i) samples[i] = vanDerCorput(i, scramble);
    for (uint32_t i = 0; i < nPixel; ++i) ldShuffle(samples + i * nSamples, nSamples, 1, r, ri);
    ldShuffle(samples, nPixel, nSamples, r, ri);
}
__device__ inline void ldScrambled2D(uint32_t nSamples, uint32_t nPixel, float *samples, const uint32_t *__restrict__ r, uint32_t &ri) {
    uint32_t s0 = r[ri++], s1 = r[ri++];
    for (uint32_t i = 0; i < nSamples * nPixel; ++i) { samples[2 * i] = vanDerCorput(i, s0); samples[2 * i + 1] = sobol2(i, s1); }
    for (uint32_t i = 0; i < nPixel; ++i) ldShuffle(samples + 2 * i * nSamples, nSamples, 2, r, ri);
    ldShuffle(samples, nPixel, 2 * nSamples, r, ri);
}

// LDPixelSample (core/montecarlo.cpp:200-252) for one pixel per thread
__global__ void __launch_bounds__(128) k_ld_pixel(FrameLayout lay, const int4 *__restrict__ win, const uint32_t *__restrict__ pixBase,
                                                 uint32_t nTasks, const uint32_t *__restrict__ words, const uint64_t *__restrict__ pixOff,
                                                 float2 *__restrict__ xy, float *__restrict__ arrays, int skipUnhit) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= pixBase[nTasks]) return;
    uint32_t lo = 0, hi = nTasks - 1;
    while (lo < hi) { uint32_t mid = (lo + hi + 1) >> 1; if (pixBase[mid] <= g) lo = mid; else hi = mid - 1; }
    int4 w = win[lo];
    uint32_t k = g - pixBase[lo], wx = (uint32_t)(w.y - w.x);
    int xPos = w.x + (int)(k % wx), yPos = w.z + (int)(k / wx);     // LDSampler::GetMoreSamples: x fastest (lowdiscrepancy.cpp:66-79)
    const uint32_t *r = words + pixOff[g];
    uint32_t ri = 0;
    const uint32_t spp = lay.spp;
    float buf[2 * FR_MAX_VALUES];
    uint64_t s0 = (uint64_t)g * spp;
    ldScrambled2D(1, spp, buf, r, ri);                               // image samples
    for (uint32_t i = 0; i < spp; ++i) xy[s0 + i] = make_float2((float)xPos + buf[2 * i], (float)yPos + buf[2 * i + 1]);
    // The integrator's arrays of a pixel are read only for camera rays that hit something (EstimateDirect).  k_ld_chain already
    // knows which pixels those are -- every hit moves the next pixel's start by 6 words -- so on the prefetch path the arrays of
    // pixels without a hit are not generated (their words are still accounted for: the offsets come from the chain).
    if (skipUnhit && g + 1 < pixBase[lo + 1] && pixOff[g + 1] - pixOff[g] == (uint64_t)lay.draws) return;
    ldScrambled2D(1, spp, buf, r, ri);                               // lens samples (pinhole: unused)
    ldScrambled1D(1, spp, buf, r, ri);                               // time samples (shutter 0..0: unused)
    float *out = arrays + s0 * lay.floats;
    uint32_t off = 0;
    for (uint32_t j = 0; j < lay.n1D; ++j) {
        uint32_t n = lay.size1D[j];
        ldScrambled1D(n, spp, buf, r, ri);
        for (uint32_t i = 0; i < spp; ++i) for (uint32_t q = 0; q < n; ++q) out[(uint64_t)i * lay.floats + off + q] = buf[n * i + q];
        off += n;
    }
    for (uint32_t j = 0; j < lay.n2D; ++j) {
        uint32_t n = lay.size2D[j];
        ldScrambled2D(n, spp, buf, r, ri);
        for (uint32_t i = 0; i < spp; ++i) for (uint32_t q = 0; q < 2 * n; ++q) out[(uint64_t)i * lay.floats + off + q] = buf[2 * n * i + q];
        off += 2 * n;
    }
}

// PerspectiveCamera::GenerateRay (cameras/perspective.cpp:79-113) without depth of field
__device__ __forceinline__ DRay cameraRay(const FrameConsts *__restrict__ C, float2 s) {
    f3 Pcamera = xformPoint(C->r2c, mk3(s.x, s.y, 0.f));
    f3 dir = normalize3(Pcamera);
    DRay r;
    r.o = xformPoint(C->c2w, mk3(0.f, 0.f, 0.f));
    r.d = xformVector(C->c2w, dir);
    r.mint = 0.f; r.maxt = INFINITY;
    return r;
}
// Render order.  The kernels after the sampler work on camera samples in an order of their own choosing: within each
// task window the pixels are visited along a Z-curve, so that the 32 consecutive samples of one Mo packet cover a
// compact 4x2-pixel block (at 4 spp) instead of a strip -- neighbouring hit points share their octree walk.  Sample j
// of the render order is sample renderToStream(j) of the reference's stream; the film does not depend on the order.
__device__ __forceinline__ uint64_t renderToStream(const uint32_t *__restrict__ order, uint32_t spp, uint64_t j) {
    return order ? (uint64_t)__ldg(&order[j / spp]) * spp + (j % spp) : j;
}
__global__ void __launch_bounds__(256) k_camera_rays(const FrameConsts *__restrict__ C, const float2 *__restrict__ xy, uint64_t n,
                                                    const uint32_t *__restrict__ order, uint32_t spp, skc_ray *__restrict__ rays) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    DRay r = cameraRay(C, xy[renderToStream(order, spp, i)]);
    float4 *dst = reinterpret_cast<float4 *>(rays + i);
    dst[0] = make_float4(r.o.x, r.o.y, r.o.z, r.d.x);
    dst[1] = make_float4(r.d.y, r.d.z, 0.f, INFINITY);
}

// "texture albedo" at the camera hits (materials/layeredskin.cpp:179-185 through Intersection::GetMultipoleBSSRDF): one thread per camera
// sample that hit the skin.  The hit's DifferentialGeometry as Triangle::Intersect builds it (shapes/trianglemesh.inl:80-135: dpdu, dpdv
// from the uv parameterisation, (u, v), the geometric normal), the camera ray's differentials (cameras/perspective.cpp:106-112, scaled by
// 1 / sqrt(spp): samplerrenderer.cpp:97), DifferentialGeometry::ComputeDifferentials (core/diffgeom.cpp:58-113), UVMapping2D and the EWA
// lookup.  Output: the albedo's RGB per sample (the Li epilogue turns it into Pow(Spectrum::FromRGB(rgb), 1 - mix)).
__device__ __forceinline__ bool solve2x2(const float A[2][2], const float B[2], float *x0, float *x1) {      // core/transform.cpp:39-49
    const float det = A[0][0] * A[1][1] - A[0][1] * A[1][0];
    if (fabsf(det) < 1e-10f) return false;
    *x0 = (A[1][1] * B[0] - A[0][1] * B[1]) / det;
    *x1 = (A[0][0] * B[1] - A[1][0] * B[0]) / det;
    return !(isnan(*x0) || isnan(*x1));
}
__device__ __forceinline__ float comp3(f3 v, int a) { return a == 0 ? v.x : a == 1 ? v.y : v.z; }
__global__ void __launch_bounds__(128) k_hit_textures(BvhView bv, MeshView mv, const __grid_constant__ AlbedoTexDev tex, const __grid_constant__ AlbedoTexDev bump,
                                                      const FrameConsts *__restrict__ C, const float2 *__restrict__ xy,
                                                      const uint32_t *__restrict__ order, uint32_t spp, uint64_t first, float diffScale,
                                                      const skc_ray *__restrict__ rays, skc_hit *__restrict__ hits, uint64_t n,
                                                      float4 *__restrict__ albedoRgb, float4 *__restrict__ frame) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *hp = reinterpret_cast<const float *>(hits + i);
    float4 out = make_float4(1.f, 1.f, 1.f, 0.f);
    if (__float_as_int(hp[0]) && __float_as_int(hp[13])) {
        const int prim = __float_as_int(hp[1]);
        const float4 *rp = reinterpret_cast<const float4 *>(rays + i);
        const float4 r0 = __ldg(rp), r1 = __ldg(rp + 1);
        DRay ray; ray.o = mk3(r0.x, r0.y, r0.z); ray.d = mk3(r0.w, r1.x, r1.y); ray.mint = r1.z; ray.maxt = r1.w;
        // the barycentrics of the hit: the same test on the same operands as the traversal's
        const float4 *pp = &bv.prims[3 * prim];
        const float4 a = __ldg(pp), b = __ldg(pp + 1), c = __ldg(pp + 2);
        float t, b1 = 0.f, b2 = 0.f;
        triIntersect(mk3(a.x, a.y, a.z), mk3(a.w, b.x, b.y), mk3(b.z, b.w, c.x), ray, &t, &b1, &b2);
        const int4 vi = __ldg(&mv.primInfo[prim]);
        f3 dpdu, dpdv, nn; float tu, tv;
        triDiffGeom(mv, vi, 1 - b1 - b2, b1, b2, &dpdu, &dpdv, &tu, &tv, &nn);
        const f3 p = mk3(hp[4], hp[5], hp[6]);
        // the offset rays
        const float2 s = xy[renderToStream(order, spp, first + i)];
        const f3 Pcamera = xformPoint(C->r2c, mk3(s.x, s.y, 0.f));
        const f3 P000 = xformPoint(C->r2c, mk3(0.f, 0.f, 0.f));
        const f3 dxCamera = sub3(xformPoint(C->r2c, mk3(1.f, 0.f, 0.f)), P000), dyCamera = sub3(xformPoint(C->r2c, mk3(0.f, 1.f, 0.f)), P000);
        f3 rxD = xformVector(C->c2w, normalize3(add3(Pcamera, dxCamera))), ryD = xformVector(C->c2w, normalize3(add3(Pcamera, dyCamera)));
        rxD = add3(ray.d, mul3(sub3(rxD, ray.d), diffScale));                 // ScaleDifferentials (core/geometry.h:412-417); the origins stay ray.o
        ryD = add3(ray.d, mul3(sub3(ryD, ray.d), diffScale));
        const f3 rxO = add3(ray.o, mul3(sub3(ray.o, ray.o), diffScale)), ryO = rxO;
        // ComputeDifferentials
        float dudx = 0.f, dvdx = 0.f, dudy = 0.f, dvdy = 0.f;
        const float d = -dot3(nn, p);
        const float tx = -(dot3(nn, rxO) + d) / dot3(nn, rxD);
        const float ty = -(dot3(nn, ryO) + d) / dot3(nn, ryD);
        if (!isnan(tx) && !isnan(ty)) {
            const f3 px = add3(rxO, mul3(rxD, tx)), py = add3(ryO, mul3(ryD, ty));
            int ax0, ax1;
            if (fabsf(nn.x) > fabsf(nn.y) && fabsf(nn.x) > fabsf(nn.z)) { ax0 = 1; ax1 = 2; }
            else if (fabsf(nn.y) > fabsf(nn.z)) { ax0 = 0; ax1 = 2; }
            else { ax0 = 0; ax1 = 1; }
            float A[2][2], Bx[2], By[2];
            A[0][0] = comp3(dpdu, ax0); A[0][1] = comp3(dpdv, ax0); A[1][0] = comp3(dpdu, ax1); A[1][1] = comp3(dpdv, ax1);
            Bx[0] = comp3(px, ax0) - comp3(p, ax0); Bx[1] = comp3(px, ax1) - comp3(p, ax1);
            By[0] = comp3(py, ax0) - comp3(p, ax0); By[1] = comp3(py, ax1) - comp3(p, ax1);
            if (!solve2x2(A, Bx, &dudx, &dvdx)) { dudx = 0.f; dvdx = 0.f; }
            if (!solve2x2(A, By, &dudy, &dvdy)) { dudy = 0.f; dvdy = 0.f; }
        }
        // UVMapping2D::Map (core/texture.cpp) and MIPMap::Lookup
        if (tex.rgb) {
            const f3 rgb = texLookup(tex, tex.su * tu + tex.du, tex.sv * tv + tex.dv, tex.su * dudx, tex.sv * dvdx, tex.su * dudy, tex.sv * dvdy);
            out = make_float4(rgb.x, rgb.y, rgb.z, 1.f);
        }
        if (bump.rgb) {     // LayeredSkin::GetBSDF -> Material::Bump: the BSDF's shading normal and dpdu (the geometric normal stays)
            f3 nnB, dpduB;
            triBump(mv, vi, 1 - b1 - b2, b1, b2, bump, dudx, dvdx, dudy, dvdy, &nnB, &dpduB);
            float *hw = reinterpret_cast<float *>(hits + i);
            hw[7] = nnB.x; hw[8] = nnB.y; hw[9] = nnB.z;
            frame[2 * i].w = dpduB.x; frame[2 * i + 1].x = dpduB.y; frame[2 * i + 1].y = dpduB.z;
        }
    }
    if (albedoRgb) albedoRgb[i] = out;
}

// Start of every pixel's words in its task's stream: off[k] = k*K + 6 * (camera rays of pixels < k that hit).
// Pixel k's own hit count depends on off[k] (through its image samples), so this is a chain -- but an almost
// data-independent one: whether a pixel's rays hit is decided by where the pixel is, and only at silhouettes by where
// exactly its samples fall.  One CTA per task iterates to the fixed point instead of walking the pixels one by one:
// every round computes all pixels' samples from the current offsets, traces all camera rays in parallel, and rescans
// the hit counts; pixel k is final after at most k+1 rounds (induction over k), in practice after 2-4 rounds.
#define LC_THREADS 256
__global__ void __launch_bounds__(LC_THREADS) k_ld_chain(BvhView bv, const FrameConsts *__restrict__ C, FrameLayout lay,
                                                        const int4 *__restrict__ win, const uint32_t *__restrict__ pixBase,
                                                        uint32_t cap, const uint32_t *__restrict__ words,
                                                        uint64_t *__restrict__ pixOff, float2 *__restrict__ xy) {
    extern __shared__ uint32_t sm[];                 // [nPix] hit counts, then [nPix] relative offsets
    __shared__ int changed;
    uint32_t t = blockIdx.x;
    int4 w = win[t];
    uint32_t p0 = pixBase[t], nPix = pixBase[t + 1] - p0, wx = (uint32_t)(w.y - w.x);
    uint32_t *hits = sm, *rel = sm + nPix;
    const uint32_t *tw = words + (uint64_t)p0 * cap; // the task's words start here (k_mt_words)
    const uint32_t spp = lay.spp;
    int tid = threadIdx.x;
    for (uint32_t k = tid; k < nPix; k += LC_THREADS) rel[k] = k * lay.draws;
    __syncthreads();
    for (uint32_t round = 0; round <= nPix; ++round) {
        // image samples of every pixel at its current offset (LDPixelSample's first array)
        for (uint32_t k = tid; k < nPix; k += LC_THREADS) {
            float buf[2 * FR_MAX_VALUES];
            uint32_t ri = 0;
            ldScrambled2D(1, spp, buf, tw + rel[k], ri);
            int xPos = w.x + (int)(k % wx), yPos = w.z + (int)(k / wx);
            for (uint32_t i = 0; i < spp; ++i) xy[(uint64_t)(p0 + k) * spp + i] = make_float2((float)xPos + buf[2 * i], (float)yPos + buf[2 * i + 1]);
            hits[k] = 0;
        }
        if (tid == 0) changed = 0;
        __syncthreads();
        // Scene::Intersect is true iff anything lies on the camera ray
        for (uint32_t j = tid; j < nPix * spp; j += LC_THREADS) {
            DRay r = cameraRay(C, xy[(uint64_t)p0 * spp + j]);
            if (bvhOccludedFast(bv, r)) atomicAdd(&hits[j / spp], 1u);
        }
        __syncthreads();
        // exclusive scan of the hit counts (warp 0, 32 pixels at a time)
        if (tid < 32) {
            uint32_t running = 0;
            bool ch = false;
            for (uint32_t b = 0; b < nPix; b += 32) {
                uint32_t k = b + tid;
                uint32_t h = k < nPix ? hits[k] : 0, incl = h;
                for (int o = 1; o < 32; o <<= 1) { uint32_t v = __shfl_up_sync(0xffffffffu, incl, o); if (tid >= o) incl += v; }
                if (k < nPix) {
                    uint32_t nr = k * lay.draws + 6 * (running + incl - h);
                    ch |= nr != rel[k];
                    rel[k] = nr;
                }
                running += __shfl_sync(0xffffffffu, incl, 31);
            }
            if (__any_sync(0xffffffffu, ch) && tid == 0) changed = 1;
        }
        __syncthreads();
        if (!changed) break;
        __syncthreads();
    }
    for (uint32_t k = tid; k < nPix; k += LC_THREADS) pixOff[p0 + k] = (uint64_t)p0 * cap + rel[k];
}

// ---------------------------------------------------------------------------------------------------------------
// direct lighting
// ---------------------------------------------------------------------------------------------------------------
#define FR_INV_PI 0.31830988618379067154f

struct FrameBsdf {      // BSDF::BSDF (core/reflection.cpp:755-762) + the lobes LayeredSkin / MatteMaterial add
    f3 nn, ng, sn, tn;
    bool matte, hasR, hasT;             // skin: Microfacet iff Kr is not black, MicrofacetTransmission iff Kt is not black
    RhoBxdf mf; float ior, lambert;     // lambert = Kd * INV_PI
};
__device__ __forceinline__ f3 bsdfToLocal(const FrameBsdf &B, f3 v) { return mk3(dot3(v, B.sn), dot3(v, B.tn), dot3(v, B.nn)); }
__device__ __forceinline__ f3 bsdfToWorld(const FrameBsdf &B, f3 v) {
    return mk3(B.sn.x * v.x + B.tn.x * v.y + B.nn.x * v.z, B.sn.y * v.x + B.tn.y * v.y + B.nn.y * v.z, B.sn.z * v.x + B.tn.z * v.y + B.nn.z * v.z);
}
// The part of BSDF::f / Sample_f that sums the lobes on the side of the geometric normal wi lies on
// (reflection.cpp:719-728,765-778).  Returns the value for unit Kr / Kt (zero <=> black) and the factors the spectral
// sum needs in comp: {D, G, F, den} reflection, {a, 1-F, -1, 1} transmission, {Kd/pi, 0, -2, 1} matte.
__device__ inline float bsdfEvalLocal(const FrameBsdf &B, f3 woW, f3 wiW, f3 wo, f3 wi, float4 *comp) {
    bool reflect = dot3(wiW, B.ng) * dot3(woW, B.ng) > 0;
    if (B.matte) { *comp = make_float4(B.lambert, 0.f, -2.f, 1.f); return reflect ? B.lambert : 0.f; }
    if (reflect) return B.hasR ? microfacetF(B.mf, wo, wi, comp) : 0.f;
    return B.hasT ? transmissionF(B.mf, B.ior, wo, wi, comp) : 0.f;
}
__device__ inline float bsdfF(const FrameBsdf &B, f3 woW, f3 wiW, float4 *comp) {
    f3 wi = bsdfToLocal(B, wiW), wo = bsdfToLocal(B, woW);
    return bsdfEvalLocal(B, woW, wiW, wo, wi, comp);
}
// BSDF::Pdf (reflection.cpp:736-752) for flags = BSDF_ALL & ~BSDF_SPECULAR
__device__ inline float bsdfPdf(const FrameBsdf &B, f3 woW, f3 wiW) {
    f3 wo = bsdfToLocal(B, woW), wi = bsdfToLocal(B, wiW);
    if (B.matte) return (0.f + (sameHemisphere(wo, wi) ? fabsf(wi.z) * FR_INV_PI : 0.f)) / 1;
    float pdf = 0.f;
    int matchingComps = 0;
    if (B.hasR) { ++matchingComps; pdf += microfacetPdf(B.mf, wo, wi); }
    if (B.hasT) { ++matchingComps; pdf += transmissionPdf(B.mf, B.ior, wo, wi); }
    return matchingComps > 0 ? pdf / matchingComps : 0.f;
}
// BSDF::Sample_f (reflection.cpp:675-733); returns the value as bsdfEvalLocal does (0 = black), *pdf may be 0
__device__ inline float bsdfSample(const FrameBsdf &B, f3 woW, float uComponent, float u1, float u2, f3 *wiW, float *pdf, float4 *comp) {
    f3 wo = bsdfToLocal(B, woW), wi;
    *pdf = 0.f;
    if (B.matte) {
        wi = cosineSampleHemisphere(u1, u2);                   // BxDF::Sample_f (reflection.cpp:370-377)
        if (wo.z < 0.) wi.z *= -1.f;
        *pdf = sameHemisphere(wo, wi) ? fabsf(wi.z) * FR_INV_PI : 0.f;
        if (*pdf == 0.f) return 0.f;
        *wiW = bsdfToWorld(B, wi);
        return bsdfEvalLocal(B, woW, *wiW, wo, wi, comp);
    }
    int matchingComps = (B.hasR ? 1 : 0) + (B.hasT ? 1 : 0);
    if (matchingComps == 0) return 0.f;
    int which = min((int)floorf(uComponent * matchingComps), matchingComps - 1);
    bool sampleR = B.hasR && which == 0;                       // bxdfs[] order: Microfacet, MicrofacetTransmission
    if (sampleR) beckmannSample(B.mf, wo, u1, u2, &wi, pdf);
    else transmissionSample(B.mf, B.ior, wo, u1, u2, &wi, pdf);
    if (*pdf == 0.f) return 0.f;
    *wiW = bsdfToWorld(B, wi);
    if (matchingComps > 1) {
        if (sampleR) *pdf += transmissionPdf(B.mf, B.ior, wo, wi);
        else *pdf += microfacetPdf(B.mf, wo, wi);
        *pdf /= matchingComps;
    }
    return bsdfEvalLocal(B, woW, *wiW, wo, wi, comp);
}
__device__ __forceinline__ float powerHeuristic(float fPdf, float gPdf) {   // core/montecarlo.h:270-273 with nf = ng = 1
    float f = 1 * fPdf, g = 1 * gPdf;
    return (f * f) / (f * f + g * g);
}

// EstimateDirect (core/integrator.cpp:108-155) for one (camera sample, light, light sample).
// terms[3t]   = BSDF factors of the light-sampling term (bsdfEvalLocal's comp; den = 0: term absent)
// terms[3t+1] = {|wi.n| * weight / lightPdf of the light-sampling term;  |wi.n|, weight, bsdfPdf of the BSDF-sampling term}
// terms[3t+2] = BSDF factors of the BSDF-sampling term (den = 0: absent)
template <bool ENV>      // ENV: the scene has an infinite light (the instance without it is the benchmarked frame's: no map parameter, no branch)
__global__ void __launch_bounds__(128) k_direct_terms(BvhView bv, const skc_light *__restrict__ lights, FrameLayout lay,
                                                     RhoBxdf mf, float ior, float kd, int hasR, int hasT,
                                                     const skc_hit *__restrict__ hits, const float4 *__restrict__ frame,
                                                     const float *__restrict__ arrays, const uint32_t *__restrict__ order, uint64_t first,
                                                     uint64_t n, float4 *__restrict__ terms, EnvMapDev env, float4 *__restrict__ termsEnv) {
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n * lay.slots) return;
    uint64_t s = t / lay.slots;
    uint32_t j = (uint32_t)(t % lay.slots), li = 0;
    while (j >= lay.nS[li]) { j -= lay.nS[li]; ++li; }
    float4 T0 = make_float4(0.f, 0.f, 0.f, 0.f), T1 = T0, T2 = T0;
    const float *hp = reinterpret_cast<const float *>(hits + s);
    if (__float_as_int(__ldg(hp))) {
        float eps = __ldg(hp + 3);
        f3 p = mk3(__ldg(hp + 4), __ldg(hp + 5), __ldg(hp + 6));
        f3 wo = mk3(__ldg(hp + 10), __ldg(hp + 11), __ldg(hp + 12));
        float4 a0 = __ldg(&frame[2 * s]), a1 = __ldg(&frame[2 * s + 1]);
        FrameBsdf B;
        B.nn = mk3(__ldg(hp + 7), __ldg(hp + 8), __ldg(hp + 9));
        B.ng = mk3(a0.x, a0.y, a0.z);
        B.sn = normalize3(mk3(a0.w, a1.x, a1.y));
        B.tn = cross3(B.nn, B.sn);
        B.matte = __float_as_int(__ldg(hp + 13)) == 0;
        B.mf = mf; B.ior = ior; B.lambert = kd * FR_INV_PI; B.hasR = hasR != 0; B.hasT = hasT != 0;
        const float *A = arrays + renderToStream(order, lay.spp, first + s) * lay.floats;
        float lu0 = __ldg(A + lay.off2L[li] + 2 * j), lu1 = __ldg(A + lay.off2L[li] + 2 * j + 1);
        float bc = __ldg(A + lay.off1B[li] + j), bu0 = __ldg(A + lay.off2B[li] + 2 * j), bu1 = __ldg(A + lay.off2B[li] + 2 * j + 1);
        const skc_light &L = lights[li];
        const int kind = lightKind(L);
        if (ENV && kind == 4) {
            // infinite light (lights/infinite.cpp): both MIS terms as for an emitter shape, but the radiance is the map's RGB in the
            // term's direction (termsEnv[2t], [2t+1]; k_frame_sum turns it into a spectrum), and the BSDF-sampled ray sees the
            // light only if it leaves the scene (integrator.cpp:144-148: `else Li = light->Le(ray)`)
            f3 wi, rgbA = mk3(0.f, 0.f, 0.f), rgbB = rgbA; float lightPdf; DRay sh;
            if (sampleInfiniteLight(L, env, p, eps, lu0, lu1, &wi, &lightPdf, &rgbA, &sh) && lightPdf > 0.) {
                float4 comp;
                float f = bsdfF(B, wo, wi, &comp);
                if (f != 0.f && !bvhOccludedFast(bv, sh)) {
                    float bp = bsdfPdf(B, wo, wi);
                    float weight = powerHeuristic(lightPdf, bp);
                    T0 = comp; T1.x = fabsf(dot3(wi, B.nn)) * weight / lightPdf;
                }
            }
            float bp;
            float4 comp2;
            float f = bsdfSample(B, wo, bc, bu0, bu1, &wi, &bp, &comp2);
            if (f != 0.f && bp > 0.) {
                float lp = infiniteLightPdf(L, env, wi);
                if (lp != 0.) {
                    float weight = powerHeuristic(bp, lp);
                    DRay ray; ray.o = p; ray.d = wi; ray.mint = eps; ray.maxt = INFINITY;
                    if (!bvhOccludedFast(bv, ray)) {
                        rgbB = infiniteLightLe(L, env, wi);
                        T2 = comp2; T1.y = fabsf(dot3(wi, B.nn)); T1.z = weight; T1.w = bp;
                    }
                }
            }
            terms[3 * t] = T0; terms[3 * t + 1] = T1; terms[3 * t + 2] = T2;
            termsEnv[2 * t] = make_float4(rgbA.x, rgbA.y, rgbA.z, 0.f); termsEnv[2 * t + 1] = make_float4(rgbB.x, rgbB.y, rgbB.z, 0.f);
            return;
        }
        if (kind != 0) {
            // delta light (integrator.cpp:114-126): Ld += f * Li * (AbsDot(wi, n) / lightPdf), lightPdf = 1, and no BSDF-sampling
            // term (:134).  T1.y / T1.z carry DistanceSquared and the spot falloff: Li = (I * falloff) / d2 in k_frame_sum.
            f3 wi; float d2, fall; DRay sh;
            if (sampleDeltaLight(L, kind, p, eps, &wi, &d2, &fall, &sh) && d2 != 0.f) {
                float4 comp;
                float f = bsdfF(B, wo, wi, &comp);
                if (f != 0.f && !bvhOccludedFast(bv, sh)) { T0 = comp; T1.x = fabsf(dot3(wi, B.nn)) / 1.f; T1.y = d2; T1.z = fall; }
            }
            terms[3 * t] = T0; terms[3 * t + 1] = T1; terms[3 * t + 2] = T2;
            return;
        }
        // ---- sample the light (integrator.cpp:114-131)
        f3 wi; float lightPdf; DRay sh;
        bool lit = sampleSphereLightL(L, p, eps, lu0, lu1, &wi, &lightPdf, &sh);
        if (lightPdf > 0. && lit) {
            float4 comp;
            float f = bsdfF(B, wo, wi, &comp);
            if (f != 0.f && !bvhOccludedFast(bv, sh)) {
                float bp = bsdfPdf(B, wo, wi);
                float weight = powerHeuristic(lightPdf, bp);
                T0 = comp; T1.x = fabsf(dot3(wi, B.nn)) * weight / lightPdf;
            }
        }
        // ---- sample the BSDF (integrator.cpp:134-153)
        float bp;
        float4 comp2;
        float f = bsdfSample(B, wo, bc, bu0, bu1, &wi, &bp, &comp2);
        if (f != 0.f && bp > 0.) {
            float lp = sphereLightPdf(L, p, wi);
            if (lp != 0.) {
                float weight = powerHeuristic(bp, lp);
                DRay ray; ray.o = p; ray.d = wi; ray.mint = eps; ray.maxt = INFINITY;
                DRay r0 = ray;
                float b1, b2;
                // Li is the light's only if the closest hit IS the light (integrator.cpp:141-145).  A ray that misses the emitter
                // sphere under the very test the traversal applies to it (same function, same record, a maxt that only shrinks
                // there) cannot have it as its closest hit: the walk is skipped.  The lobe rarely points at a small emitter, so
                // this removes nearly all closest-hit walks of the BSDF-sampling term.
                float tLight;
                int prim = sphereIntersect(L, r0, &tLight) ? bvhClosest<false>(bv, ray, &b1, &b2, nullptr) : -1;
                bool emits = false;
                if (prim >= 0) {
                    float4 c = __ldg(&bv.prims[3 * prim + 2]);
                    if (__float_as_int(c.y) != 0 && (uint32_t)__float_as_int(c.z) == li) {      // lightIsect.primitive->GetAreaLight() == light
                        float ts; f3 ns;
                        if (sphereIntersectDg(L, r0, &ts, &ns)) {
                            if (L.reverseOrientation != 0.f) ns = mul3(ns, -1.f);
                            emits = dot3(ns, mul3(wi, -1.f)) > 0.f;                             // DiffuseAreaLight::L (lights/diffuse.h:51-53)
                        }
                    }
                }
                if (emits) { T2 = comp2; T1.y = fabsf(dot3(wi, B.nn)); T1.z = weight; T1.w = bp; }
            }
        }
    }
    terms[3 * t] = T0; terms[3 * t + 1] = T1; terms[3 * t + 2] = T2;
}

// ---------------------------------------------------------------------------------------------------------------
// spectral sum + film
// ---------------------------------------------------------------------------------------------------------------
#define FS_WARPS 8
// value of the BSDF in one spectral channel from the factors k_direct_terms stored
__device__ __forceinline__ float lobeValue(float4 c, float Kr, float Kt) {
    if (c.z == -1.f) return (c.x * Kt) * c.y;          // MicrofacetTransmission::f: a * T * (1 - F)
    if (c.z == -2.f) return c.x;                       // Lambertian::f
    return ((Kr * c.x) * c.y) * c.z / c.w;             // Microfacet::f: R * D * G * F / (4 cos cos)
}
// UniformSampleAllLights (integrator.cpp:47-75) of one camera sample in one spectral channel, from k_direct_terms' factors
__device__ __forceinline__ float directSum(const skc_light *__restrict__ lights, const FrameLayout &lay, const float4 *__restrict__ terms,
                                           uint64_t s, int lane, float Kr, float Kt, const float4 *__restrict__ termsEnv, const float *__restrict__ illum) {
    float Ldir = 0.f;
    uint64_t t = s * lay.slots;
    for (uint32_t i = 0; i < lay.nLights; ++i) {
        float Le = __ldg(&lights[i].Lemit[lane]);
        const int kind = lightKind(lights[i]);
        const bool delta = kind >= 1 && kind <= 3;
        float Ld = 0.f;
        for (uint32_t j = 0; j < lay.nS[i]; ++j, ++t) {
            float4 T0 = __ldg(&terms[3 * t]), T1 = __ldg(&terms[3 * t + 1]), T2 = __ldg(&terms[3 * t + 2]);
            float est = 0.f;
            if (kind == 4) {      // infinite light: the two terms carry their own radiance (the map's RGB -> Spectrum(.., SPECTRUM_ILLUMINANT))
                if (T0.w != 0.f) { const float4 c = __ldg(&termsEnv[2 * t]); est += (lobeValue(T0, Kr, Kt) * envSpectrum(illum, c.x, c.y, c.z, lane)) * T1.x; }
                if (T2.w != 0.f) { const float4 c = __ldg(&termsEnv[2 * t + 1]); est += (((lobeValue(T2, Kr, Kt) * envSpectrum(illum, c.x, c.y, c.z, lane)) * T1.y) * T1.z) / T1.w; }
                Ld += est;
                continue;
            }
            if (delta) {                                                   // Li = Intensity * Falloff / DistanceSquared (lights/point.cpp:55, spot.cpp:58)
                if (T0.w != 0.f) est += (lobeValue(T0, Kr, Kt) * ((Le * T1.z) / T1.y)) * T1.x;
                Ld += est;
                continue;
            }
            if (T0.w != 0.f) est += (lobeValue(T0, Kr, Kt) * Le) * T1.x;
            if (T2.w != 0.f) est += (((lobeValue(T2, Kr, Kt) * Le) * T1.y) * T1.z) / T1.w;
            Ld += est;
        }
        Ldir += Ld / (float)lay.nS[i];
    }
    return Ldir;
}

// The direct-lighting sum of every sample as its own pass (prefetch flow: it runs on the prefetch stream beside Mo, and
// k_frame_sum then only adds the row): Ldir[s][lane], zero for samples that hit nothing.
__global__ void __launch_bounds__(FS_WARPS * 32) k_direct_sum(const skc_light *__restrict__ lights, FrameLayout lay, const FrameConsts *__restrict__ C,
                                                             const skc_hit *__restrict__ hits, const float4 *__restrict__ terms, uint64_t n,
                                                             float *__restrict__ Ldir, const float4 *__restrict__ termsEnv, const float *__restrict__ illum) {
    int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint64_t s = (uint64_t)blockIdx.x * FS_WARPS + warp;
    if (s >= n || lane >= SKC_NS) return;
    float v = 0.f;
    if (__float_as_int(__ldg(reinterpret_cast<const float *>(hits + s)))) v = directSum(lights, lay, terms, s, lane, C->Kr[lane], C->Kt[lane], termsEnv, illum);
    Ldir[s * SKC_NS + lane] = v;
}

__global__ void __launch_bounds__(FS_WARPS * 32) k_frame_sum(BvhView bv, const skc_light *__restrict__ lights, FrameLayout lay,
                                                            const FrameConsts *__restrict__ C, const skc_hit *__restrict__ hits,
                                                            const float4 *__restrict__ terms, const float2 *__restrict__ xy,
                                                            const uint32_t *__restrict__ order, uint64_t first,
                                                            uint64_t n, float *__restrict__ Lrows, float4 *__restrict__ film,
                                                            const float *__restrict__ LdirRows, const float4 *__restrict__ termsEnv, EnvMapDev env, int envLight,
                                                            const skc_ray *__restrict__ rays) {
    __shared__ float sL[FS_WARPS][32];
    int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint64_t s = (uint64_t)blockIdx.x * FS_WARPS + warp;
    if (s >= n) return;
    const float *hp = reinterpret_cast<const float *>(hits + s);
    float Lc = 0.f;
    if (!__float_as_int(__ldg(hp)) && envLight < 0) {
        // A camera ray that leaves a scene without an infinite light: L = 0 exactly, so the sanity checks pass, XYZ = 0, and
        // AddSample adds wt * 0 to the pixel's X, Y, Z (no change to any float) and wt to its weight sum -- only that is done.
        if (lane < SKC_NS) Lrows[s * SKC_NS + lane] = 0.f;
        if (lane == 0) {
            float2 p = xy[renderToStream(order, lay.spp, first + s)];
            float dimageX = p.x - 0.5f, dimageY = p.y - 0.5f;
            int x0 = (int)ceilf(dimageX - C->filterXWidth), x1 = (int)floorf(dimageX + C->filterXWidth);
            int y0 = (int)ceilf(dimageY - C->filterYWidth), y1 = (int)floorf(dimageY + C->filterYWidth);
            x0 = max(x0, 0); x1 = min(x1, C->xres - 1); y0 = max(y0, 0); y1 = min(y1, C->yres - 1);
            for (int yy2 = y0; yy2 <= y1; ++yy2) {
                float fy = fabsf((yy2 - dimageY) * C->invYWidth * 16);
                int iy = min((int)floorf(fy), 15);
                for (int xx = x0; xx <= x1; ++xx) {
                    float fx = fabsf((xx - dimageX) * C->invXWidth * 16);
                    int ix = min((int)floorf(fx), 15);
                    atomicAdd(reinterpret_cast<float *>(film + (size_t)yy2 * C->xres + xx) + 3, C->filterTable[iy * 16 + ix]);
                }
            }
        }
        return;
    }
    if (__float_as_int(__ldg(hp)) && lane < SKC_NS) {
        int prim = __float_as_int(__ldg(hp + 1));
        float4 c = __ldg(&bv.prims[3 * prim + 2]);
        if (__float_as_int(c.y) != 0) {      // emitter: L += isect.Le(wo) (multipolesubsurface.cpp:260)
            f3 nn = mk3(__ldg(hp + 7), __ldg(hp + 8), __ldg(hp + 9)), wo = mk3(__ldg(hp + 10), __ldg(hp + 11), __ldg(hp + 12));
            if (dot3(nn, wo) > 0.f) Lc += bv.spheres[__float_as_int(c.z)].Lemit[lane];
        }
        Lc += Lrows[s * SKC_NS + lane];       // subsurface term (0 for non-subsurface hits)
        // UniformSampleAllLights (integrator.cpp:47-75): summed here, or already by k_direct_sum (same arithmetic, same order)
        float Ldir = LdirRows ? __ldg(&LdirRows[s * SKC_NS + lane]) : directSum(lights, lay, terms, s, lane, C->Kr[lane], C->Kt[lane], termsEnv, env.illum);
        Lc += Ldir;
    } else if (envLight >= 0 && lane < SKC_NS && !__float_as_int(__ldg(hp))) {
        // a camera ray that leaves the scene sees the infinite light (samplerrenderer.cpp:239-246: Li += lights[i]->Le(ray))
        const float *rp = reinterpret_cast<const float *>(rays + s);
        const f3 rgb = infiniteLightLe(lights[envLight], env, mk3(__ldg(rp + 3), __ldg(rp + 4), __ldg(rp + 5)));
        Lc += envSpectrum(env.illum, rgb.x, rgb.y, rgb.z, lane);
    }
    sL[warp][lane] = lane < SKC_NS ? Lc : 0.f;
    __syncwarp();
    // radiance sanity checks of SamplerRendererTask::Run (samplerrenderer.cpp:117-135)
    float yy = 0.f;
    bool nan = false;
    for (int i = 0; i < SKC_NS; ++i) { float v = sL[warp][i]; nan |= isnan(v); yy += C->cie[1][i] * v; }
    float y = yy * 300.f / (106.856895f * SKC_NS);
    __syncwarp();
    if (nan || y < -1e-5 || isinf(y)) { Lc = 0.f; sL[warp][lane] = 0.f; }
    __syncwarp();
    if (lane < SKC_NS) Lrows[s * SKC_NS + lane] = Lc;
    // ImageFilm::AddSample (film/image.cpp:77-137)
    float xyzc = 0.f;
    if (lane < 3) {
        for (int i = 0; i < SKC_NS; ++i) xyzc += C->cie[lane][i] * sL[warp][i];    // SampledSpectrum::ToXYZ (spectrum.h:374-386)
        xyzc *= C->xyzScale;
    }
    float X = __shfl_sync(0xffffffffu, xyzc, 0), Y = __shfl_sync(0xffffffffu, xyzc, 1), Z = __shfl_sync(0xffffffffu, xyzc, 2);
    if (lane == 0) {
        float2 p = xy[renderToStream(order, lay.spp, first + s)];
        float dimageX = p.x - 0.5f, dimageY = p.y - 0.5f;
        int x0 = (int)ceilf(dimageX - C->filterXWidth), x1 = (int)floorf(dimageX + C->filterXWidth);
        int y0 = (int)ceilf(dimageY - C->filterYWidth), y1 = (int)floorf(dimageY + C->filterYWidth);
        x0 = max(x0, 0); x1 = min(x1, C->xres - 1); y0 = max(y0, 0); y1 = min(y1, C->yres - 1);
        for (int yy2 = y0; yy2 <= y1; ++yy2) {
            float fy = fabsf((yy2 - dimageY) * C->invYWidth * 16);
            int iy = min((int)floorf(fy), 15);
            for (int xx = x0; xx <= x1; ++xx) {
                float fx = fabsf((xx - dimageX) * C->invXWidth * 16);
                int ix = min((int)floorf(fx), 15);
                float wt = C->filterTable[iy * 16 + ix];
                float *px = reinterpret_cast<float *>(film + (size_t)yy2 * C->xres + xx);
                atomicAdd(px, wt * X); atomicAdd(px + 1, wt * Y); atomicAdd(px + 2, wt * Z); atomicAdd(px + 3, wt);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------------
static inline int floor2Int(float v) { return (int)floorf(v); }
static inline int ceil2Int(float v) { return (int)ceilf(v); }
static inline float lerpf(float t, float a, float b) { return (1.f - t) * a + t * b; }

extern "C" int skc_frame_setup(skc_ctx *ctx, const skc_frame_desc *d, uint64_t *nSamplesOut) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!d || d->xres <= 0 || d->yres <= 0 || d->spp <= 0 || (d->spp & (d->spp - 1)) || d->nTasks <= 0)
        return skc_fail(ctx, "skc_frame_setup: bad frame description (pixelsamples must be a power of two)");
    if (!ctx->nLights || ctx->nLights > FR_MAX_LIGHTS) return skc_fail(ctx, "skc_frame_setup: %u lights (1..%d supported; skc_upload_lights first)", ctx->nLights, FR_MAX_LIGHTS);
    skc_frame_release(ctx);
    skc_frame_state *f = new skc_frame_state();
    ctx->frame = f;
    f->desc = *d;
    // ---- Sample layout: MultipoleSubsurfaceIntegrator::RequestSamples (multipolesubsurface.cpp:154-168) then the volume
    // integrator's arrays; LightSampleOffsets / BSDFSampleOffsets each add one 1-D and one 2-D array of nSamples
    FrameLayout &L = f->lay;
    memset(&L, 0, sizeof L);
    L.spp = (uint32_t)d->spp; L.nLights = ctx->nLights;
    uint32_t off = 0;
    for (uint32_t i = 0; i < ctx->nLights; ++i) {
        uint32_t ns = ctx->lightNS[i];          // LDSampler::RoundSize = RoundUpPow2 (lowdiscrepancy.h:52)
        if ((uint64_t)ns * L.spp > FR_MAX_VALUES) return skc_fail(ctx, "skc_frame_setup: nsamples %u x pixelsamples %u exceeds %d", ns, L.spp, FR_MAX_VALUES);
        L.nS[i] = ns; L.slots += ns;
        L.size1D[L.n1D++] = ns; L.size1D[L.n1D++] = ns;
        L.size2D[L.n2D++] = ns; L.size2D[L.n2D++] = ns;
    }
    for (int e = 0; e < d->extra1D; ++e) { if (L.n1D >= FR_MAX_ARRAYS) return skc_fail(ctx, "skc_frame_setup: too many sample arrays"); L.size1D[L.n1D++] = 1; }
    for (uint32_t i = 0; i < ctx->nLights; ++i) { L.off1L[i] = off; off += L.nS[i]; L.off1B[i] = off; off += L.nS[i]; }
    for (int e = 0; e < d->extra1D; ++e) off += 1;
    for (uint32_t i = 0; i < ctx->nLights; ++i) { L.off2L[i] = off; off += 2 * L.nS[i]; L.off2B[i] = off; off += 2 * L.nS[i]; }
    L.floats = off;
    // MT words per pixel (montecarlo.h:312-331): scrambles + nPixel * nSamples + nPixel draws per array
    L.draws = 2 * (2 + L.spp + L.spp) + (1 + L.spp + L.spp);
    for (uint32_t j = 0; j < L.n1D; ++j) L.draws += 1 + L.spp * L.size1D[j] + L.spp;
    for (uint32_t j = 0; j < L.n2D; ++j) L.draws += 2 + L.spp * L.size2D[j] + L.spp;
    // ---- tasks: Film::GetSampleExtent (film/image.cpp:158-167), Sampler::ComputeSubWindow (core/sampler.cpp:56-75),
    // emitted in the order SamplerRenderer::Render enqueues them (taskNum = nTasks-1-i, samplerrenderer.cpp:211-215)
    int xs = floor2Int(0 + 0.5f - d->filterXWidth), xe = ceil2Int(0 + 0.5f + d->xres + d->filterXWidth);
    int ys = floor2Int(0 + 0.5f - d->filterYWidth), ye = ceil2Int(0 + 0.5f + d->yres + d->filterYWidth);
    int count = d->nTasks;
    int dx = xe - xs, dy = ye - ys, nx = count, ny = 1;
    while ((nx & 0x1) == 0 && 2 * dx * ny < dy * nx) { nx >>= 1; ny <<= 1; }
    f->pixBase.push_back(0);
    for (int i = 0; i < count; ++i) {
        int num = count - 1 - i;
        int xo = num % nx, yo = num / nx;
        float tx0 = float(xo) / float(nx), tx1 = float(xo + 1) / float(nx);
        float ty0 = float(yo) / float(ny), ty1 = float(yo + 1) / float(ny);
        int x0 = floor2Int(lerpf(tx0, (float)xs, (float)xe)), x1 = floor2Int(lerpf(tx1, (float)xs, (float)xe));
        int y0 = floor2Int(lerpf(ty0, (float)ys, (float)ye)), y1 = floor2Int(lerpf(ty1, (float)ys, (float)ye));
        if (x0 == x1 || y0 == y1) continue;                   // GetSubSampler returns NULL
        // several GPUs: windows dealt round-robin along each row of windows, every row shifted by 3 against the one above
        // (plain emission order would give each rank whole columns of the film whenever nx is a multiple of the world size)
        if (d->taskWorld > 1 && (xo + 3 * yo) % d->taskWorld != d->taskRank) continue;
        f->win.push_back(make_int4(x0, x1, y0, y1));
        f->seed.push_back((uint32_t)num);                     // RNG rng(taskNum), samplerrenderer.cpp:73
        f->pixBase.push_back(f->pixBase.back() + (uint32_t)((x1 - x0) * (y1 - y0)));
    }
    f->nPixels = f->pixBase.back();
    std::vector<uint32_t> order(f->nPixels);
    {   // Z-curve over the film (not per window: task windows are not multiples of the 4x2 block)
        auto part1by1 = [](uint32_t v) { v &= 0xffff; v = (v | (v << 8)) & 0x00ff00ffu; v = (v | (v << 4)) & 0x0f0f0f0fu; v = (v | (v << 2)) & 0x33333333u; v = (v | (v << 1)) & 0x55555555u; return v; };
        std::vector<std::pair<uint32_t, uint32_t>> keyed(f->nPixels);
        for (size_t t = 0; t < f->win.size(); ++t) {
            int4 w = f->win[t];
            uint32_t wx = (uint32_t)(w.y - w.x), np = f->pixBase[t + 1] - f->pixBase[t];
            for (uint32_t k = 0; k < np; ++k) {
                uint32_t x = (uint32_t)(w.x - xs) + k % wx, y = (uint32_t)(w.z - ys) + k / wx;
                keyed[f->pixBase[t] + k] = std::make_pair(part1by1(x) | (part1by1(y) << 1), f->pixBase[t] + k);
            }
        }
        // the keys are distinct (one per film position): two stable 16-bit counting passes give the sorted order (std::sort of the
        // 0.9 M pairs of a 1280x720 film took 35 ms of every skc_frame_setup)
        std::vector<std::pair<uint32_t, uint32_t>> tmp(keyed.size());
        for (int pass = 0; pass < 2; ++pass) {
            std::vector<uint32_t> start(65537, 0u);
            const int shift = 16 * pass;
            for (const auto &e : keyed) ++start[((e.first >> shift) & 0xffffu) + 1];
            for (int b = 0; b < 65536; ++b) start[b + 1] += start[b];
            for (const auto &e : keyed) tmp[start[(e.first >> shift) & 0xffffu]++] = e;
            keyed.swap(tmp);
        }
        for (size_t k = 0; k < keyed.size(); ++k) order[k] = keyed[k].second;
    }
    f->nSamples = f->nPixels * L.spp;
    if (!f->nPixels) return skc_fail(ctx, "skc_frame_setup: empty sample extent");
    size_t nT = f->win.size();
    SKC_CUDA(cudaMalloc((void **)&f->dWin, nT * sizeof(int4)));
    SKC_CUDA(cudaMalloc((void **)&f->dSeed, nT * sizeof(uint32_t)));
    SKC_CUDA(cudaMalloc((void **)&f->dPixBase, (nT + 1) * sizeof(uint32_t)));
    SKC_CUDA(cudaMalloc((void **)&f->dOrder, f->nPixels * sizeof(uint32_t)));
    SKC_CUDA(cudaMemcpyAsync(f->dOrder, order.data(), f->nPixels * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    SKC_CUDA(cudaMalloc((void **)&f->dXY, f->nSamples * sizeof(float2)));
    SKC_CUDA(cudaMalloc((void **)&f->dArrays, f->nSamples * L.floats * sizeof(float)));
    SKC_CUDA(cudaMalloc((void **)&f->dFilm, (size_t)d->xres * d->yres * sizeof(float4)));
    SKC_CUDA(cudaMalloc((void **)&f->dConsts, sizeof(FrameConsts)));
    FrameConsts hc;
    memset(&hc, 0, sizeof hc);
    memcpy(hc.c2w, d->cameraToWorld, 64); memcpy(hc.r2c, d->rasterToCamera, 64);
    memcpy(hc.filterTable, d->filterTable, sizeof hc.filterTable);
    for (int k = 0; k < 3; ++k) for (int i = 0; i < SKC_NS; ++i) hc.cie[k][i] = d->cieXYZ[k][i];
    for (int i = 0; i < SKC_NS; ++i) { hc.Kr[i] = d->Kr[i]; hc.Kt[i] = d->Kt[i]; }
    hc.xyzScale = d->xyzScale; hc.filterXWidth = d->filterXWidth; hc.filterYWidth = d->filterYWidth;
    hc.invXWidth = 1.f / d->filterXWidth; hc.invYWidth = 1.f / d->filterYWidth;      // Filter::invXWidth (core/filter.h:50-52)
    hc.xres = d->xres; hc.yres = d->yres;
    SKC_CUDA(cudaMemcpyAsync(f->dWin, f->win.data(), nT * sizeof(int4), cudaMemcpyHostToDevice, ctx->stream));
    SKC_CUDA(cudaMemcpyAsync(f->dSeed, f->seed.data(), nT * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    SKC_CUDA(cudaMemcpyAsync(f->dPixBase, f->pixBase.data(), (nT + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    SKC_CUDA(cudaMemcpyAsync(f->dConsts, &hc, sizeof hc, cudaMemcpyHostToDevice, ctx->stream));
    SKC_CUDA(cudaMemsetAsync(f->dFilm, 0, (size_t)d->xres * d->yres * sizeof(float4), ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    if (nSamplesOut) *nSamplesOut = f->nSamples;
    return 0;
}

extern "C" uint32_t skc_frame_sample_floats(skc_ctx *ctx) { return ctx && ctx->frame ? ctx->frame->lay.floats : 0; }

extern "C" int skc_frame_samples(skc_ctx *ctx) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    skc_frame_state *f = ctx->frame;
    if (!f) return skc_fail(ctx, "skc_frame_samples: skc_frame_setup first");
    uint32_t cap = f->lay.draws + 6 * f->lay.spp;        // words a pixel can take from its task's stream
    uint32_t *dWords = (uint32_t *)skc_scratch(ctx, 12, f->nPixels * cap * sizeof(uint32_t));
    uint64_t *dPixOff = (uint64_t *)skc_scratch(ctx, 15, f->nPixels * sizeof(uint64_t));
    if (!dWords || !dPixOff) return 1;
    if (!ctx->bvhNodes) return skc_fail(ctx, "skc_frame_samples: no BVH uploaded (the sampler stream depends on which camera rays hit)");
    if (ctx->resRaysCount != f->nSamples) {
        if (ctx->resRays) { cudaFree(ctx->resRays); ctx->resRays = nullptr; }
        ctx->resRaysCount = 0;
        SKC_CUDA(cudaMalloc((void **)&ctx->resRays, f->nSamples * sizeof(skc_ray)));
        ctx->resRaysCount = f->nSamples;
    }
    uint32_t nT = (uint32_t)f->win.size();
    skc_stage_begin(ctx);
    k_mt_words<<<nT, 256, 0, ctx->stream>>>(f->dSeed, f->dPixBase, cap, dWords);
    SKC_LAUNCHED("k_mt_words");
    uint32_t maxPix = 0;
    for (size_t t = 0; t < f->win.size(); ++t) maxPix = std::max(maxPix, f->pixBase[t + 1] - f->pixBase[t]);
    size_t chainSmem = (size_t)maxPix * 2 * sizeof(uint32_t);
    if (chainSmem > 200 * 1024) return skc_fail(ctx, "skc_frame_samples: a render task of %u pixels exceeds the sampler kernel's shared memory", maxPix);
    if (chainSmem > 48 * 1024) SKC_CUDA(cudaFuncSetAttribute(k_ld_chain, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)chainSmem));
    k_ld_chain<<<nT, LC_THREADS, chainSmem, ctx->stream>>>(skc_bvh_view(ctx), f->dConsts, f->lay, f->dWin, f->dPixBase, cap, dWords, dPixOff, f->dXY);
    SKC_LAUNCHED("k_ld_chain");
    k_ld_pixel<<<(unsigned)((f->nPixels + 127) / 128), 128, 0, ctx->stream>>>(f->lay, f->dWin, f->dPixBase, nT, dWords, dPixOff, f->dXY, f->dArrays, f->lazyArrays ? 1 : 0);
    SKC_LAUNCHED("k_ld_pixel");
    k_camera_rays<<<(unsigned)((f->nSamples + 255) / 256), 256, 0, ctx->stream>>>(f->dConsts, f->dXY, f->nSamples, f->dOrder, f->lay.spp, ctx->resRays);
    SKC_LAUNCHED("k_camera_rays");
    skc_stage_end(ctx, 9);
    f->samplesReady = true;
    return 0;
}

extern "C" int skc_frame_get_samples(skc_ctx *ctx, uint64_t first, uint64_t count, float *xy, float *arrays, skc_ray *rays) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    skc_frame_state *f = ctx->frame;
    if (!f || !f->samplesReady) return skc_fail(ctx, "skc_frame_get_samples: skc_frame_samples first");
    if (first + count > f->nSamples) return skc_fail(ctx, "skc_frame_get_samples: bad range");
    if (!count) return 0;
    if (xy) SKC_CUDA(cudaMemcpyAsync(xy, f->dXY + first, count * sizeof(float2), cudaMemcpyDeviceToHost, ctx->stream));
    if (arrays) SKC_CUDA(cudaMemcpyAsync(arrays, f->dArrays + first * f->lay.floats, count * f->lay.floats * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    if (rays) {   // stream order (the resident rays are in render order)
        skc_ray *tmp = (skc_ray *)skc_scratch(ctx, 0, count * sizeof(skc_ray));
        if (!tmp) return 1;
        k_camera_rays<<<(unsigned)((count + 255) / 256), 256, 0, ctx->stream>>>(f->dConsts, f->dXY + first, count, nullptr, f->lay.spp, tmp);
        SKC_LAUNCHED("k_camera_rays");
        SKC_CUDA(cudaMemcpyAsync(rays, tmp, count * sizeof(skc_ray), cudaMemcpyDeviceToHost, ctx->stream));
    }
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// Scene::Intersect for the camera rays and EstimateDirect's terms (everything of Li that does not need the octree)
// index of the scene's infinite light (-1: none) and, with one, the per-term RGB array of `count` samples (scratch slot 22)
static int frameEnv(skc_ctx *ctx, const FrameLayout &lay, uint64_t count, int *light, float4 **termsEnv) {
    *light = -1; *termsEnv = nullptr;
    for (uint32_t j = 0; j < ctx->nLights && j < 16; ++j) if (ctx->lightKinds[j] == 4) *light = (int)j;
    if (*light < 0) return 0;
    if (!ctx->env.rgb) return skc_fail(ctx, "the scene has an infinite light but no environment map was uploaded (skc_upload_envmap)");
    *termsEnv = (float4 *)skc_scratch(ctx, 22, count * lay.slots * 2 * sizeof(float4));
    return *termsEnv ? 0 : 1;
}

static int frameHitsAndDirect(skc_ctx *ctx, skc_frame_state *f, uint64_t first, uint64_t count, skc_hit *dHits, float4 *dFrame, float4 *dTerms,
                              cudaEvent_t hitsDone = nullptr, float *dLdir = nullptr) {
    int envLight; float4 *dTermsEnv;
    if (frameEnv(ctx, f->lay, count, &envLight, &dTermsEnv)) return 1;
    const FrameLayout &lay = f->lay;
    skc_stage_begin(ctx);
    if (skc_trace_closest_dev(ctx, ctx->resRays + first, count, dHits, nullptr, dFrame)) return 1;
    if (ctx->albedoTex.rgb || ctx->bumpTex.rgb) {      // "texture albedo": the albedo of every skin hit, for the Li epilogue of the Mo kernel; "texture bumpmap": its bumped shading frame
        float4 *dAlb = ctx->albedoTex.rgb ? (float4 *)skc_scratch(ctx, 23, count * sizeof(float4)) : nullptr;
        if (ctx->albedoTex.rgb && !dAlb) return 1;
        k_hit_textures<<<(unsigned)((count + 127) / 128), 128, 0, ctx->stream>>>(skc_bvh_view(ctx), skc_mesh_view(ctx), ctx->albedoTex, ctx->bumpTex, f->dConsts, f->dXY, f->dOrder, lay.spp, first,
                                                                              1.f / sqrtf((float)lay.spp), ctx->resRays + first, dHits, count, dAlb, dFrame);
        SKC_LAUNCHED("k_hit_textures");
    }
    skc_stage_end(ctx, 3);
    if (hitsDone) SKC_CUDA(cudaEventRecord(hitsDone, ctx->stream));
    RhoBxdf mf;
    float rms = f->desc.roughness < 1e-3f ? 1e-3f : f->desc.roughness;      // Beckmann::Beckmann (reflection.h:509-514)
    mf.rms2 = rms * rms; mf.rcpRMS2 = 1 / mf.rms2; mf.etaI = 1.f; mf.etaT = f->desc.ior; mf.fixedFresnel = 0;
    int hasR = 0, hasT = 0;                 // Spectrum::IsBlack (layeredskin.cpp:156-161)
    for (int i = 0; i < SKC_NS; ++i) { hasR |= f->desc.Kr[i] != 0.f; hasT |= f->desc.Kt[i] != 0.f; }
    skc_stage_begin(ctx);
    uint64_t nThreads = count * lay.slots;
    if (envLight >= 0)
        k_direct_terms<true><<<(unsigned)((nThreads + 127) / 128), 128, 0, ctx->stream>>>(skc_bvh_view(ctx), ctx->lights, lay, mf, f->desc.ior, f->desc.matteKd, hasR, hasT,
                                                                                          dHits, dFrame, f->dArrays, f->dOrder, first, count, dTerms, ctx->env, dTermsEnv);
    else
        k_direct_terms<false><<<(unsigned)((nThreads + 127) / 128), 128, 0, ctx->stream>>>(skc_bvh_view(ctx), ctx->lights, lay, mf, f->desc.ior, f->desc.matteKd, hasR, hasT,
                                                                                           dHits, dFrame, f->dArrays, f->dOrder, first, count, dTerms, EnvMapDev(), nullptr);
    SKC_LAUNCHED("k_direct_terms");
    if (dLdir) {
        k_direct_sum<<<(unsigned)((count + FS_WARPS - 1) / FS_WARPS), FS_WARPS * 32, 0, ctx->stream>>>(ctx->lights, lay, f->dConsts, dHits, dTerms, count, dLdir, dTermsEnv, ctx->env.illum);
        SKC_LAUNCHED("k_direct_sum");
    }
    skc_stage_end(ctx, 10);
    return 0;
}

// The sampler stream, the camera rays, their closest hits and the direct-lighting terms depend on the scene and the film
// only -- not on Preprocess.  skc_frame_prefetch enqueues them for the whole frame on the context's second stream and
// returns at once; the caller then runs Preprocess (tessellation, irradiance pass, octree) on the main stream, and
// skc_frame_render(0, nSamples, ...) waits for the prefetched part on the device and goes on with Mo and the film.  The
// latency-bound sampler kernels and the BVH walks of direct lighting fill issue slots the irradiance kernels leave empty.
// Same kernels, same inputs: bit-identical results.  Stage times of the prefetched part: "samples" = the whole prefetch as
// timed on its own stream (it overlaps the other stages), "closest" and "direct" = 0.
extern "C" int skc_frame_prefetch(skc_ctx *ctx) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    skc_frame_state *f = ctx->frame;
    if (!f) return skc_fail(ctx, "skc_frame_prefetch: skc_frame_setup first");
    if (!f->evP0) { SKC_CUDA(cudaEventCreate(&f->evP0)); SKC_CUDA(cudaEventCreate(&f->evP1)); SKC_CUDA(cudaEventCreateWithFlags(&f->evDone, cudaEventDisableTiming));
                     SKC_CUDA(cudaEventCreateWithFlags(&f->evHits, cudaEventDisableTiming)); }
    const uint64_t count = f->nSamples;
    skc_hit *dHits = (skc_hit *)skc_scratch(ctx, 17, count * sizeof(skc_hit));
    float4 *dFrame = (float4 *)skc_scratch(ctx, 13, count * 2 * sizeof(float4));
    float4 *dTerms = (float4 *)skc_scratch(ctx, 14, count * f->lay.slots * 3 * sizeof(float4));
    float *dLdir = (float *)skc_scratch(ctx, 19, count * SKC_NS * sizeof(float));
    if (!dHits || !dFrame || !dTerms || !dLdir) return 1;
    SKC_CUDA(cudaEventRecord(f->evDone, ctx->stream));                       // earlier consumers of these buffers
    SKC_CUDA(cudaStreamWaitEvent(ctx->prefetchStream, f->evDone, 0));
    cudaStream_t mainStream = ctx->stream;
    ctx->stream = ctx->prefetchStream; ctx->stageOff = true;
    cudaEventRecord(f->evP0, ctx->stream);
    f->lazyArrays = !getenv("SKC_EAGER_ARRAYS");
    int rc = skc_frame_samples(ctx);
    f->lazyArrays = false;
    if (!rc) rc = frameHitsAndDirect(ctx, f, 0, count, dHits, dFrame, dTerms, f->evHits, dLdir);
    cudaEventRecord(f->evP1, ctx->stream);
    ctx->stream = mainStream; ctx->stageOff = false;
    if (rc) return 1;
    f->prefetched = true;
    return 0;
}

extern "C" int skc_frame_render(skc_ctx *ctx, uint64_t first, uint64_t count, float maxError, float mix, float *L30) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    skc_frame_state *f = ctx->frame;
    if (!f || !f->samplesReady) return skc_fail(ctx, "skc_frame_render: skc_frame_samples first");
    if (first + count > f->nSamples) return skc_fail(ctx, "skc_frame_render: bad range");
    if (!count) return 0;
    const FrameLayout &lay = f->lay;
    skc_hit *dHits = (skc_hit *)skc_scratch(ctx, 17, count * sizeof(skc_hit));
    float *dL = (float *)skc_scratch(ctx, 2, count * SKC_NS * sizeof(float));
    float4 *dFrame = (float4 *)skc_scratch(ctx, 13, count * 2 * sizeof(float4));
    float4 *dTerms = (float4 *)skc_scratch(ctx, 14, count * lay.slots * 3 * sizeof(float4));
    if (!dHits || !dL || !dFrame || !dTerms) return 1;
    const bool usePrefetch = f->prefetched && first == 0 && count == f->nSamples;
    // hits, frames and direct-lighting terms are (being) made on prefetchStream: Mo needs the hits only, the spectral sum the terms too
    if (f->prefetched) SKC_CUDA(cudaStreamWaitEvent(ctx->stream, usePrefetch ? f->evHits : f->evP1, 0));
    f->prefetched = false;
    if (!usePrefetch) if (frameHitsAndDirect(ctx, f, first, count, dHits, dFrame, dTerms)) return 1;
    if (ctx->otGeo && ctx->otNodes) {
        ctx->hitAlbedo = ctx->albedoTex.rgb ? (const float4 *)skc_scratch(ctx, 23, count * sizeof(float4)) : nullptr;      // filled by frameHitsAndDirect
        const int rcLi = skc_li_fused_dev(ctx, dHits, count, maxError, mix, dL);
        ctx->hitAlbedo = nullptr;
        if (rcLi) return 1;
    } else {
        SKC_CUDA(cudaMemsetAsync(dL, 0, count * SKC_NS * sizeof(float), ctx->stream));    // no octree: `if (bssrdf && octree)` (multipolesubsurface.cpp:267)
    }
    if (usePrefetch) SKC_CUDA(cudaStreamWaitEvent(ctx->stream, f->evP1, 0));
    int envLight; float4 *dTermsEnv;
    if (frameEnv(ctx, lay, count, &envLight, &dTermsEnv)) return 1;
    skc_stage_begin(ctx);
    k_frame_sum<<<(unsigned)((count + FS_WARPS - 1) / FS_WARPS), FS_WARPS * 32, 0, ctx->stream>>>(skc_bvh_view(ctx), ctx->lights, lay, f->dConsts, dHits, dTerms,
                                                                                                 f->dXY, f->dOrder, first, count, dL, f->dFilm,
                                                                                                 usePrefetch ? (const float *)skc_scratch(ctx, 19, count * SKC_NS * sizeof(float)) : nullptr,
                                                                                                 dTermsEnv, ctx->env, envLight, ctx->resRays + first);
    SKC_LAUNCHED("k_frame_sum");
    skc_stage_end(ctx, 11);
    if (usePrefetch) {      // the prefetched part, as timed on its own stream
        float ms = 0.f;
        cudaEventElapsedTime(&ms, f->evP0, f->evP1);
        if (ctx->stageAccumulate) ctx->stageMs[9] += ms; else ctx->stageMs[9] = ms;
        if (!ctx->stageAccumulate) ctx->stageMs[3] = ctx->stageMs[10] = 0.f;
    }
    if (L30) {
        SKC_CUDA(cudaMemcpyAsync(L30, dL, count * SKC_NS * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
        SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return 0;
}

extern "C" int skc_frame_film_clear(skc_ctx *ctx) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    skc_frame_state *f = ctx->frame;
    if (!f) return skc_fail(ctx, "skc_frame_film_clear: skc_frame_setup first");
    SKC_CUDA(cudaMemsetAsync(f->dFilm, 0, (size_t)f->desc.xres * f->desc.yres * sizeof(float4), ctx->stream));
    return 0;
}
extern "C" void *skc_frame_film_dev(skc_ctx *ctx) { return ctx && ctx->frame ? ctx->frame->dFilm : nullptr; }
// the film as floats, for the reduce of skc_comm.cu
void *skc_frame_film_floats(skc_ctx *ctx, uint64_t *floats) {
    if (!ctx || !ctx->frame || !ctx->frame->dFilm) return nullptr;
    *floats = (uint64_t)ctx->frame->desc.xres * ctx->frame->desc.yres * 4;
    return ctx->frame->dFilm;
}
extern "C" int skc_frame_film_get(skc_ctx *ctx, float *pixels4) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    skc_frame_state *f = ctx->frame;
    if (!f) return skc_fail(ctx, "skc_frame_film_get: skc_frame_setup first");
    SKC_CUDA(cudaMemcpyAsync(pixels4, f->dFilm, (size_t)f->desc.xres * f->desc.yres * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// ImageFilm::WriteImage's pixel arithmetic (film/image.cpp:178-213): XYZ -> RGB (core/spectrum.h:51-55), divide by the weight
// sum, clamp at 0.  fp32, no contraction: the same values as host/skin_integrator.cpp skh_film_to_rgb.
__global__ void __launch_bounds__(256) k_film_to_rgb(const float4 *__restrict__ film, uint32_t n, float *__restrict__ rgb) {
    uint32_t i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    float4 p = __ldg(&film[i]);
    float r = 3.240479f * p.x - 1.537150f * p.y - 0.498535f * p.z;
    float g = -0.969256f * p.x + 1.875991f * p.y + 0.041556f * p.z;
    float b = 0.055648f * p.x - 0.204043f * p.y + 1.057311f * p.z;
    if (p.w != 0.f) {
        float invWt = 1.f / p.w;
        r = fmaxf(0.f, r * invWt); g = fmaxf(0.f, g * invWt); b = fmaxf(0.f, b * invWt);
    }
    rgb[3 * i] = r; rgb[3 * i + 1] = g; rgb[3 * i + 2] = b;
}

extern "C" int skc_frame_film_rgb(skc_ctx *ctx, float *rgb3) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    skc_frame_state *f = ctx->frame;
    if (!f) return skc_fail(ctx, "skc_frame_film_rgb: skc_frame_setup first");
    uint32_t n = (uint32_t)f->desc.xres * (uint32_t)f->desc.yres;
    float *d = (float *)skc_scratch(ctx, 18, (size_t)n * 3 * sizeof(float));
    if (!d) return 1;
    k_film_to_rgb<<<(n + 255) / 256, 256, 0, ctx->stream>>>(f->dFilm, n, d);
    SKC_LAUNCHED("k_film_to_rgb");
    SKC_CUDA(cudaMemcpyAsync(rgb3, d, (size_t)n * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

extern "C" int skc_frame_get_order(skc_ctx *ctx, uint64_t first, uint64_t count, uint64_t *streamIndex) {
    if (!ctx) return 1;
    skc_frame_state *f = ctx->frame;
    if (!f) return skc_fail(ctx, "skc_frame_get_order: skc_frame_setup first");
    if (first + count > f->nSamples) return skc_fail(ctx, "skc_frame_get_order: bad range");
    std::vector<uint32_t> order(f->nPixels);
    cudaSetDevice(ctx->device);
    SKC_CUDA(cudaMemcpy(order.data(), f->dOrder, f->nPixels * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    uint32_t spp = f->lay.spp;
    for (uint64_t j = 0; j < count; ++j) streamIndex[j] = (uint64_t)order[(first + j) / spp] * spp + (first + j) % spp;
    return 0;
}

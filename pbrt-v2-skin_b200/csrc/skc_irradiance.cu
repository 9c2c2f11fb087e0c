// skc_irradiance.cu -- MultipoleSubsurfaceIntegrator::Preprocess, irradiance pass.
//
// Reference: integrators/multipolesubsurface.cpp:72-152 (IrradianceTask::Run),
// lights/diffuse.cpp:69-81, core/light.cpp:145-174 (ShapeSet::Sample/Pdf),
// shapes/sphere.cpp:236-273 (Sphere::Sample/Pdf), core/light.h:87-92 (SetSegment),
// core/montecarlo.h:278-301 ((0,2)-sequence), core/rng.cpp:43-110 (MT19937),
// core/multipole.cpp:458-463 (rho_hd lookup).
//
// Two kernels:
//   k_scrambles  one CTA per IrradianceTask: regenerates the task's MT19937 stream
//                (RNG(taskNum*47)) 624 words at a time in shared memory and writes the three
//                scramble words of every (point, light) pair -- the stream position depends
//                only on (point index within task, light), SURVEY.md appendix A6.
//   k_irradiance one thread per point for the geometric part (light sampling + shadow ray
//                through the BVH), then the warp transposes: lane = wavelength, and the
//                30-wide spectral accumulation runs with coalesced 120 B rows of the rho
//                table, Lemit and the output E.
#include <algorithm>
#include <utility>
#include <vector>
#include "skc_light.cuh"
#include "skc_texture.cuh"
#include "skc_internal.h"

// grid = tasks [taskFirst, taskFirst + gridDim.x); block = 256 threads
__global__ void __launch_bounds__(256) k_scrambles(uint32_t taskFirst, uint32_t nTasks, uint64_t nTotal,
                                                  uint64_t first, uint64_t count, uint32_t wordsPerPoint,
                                                  uint32_t *__restrict__ out) {
    __shared__ uint32_t mt[MT_N];
    uint32_t task = taskFirst + blockIdx.x;
    // multipolesubsurface.cpp:73-74 (the reference truncates through int)
    uint64_t idxStart = (uint64_t)(int)((uint64_t)task * nTotal / nTasks);
    uint64_t idxEnd = (uint64_t)(int)((uint64_t)(task + 1) * nTotal / nTasks);
    if (idxStart == idxEnd) return;
    uint64_t nWords = (idxEnd - idxStart) * wordsPerPoint;
    int tid = threadIdx.x;
    if (tid == 0) {   // RNG::Seed, core/rng.cpp:43-55
        uint32_t s = task * 47u;
        mt[0] = s;
        for (int i = 1; i < MT_N; ++i) { s = 1812433253u * (s ^ (s >> 30)) + (uint32_t)i; mt[i] = s; }
    }
    __syncthreads();
    for (uint64_t base = 0; base < nWords; base += MT_N) {
        // regenerate the table (core/rng.cpp:84-103) in three dependency-free phases
        uint32_t v = 0;
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
                uint64_t point = idxStart + w / wordsPerPoint;
                if (point >= first && point < first + count) {
                    uint32_t y = mt[k];       // tempering, core/rng.cpp:105-110
                    y ^= (y >> 11);
                    y ^= (y << 7) & 0x9d2c5680u;
                    y ^= (y << 15) & 0xefc60000u;
                    y ^= (y >> 18);
                    out[(point - first) * wordsPerPoint + w % wordsPerPoint] = y;
                }
            }
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------
// The pass runs as three kernels per batch of points (the first version did everything in one
// thread-per-point kernel: 93 registers + 41 KB of shared accumulators held it at 20 warps/SM and
// ncu showed it latency-bound in the BVH loop, ~50 % of the stall samples on node fetch + slab test):
//   k_irr_sample    one thread per (point, light, sample): scrambled (0,2) sample -> Sample_L geometry ->
//                   (cos, pdf) candidate, and the shadow segments that need a visibility test are
//                   compacted into a ray list (one atomic per CTA)
//   k_irr_shadow    one thread per listed shadow ray, any-hit BVH walk at full occupancy; an occluded
//                   ray clears its sample's pdf
//   k_irr_spectral  one warp per run of points, lane = wavelength: the reference's sum
//                   sum_lights (1/nS) sum_s (1 - rho_hd(cos)) * Lemit * cos / pdf  in its own order,
//                   128 B coalesced rows for rho, Lemit and E
#define IRR_BLOCK 256
#define IRR_MAX_LIGHTS 16
#define IRR_SPECTRAL_POINTS 8        // points per warp in k_irr_spectral

struct IrrSlots { uint32_t nLights, S; uint32_t nS[IRR_MAX_LIGHTS]; };   // S = sum of RoundUpPow2(nSamples)

// DELTA: the scene has point / spot / distant lights (row N4); without them the instance is the sphere-emitter code alone
template <bool COUNT, bool DELTA>
__global__ void __launch_bounds__(IRR_BLOCK, 4) k_irr_sample(const skc_light *__restrict__ lights, IrrSlots sl,
        const skc_surface_point *__restrict__ pts, const uint32_t *__restrict__ scr, uint64_t ptBase, uint32_t nPts,
        float2 *__restrict__ cosPdf, float4 *__restrict__ rayO, float4 *__restrict__ rayD, uint32_t *__restrict__ rayTarget,
        uint32_t *__restrict__ nRays, float4 *__restrict__ posArea, uint64_t outBase, unsigned long long *cnt, EnvMapDev env, float4 *__restrict__ envRgb) {
    __shared__ uint32_t sWarpCount[IRR_BLOCK / 32];
    __shared__ uint32_t sBase;
    // thread order is LIGHT-major: all samples towards light 0 of the batch's points first, then light 1, ... so the
    // shadow segments compacted from one warp point at the same emitter and walk the same BVH nodes (with the
    // lights interleaved per point the shadow kernel ran at 14 of 32 lanes)
    uint32_t t = blockIdx.x * IRR_BLOCK + threadIdx.x;
    bool need = false;
    DRay sh; sh.o = mk3(0, 0, 0); sh.d = mk3(0, 0, 1); sh.mint = 0.f; sh.maxt = 0.f;
    uint32_t target = 0;
    if (t < nPts * sl.S) {
        uint32_t j = 0, r = t, slotBase = 0;
        while (r >= nPts * sl.nS[j]) { r -= nPts * sl.nS[j]; slotBase += sl.nS[j]; ++j; }
        uint32_t point = r / sl.nS[j], s = r - point * sl.nS[j];
        target = point * sl.S + slotBase + s;
        uint64_t i = ptBase + point;
        const float *sp = reinterpret_cast<const float *>(pts + i);
        f3 p = mk3(__ldg(sp), __ldg(sp + 1), __ldg(sp + 2));
        f3 n = mk3(__ldg(sp + 3), __ldg(sp + 4), __ldg(sp + 5));
        float eps = __ldg(sp + 10);
        if (slotBase + s == 0) posArea[outBase + i] = make_float4(p.x, p.y, p.z, __ldg(sp + 9));
        // the third scramble word (uComponent) picks the shape of a ShapeSet; one sphere -> shape 0
        const uint32_t *w = scr + (i * sl.nLights + j) * 3;
        float u0 = vanDerCorput(s, __ldg(w)), u1 = sobol2(s, __ldg(w + 1));
        f3 wi; float lp;
        float cosT = 0.f, pdf = 0.f;
        const int kind = DELTA ? lightKind(lights[j]) : 0;
        if (kind == 0) {
            // The whole emitter below the point's horizon: every point the sampler can return -- on the sphere (cone or uniform
            // sampling) or, for a grazing ray that misses it numerically, within rounding of it -- gives Dot(wi, n) <= 0, so the
            // sample is the reference's `continue` (multipolesubsurface.cpp:121) whatever (u0, u1) are: {0, 0}, no shadow ray.
            // The margin (1 % of the radius + 1e-5 of the distance) is orders above the rounding of the sampled direction.
            const f3 toCentre = sub3(xformPoint(lights[j].o2w, mk3(0.f, 0.f, 0.f)), p);
            const float reach = fabsf(toCentre.x) + fabsf(toCentre.y) + fabsf(toCentre.z);
            if (dot3(toCentre, n) <= -1.01f * lights[j].radius - 1e-5f * reach) {
                // nothing to do
            } else if (sampleSphereLight(lights[j], p, eps, n, u0, u1, &wi, &lp, &sh)) {
                need = true;
                cosT = fminf(fabsf(dot3(wi, n)), 1.f); pdf = lp;
            }
        } else if (kind == 4) {      // infinite light (lights/infinite.cpp:195-221): the map's RGB in the sampled direction goes to envRgb
            f3 rgb;
            if (sampleInfiniteLight(lights[j], env, p, eps, u0, u1, &wi, &lp, &rgb, &sh) && dot3(wi, n) > 0.f && lp != 0.f) {
                need = true;
                cosT = fminf(fabsf(dot3(wi, n)), 1.f); pdf = lp;
                envRgb[target] = make_float4(rgb.x, rgb.y, rgb.z, 0.f);
            }
        } else {      // delta light: lightPdf = 1; the pdf slot carries DistanceSquared(lightPos, p) for k_irr_spectral's Li
            float d2, fall;
            bool lit = sampleDeltaLight(lights[j], kind, p, eps, &wi, &d2, &fall, &sh);
            if (dot3(wi, n) > 0.f && lit && d2 != 0.f) {
                need = true;
                cosT = fminf(fabsf(dot3(wi, n)), 1.f); pdf = d2;
            }
        }
        cosPdf[target] = make_float2(cosT, pdf);
    }
    // compact the shadow segments: warp ballot -> CTA total -> one atomic per CTA
    unsigned m = __ballot_sync(0xffffffffu, need);
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) sWarpCount[warp] = __popc(m);
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t tot = 0;
        for (int k = 0; k < IRR_BLOCK / 32; ++k) { uint32_t c = sWarpCount[k]; sWarpCount[k] = tot; tot += c; }
        sBase = tot ? atomicAdd(nRays, tot) : 0u;
        if (COUNT && tot) atomicAdd(&cnt[2], (unsigned long long)tot);
    }
    __syncthreads();
    if (need) {
        uint32_t r = sBase + sWarpCount[warp] + __popc(m & ((1u << lane) - 1u));
        rayO[r] = make_float4(sh.o.x, sh.o.y, sh.o.z, sh.mint);
        rayD[r] = make_float4(sh.d.x, sh.d.y, sh.d.z, sh.maxt);
        rayTarget[r] = target;
    }
}

template <bool COUNT>
__global__ void __launch_bounds__(128) k_irr_shadow(BvhView bv, const float4 *__restrict__ rayO, const float4 *__restrict__ rayD,
        const uint32_t *__restrict__ rayTarget, const uint32_t *__restrict__ nRays, float2 *__restrict__ cosPdf,
        unsigned long long *cnt) {
    uint32_t r = blockIdx.x * 128 + threadIdx.x;
    if (r >= __ldg(nRays)) return;
    float4 o = __ldg(&rayO[r]), d = __ldg(&rayD[r]);
    DRay ray; ray.o = mk3(o.x, o.y, o.z); ray.d = mk3(d.x, d.y, d.z); ray.mint = o.w; ray.maxt = d.w;
    if (COUNT ? bvhOccluded<true>(bv, ray, cnt) : bvhOccludedFast(bv, ray)) cosPdf[__ldg(&rayTarget[r])] = make_float2(0.f, 0.f);
}

template <bool DELTA>
__global__ void __launch_bounds__(256) k_irr_spectral(const skc_light *__restrict__ lights, IrrSlots sl, const float2 *__restrict__ cosPdf,
        uint64_t ptBase, uint32_t nPts, const float *__restrict__ rhoHd, const float *__restrict__ albedo, float mix,
        float *__restrict__ Eout, uint64_t outBase, const float4 *__restrict__ posArea, const float *__restrict__ albedoPowMix,
        const float4 *__restrict__ envRgb, const float *__restrict__ envIllum, AlbedoTexDev tex, const skc_surface_point *__restrict__ pts) {
    int lane = threadIdx.x & 31;
    uint32_t warpId = (blockIdx.x * 256 + threadIdx.x) >> 5;
    uint32_t p0 = warpId * IRR_SPECTRAL_POINTS;
    const int lam = lane < SKC_NS ? lane : 0;
    // the per-light sample counts in shared memory: indexing the kernel parameter with the loop variable compiles to a
    // 16-way select chain (40 instructions per point and light)
    __shared__ uint32_t sNS[IRR_MAX_LIGHTS];
    if (threadIdx.x < IRR_MAX_LIGHTS) sNS[threadIdx.x] = sl.nS[threadIdx.x];
    __syncthreads();
    // the first two lights' emission at this lane's wavelength stays in registers for the warp's points
    const float Lc0 = __ldg(&lights[0].Lemit[lam]), Lc1 = sl.nLights > 1 ? __ldg(&lights[1].Lemit[lam]) : 0.f;
    const float rN0 = 1.f / (float)(int)sl.nS[0], rN1 = sl.nLights > 1 ? 1.f / (float)(int)sl.nS[1] : 0.f;
    for (uint32_t k = 0; k < IRR_SPECTRAL_POINTS; ++k) {
        uint32_t point = p0 + k;
        if (point >= nPts) return;
        const float2 *cp = cosPdf + (size_t)point * sl.S;
        const float4 *ep = DELTA && envRgb ? envRgb + (size_t)point * sl.S : nullptr;      // per-sample RGB of the infinite light's samples
        float E = 0.f;
        for (uint32_t j = 0; j < sl.nLights; ++j) {
            uint32_t nS = sNS[j];
            float Lc = j < 2 ? (j == 0 ? Lc0 : Lc1) : __ldg(&lights[j].Lemit[lam]);
            float el = 0.f;
            const int kind = DELTA ? lightKind(lights[j]) : 0;
            float fall = 1.f;
            if (kind == 2) {      // SpotLight::Sample_L's Falloff(-wi), recomputed from the point (k_irr_sample keeps {cos, d2})
                float4 pa = __ldg(&posArea[outBase + ptBase + point]);
                f3 pp = mk3(pa.x, pa.y, pa.z);
                f3 lightPos = xformPoint(lights[j].o2w, mk3(0.f, 0.f, 0.f));
                fall = spotFalloff(lights[j], mul3(normalize3(sub3(lightPos, pp)), -1.f));
            }
            for (uint32_t sBase = 0; sBase < nS; sBase += 32) {
                uint32_t chunk = min(32u, nS - sBase);
                float2 mine = lane < chunk ? __ldg(&cp[sBase + lane]) : make_float2(0.f, 0.f);
                unsigned lit = __ballot_sync(0xffffffffu, mine.y != 0.f);
                while (lit) {                 // ascending sample order, as the reference's loop
                    int c = __ffs(lit) - 1; lit &= lit - 1;
                    float cosT = __shfl_sync(0xffffffffu, mine.x, c), pdf = __shfl_sync(0xffffffffu, mine.y, c);
                    // MultipoleBSSRDFData::rho (core/multipole.cpp:458-463)
                    float fid = cosT * (float)(SKC_RHO_LEN - 1);
                    int id = min(max((int)fid, 0), SKC_RHO_LEN - 2);
                    float lerp = fid - (float)id;
                    float r0 = __ldg(&rhoHd[id * SKC_EPAD + lane]), r1 = __ldg(&rhoHd[(id + 1) * SKC_EPAD + lane]);
                    float rho = (1.f - lerp) * r0 + lerp * r1;
                    float Ft = 1.f - rho;
                    float Li = Lc;
                    if (DELTA && kind == 4) { const float4 c4 = __ldg(&ep[sBase + c]); Li = envSpectrum(envIllum, c4.x, c4.y, c4.z, lam); }      // Spectrum(RGB, SPECTRUM_ILLUMINANT)
                    else if (DELTA && kind != 0) { Li = (Lc * fall) / pdf; pdf = 1.f; }      // Intensity * Falloff / DistanceSquared; lightPdf = 1
                    el += ((Ft * Li) * cosT) / pdf;      // multipolesubsurface.cpp:133
                }
            }
            cp += nS;
            if (ep) ep += nS;
            E += el * (j < 2 ? (j == 0 ? rN0 : rN1) : 1.f / (float)(int)nS);      // :136, `/ nSamples`: nS is a power of two (RoundUpPow2, :101), so the product with its reciprocal is the quotient
        }
        uint64_t pi = ptBase + point;
        float e = lane < SKC_NS ? E : 0.f;            // lanes 30,31 write the zero pad of the 128 B row
        if (albedo && lane < SKC_NS) e *= (float)pow((double)__ldg(&albedo[pi * SKC_NS + lane]), (double)mix);   // :139-140
        else if (DELTA && tex.rgb && lane < SKC_NS) {      // "texture albedo": ImageTexture::Evaluate at the point's (u, v) with zero differentials
            const float *sp = reinterpret_cast<const float *>(pts + pi);
            const f3 c = texBilinear(tex, 0, tex.su * __ldg(sp + 6) + tex.du, tex.sv * __ldg(sp + 7) + tex.dv);
            e *= skc_powf(texReflectance(tex.refl, c.x, c.y, c.z, lane), mix);                     // E *= Pow(albedo, mix)
        }
        else if (albedoPowMix && lane < SKC_NS) e *= __ldg(&albedoPowMix[lane]);      // a constant albedo: Pow(albedo, mix) made on the host with libm's powf
        Eout[(outBase + pi) * SKC_EPAD + lane] = e;
    }
}

// AoS export of the device point set (IrradiancePoint records for the host / the oracle)
__global__ void k_export_points(const skc_surface_point *__restrict__ pts, const float *__restrict__ E, uint64_t base,
                                uint64_t count, skc_irr_point *__restrict__ out) {
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t i = t >> 5; int lane = t & 31;
    if (i >= count) return;
    float *o = reinterpret_cast<float *>(out + i);
    const float *sp = reinterpret_cast<const float *>(pts + i);
    if (lane < 6) o[lane] = sp[lane];
    if (lane < SKC_NS) o[6 + lane] = E[(base + i) * SKC_EPAD + lane];
    if (lane == 30) o[36] = sp[9];
    if (lane == 31) o[37] = sp[10];
}

__global__ void k_import_points(const skc_irr_point *__restrict__ in, uint64_t count, float4 *__restrict__ posArea, float *__restrict__ E) {
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t i = t >> 5; int lane = t & 31;
    if (i >= count) return;
    const float *s = reinterpret_cast<const float *>(in + i);
    E[i * SKC_EPAD + lane] = lane < SKC_NS ? s[6 + lane] : 0.f;
    if (lane == 0) posArea[i] = make_float4(s[0], s[1], s[2], s[36]);
}

int skc_ensure_points(skc_ctx *ctx, uint64_t n) {
    if (ctx->ipCap >= n && ctx->ipPosArea) return 0;
    float4 *pa = nullptr; float *e = nullptr;
    SKC_CUDA(cudaMalloc((void **)&pa, n * sizeof(float4)));
    SKC_CUDA(cudaMalloc((void **)&e, n * SKC_EPAD * sizeof(float)));
    if (ctx->ipPosArea && ctx->ipCount) {
        uint64_t keep = ctx->ipCount < n ? ctx->ipCount : n;
        SKC_CUDA(cudaMemcpyAsync(pa, ctx->ipPosArea, keep * sizeof(float4), cudaMemcpyDeviceToDevice, ctx->stream));
        SKC_CUDA(cudaMemcpyAsync(e, ctx->ipE, keep * SKC_EPAD * sizeof(float), cudaMemcpyDeviceToDevice, ctx->stream));
        SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    if (ctx->ipPosArea) cudaFree(ctx->ipPosArea);
    if (ctx->ipE) cudaFree(ctx->ipE);
    ctx->ipPosArea = pa; ctx->ipE = e; ctx->ipCap = n;
    ctx->otNodes = 0; ctx->otTopologyReady = false;      // an octree built over the old arrays read its E rows from them
    return 0;
}

extern "C" int skc_upload_lights(skc_ctx *ctx, const skc_light *lights, uint32_t nLights) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (ctx->lights && ctx->nLights != nLights) { cudaFree(ctx->lights); ctx->lights = nullptr; }      // same count: the array is reused
    ctx->nLights = 0;
    if (!nLights) { if (ctx->lights) { cudaFree(ctx->lights); ctx->lights = nullptr; } return 0; }
    if (!ctx->lights) SKC_CUDA(cudaMalloc((void **)&ctx->lights, nLights * sizeof(skc_light)));
    SKC_CUDA(cudaMemcpyAsync(ctx->lights, lights, nLights * sizeof(skc_light), cudaMemcpyHostToDevice, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->nLights = nLights;
    for (uint32_t j = 0; j < nLights && j < IRR_MAX_LIGHTS; ++j) {
        int ns = (int)lights[j].nSamples;
        ctx->lightKinds[j] = lights[j].radius < 0.f ? (int)(-lights[j].radius) : 0;
        uint32_t nS = 1; while ((int)nS < ns) nS <<= 1;    // RoundUpPow2 (multipolesubsurface.cpp:101)
        ctx->lightNS[j] = nS;
    }
    ctx->hasDeltaLights = false;
    for (uint32_t j = 0; j < nLights; ++j) ctx->hasDeltaLights |= lights[j].radius < 0.f;
    return 0;
}

// device-level pass over surface points already on the device (dPts[0] = point `first`)
// hostPts (optional): the points of [first, first+count) still in host memory -- they are uploaded batch by batch on the
// copy stream, each batch's kernels waiting only for their own slice, so the 1 GB upload hides behind the compute
static int irradianceDev(skc_ctx *ctx, skc_surface_point *dPts, uint64_t nTotal, uint64_t first, uint64_t count,
                         uint32_t nTasks, float mix, const float *dAlb, unsigned long long *cntDev,
                         const skc_surface_point *hostPts = nullptr, bool deferTiming = false, uint64_t *batchPtsOut = nullptr) {
    if (!ctx->bvhNodes) return skc_fail(ctx, "skc_irradiance: no BVH uploaded");
    if (!ctx->nLights) return skc_fail(ctx, "skc_irradiance: no lights uploaded");
    if (!ctx->rhoHd) return skc_fail(ctx, "skc_irradiance: no rho table uploaded (skc_upload_profile)");
    if (first + count > nTotal || !nTasks) return skc_fail(ctx, "skc_irradiance: bad range");
    if (nTotal > 0x7fffffffull) return skc_fail(ctx, "skc_irradiance: the reference's int slice arithmetic caps the point count at 2^31-1");
    if (skc_ensure_points(ctx, nTotal)) return 1;
    if (ctx->ipCount < nTotal) ctx->ipCount = nTotal;
    if (!count) return 0;
    uint32_t wpp = 3 * ctx->nLights;
    uint32_t *dScr = (uint32_t *)skc_scratch(ctx, 1, count * wpp * sizeof(uint32_t));
    if (!dScr) return 1;
    // tasks overlapping [first, first+count): slice t covers [t*N/nT, (t+1)*N/nT)
    uint32_t t0 = 0, t1 = nTasks - 1;
    {
        uint64_t lo = 0, hi = nTasks - 1;
        while (lo < hi) { uint64_t m = (lo + hi) / 2; if ((m + 1) * nTotal / nTasks > first) hi = m; else lo = m + 1; }
        t0 = (uint32_t)lo;
        lo = 0; hi = nTasks - 1;
        while (lo < hi) { uint64_t m = (lo + hi + 1) / 2; if (m * nTotal / nTasks < first + count) lo = m; else hi = m - 1; }
        t1 = (uint32_t)lo;
    }
    // deferTiming: no host synchronisation inside the pass (the caller overlaps other work and collects the time later)
    if (deferTiming) SKC_CUDA(cudaEventRecord(ctx->defEv0, ctx->stream)); else skc_stage_begin(ctx);
    k_scrambles<<<t1 - t0 + 1, 256, 0, ctx->stream>>>(t0, nTasks, nTotal, first, count, wpp, dScr);
    SKC_LAUNCHED("k_scrambles");
    if (!deferTiming) skc_stage_end(ctx, 0);
    if (ctx->nLights > IRR_MAX_LIGHTS) return skc_fail(ctx, "skc_irradiance: %u lights, at most %d supported", ctx->nLights, IRR_MAX_LIGHTS);
    IrrSlots sl; sl.nLights = ctx->nLights; sl.S = 0;
    for (uint32_t j = 0; j < IRR_MAX_LIGHTS; ++j) { sl.nS[j] = j < ctx->nLights ? ctx->lightNS[j] : 0xffffffffu; if (j < ctx->nLights) sl.S += sl.nS[j]; }
    // batches of at most 2^25 (point, light, sample) slots: 256 MB of (cos, pdf) + at most 1.2 GB of shadow segments
    const uint64_t maxSlots = 1ull << 25;
    uint64_t batchPts = maxSlots / sl.S; if (!batchPts) batchPts = 1;
    if (batchPts > count) batchPts = count;
    uint64_t nBatches = (count + batchPts - 1) / batchPts, slots = batchPts * sl.S;
    if (batchPtsOut) *batchPtsOut = batchPts;
    float2 *dCosPdf = (float2 *)skc_scratch(ctx, 4, slots * sizeof(float2));
    char *dRays = (char *)skc_scratch(ctx, 5, slots * 36);
    uint32_t *dNRays = (uint32_t *)skc_scratch(ctx, 6, nBatches * sizeof(uint32_t));
    if (!dCosPdf || !dRays || !dNRays) return 1;
    // an infinite light's samples carry the map's RGB in their direction (16 B per slot; scenes without one do not pay for it)
    float4 *dEnvRgb = nullptr;
    for (uint32_t j = 0; j < ctx->nLights; ++j) if (ctx->lightKinds[j] == 4) {
        if (!ctx->env.rgb) return skc_fail(ctx, "skc_irradiance: the scene has an infinite light but no environment map was uploaded (skc_upload_envmap)");
        if (!(dEnvRgb = (float4 *)skc_scratch(ctx, 21, slots * sizeof(float4)))) return 1;
    }
    float4 *rayO = (float4 *)dRays, *rayD = rayO + slots;
    uint32_t *rayTarget = (uint32_t *)(rayD + slots);
    if (hostPts) {
        cudaEvent_t e0 = skc_event(ctx, 0);
        if (!e0 || !skc_event(ctx, nBatches)) return 1;
        SKC_CUDA(cudaEventRecord(e0, ctx->stream));                 // earlier work on the scratch buffer
        SKC_CUDA(cudaStreamWaitEvent(ctx->copyStream, e0, 0));
        for (uint64_t b = 0; b < nBatches; ++b) {
            uint64_t pb = b * batchPts, np = std::min<uint64_t>(batchPts, count - pb);
            SKC_CUDA(cudaMemcpyAsync(dPts + pb, hostPts + pb, np * sizeof(skc_surface_point), cudaMemcpyHostToDevice, ctx->copyStream));
            SKC_CUDA(cudaEventRecord(skc_event(ctx, 1 + b), ctx->copyStream));
        }
    }
    if (!deferTiming) skc_stage_begin(ctx);
    SKC_CUDA(cudaMemsetAsync(dNRays, 0, nBatches * sizeof(uint32_t), ctx->stream));

    for (uint64_t b = 0; b < nBatches; ++b) {
        uint64_t pb = b * batchPts;
        uint32_t np = (uint32_t)std::min<uint64_t>(batchPts, count - pb);
        if (hostPts) SKC_CUDA(cudaStreamWaitEvent(ctx->stream, skc_event(ctx, 1 + b), 0));
        uint64_t ns = (uint64_t)np * sl.S;
        unsigned gSample = (unsigned)((ns + IRR_BLOCK - 1) / IRR_BLOCK), gShadow = (unsigned)((ns + 127) / 128);
        unsigned gSpec = (unsigned)(((uint64_t)np + IRR_SPECTRAL_POINTS * 8 - 1) / (IRR_SPECTRAL_POINTS * 8));
        if (cntDev) {
            if (ctx->hasDeltaLights) k_irr_sample<true, true><<<gSample, IRR_BLOCK, 0, ctx->stream>>>(ctx->lights, sl, dPts, dScr, pb, np, dCosPdf, rayO, rayD, rayTarget, dNRays + b, ctx->ipPosArea, first, cntDev, ctx->env, dEnvRgb);
            else k_irr_sample<true, false><<<gSample, IRR_BLOCK, 0, ctx->stream>>>(ctx->lights, sl, dPts, dScr, pb, np, dCosPdf, rayO, rayD, rayTarget, dNRays + b, ctx->ipPosArea, first, cntDev, ctx->env, nullptr);
            k_irr_shadow<true><<<gShadow, 128, 0, ctx->stream>>>(skc_bvh_view(ctx), rayO, rayD, rayTarget, dNRays + b, dCosPdf, cntDev);
        } else {
            if (ctx->hasDeltaLights) k_irr_sample<false, true><<<gSample, IRR_BLOCK, 0, ctx->stream>>>(ctx->lights, sl, dPts, dScr, pb, np, dCosPdf, rayO, rayD, rayTarget, dNRays + b, ctx->ipPosArea, first, nullptr, ctx->env, dEnvRgb);
            else k_irr_sample<false, false><<<gSample, IRR_BLOCK, 0, ctx->stream>>>(ctx->lights, sl, dPts, dScr, pb, np, dCosPdf, rayO, rayD, rayTarget, dNRays + b, ctx->ipPosArea, first, nullptr, ctx->env, nullptr);
            k_irr_shadow<false><<<gShadow, 128, 0, ctx->stream>>>(skc_bvh_view(ctx), rayO, rayD, rayTarget, dNRays + b, dCosPdf, nullptr);
        }
        if (ctx->hasDeltaLights || ctx->albedoTex.rgb) k_irr_spectral<true><<<gSpec, 256, 0, ctx->stream>>>(ctx->lights, sl, dCosPdf, pb, np, ctx->rhoHd, dAlb, mix, ctx->ipE, first, ctx->ipPosArea, ctx->albedoFactors, dEnvRgb, ctx->env.illum, ctx->albedoTex, dPts);
        else k_irr_spectral<false><<<gSpec, 256, 0, ctx->stream>>>(ctx->lights, sl, dCosPdf, pb, np, ctx->rhoHd, dAlb, mix, ctx->ipE, first, ctx->ipPosArea, ctx->albedoFactors, nullptr, nullptr, AlbedoTexDev(), nullptr);
        ctx->launches += 2;
        SKC_LAUNCHED("k_irr_spectral");
    }
    if (deferTiming) SKC_CUDA(cudaEventRecord(ctx->defEv1, ctx->stream)); else skc_stage_end(ctx, 1);
    return 0;
}

// ---- Preprocess as one call: irradiance pass on `stream`, octree topology on `auxStream` -------------------------
// The octree's cell decomposition depends only on the point positions, which are known before any irradiance is:
// MultipoleSubsurfaceIntegrator::Preprocess (multipolesubsurface.cpp:170-238) = irradiance tasks, then Insert of every
// point, then InitHierarchy.  Here the positions of ALL points are written first, the topology pass runs on the
// auxiliary stream while the irradiance kernels (issue bound, little memory traffic) run on the main one, and
// skc_build_octree_finalize (leaf gather + InitHierarchy) follows when both are done.  In a multi-GPU run every rank
// builds the topology of the full point set this way while it computes only its own shard of irradiance.
__global__ void __launch_bounds__(256) k_positions(const skc_surface_point *__restrict__ pts, uint64_t n, float4 *__restrict__ posArea) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *sp = reinterpret_cast<const float *>(pts + i);
    posArea[i] = make_float4(__ldg(sp), __ldg(sp + 1), __ldg(sp + 2), __ldg(sp + 9));
}

static int topologyOnAux(skc_ctx *ctx, cudaEvent_t positionsReady) {
    SKC_CUDA(cudaStreamWaitEvent(ctx->auxStream, positionsReady, 0));
    std::swap(ctx->stream, ctx->auxStream);
    int rc = skc_build_octree_topology(ctx);
    std::swap(ctx->stream, ctx->auxStream);
    return rc;
}

static int collectDeferred(skc_ctx *ctx) {
    SKC_CUDA(cudaEventSynchronize(ctx->defEv1));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->defEv0, ctx->defEv1);
    ctx->stageMs[0] = 0.f; ctx->stageMs[1] = ms;     // scramble time is inside the pass here
    return 0;
}

extern "C" int skc_preprocess_resident(skc_ctx *ctx, uint64_t first, uint64_t count, uint32_t nTasks, float mix) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!ctx->resPts) return skc_fail(ctx, "skc_preprocess_resident: no surface points set");
    uint64_t n = ctx->resPtsCount;
    if (first + count > n) return skc_fail(ctx, "skc_preprocess_resident: bad range");
    if (skc_ensure_points(ctx, n)) return 1;
    ctx->ipCount = n;
    cudaEvent_t ready = skc_event(ctx, 0);
    if (!ready) return 1;
    k_positions<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(ctx->resPts, n, ctx->ipPosArea);
    SKC_LAUNCHED("k_positions");
    SKC_CUDA(cudaEventRecord(ready, ctx->stream));
    if (count && irradianceDev(ctx, ctx->resPts + first, n, first, count, nTasks, mix, nullptr, nullptr, nullptr, true)) return 1;
    int rc = topologyOnAux(ctx, ready);
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    if (rc) return rc;
    SKC_CUDA(e);
    if (count) return collectDeferred(ctx);
    return 0;
}

extern "C" int skc_preprocess(skc_ctx *ctx, const skc_surface_point *points, uint64_t nTotal, uint64_t first, uint64_t count,
                              uint32_t nTasks, float mix) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (first + count > nTotal || !nTotal) return skc_fail(ctx, "skc_preprocess: bad range");
    if (skc_ensure_points(ctx, nTotal)) return 1;
    ctx->ipCount = nTotal;
    // all points go up (the topology needs every position); the irradiance pass runs on [first, first+count) and its
    // batches wait only for their own slice of the upload
    skc_surface_point *dAll = (skc_surface_point *)skc_scratch(ctx, 0, nTotal * sizeof(skc_surface_point));
    if (!dAll) return 1;
    uint64_t batchPts = count ? count : 1, nB = 0;
    if (count) {
        if (irradianceDev(ctx, dAll + first, nTotal, first, count, nTasks, mix, nullptr, nullptr, points + first, true, &batchPts)) return 1;
        nB = (count + batchPts - 1) / batchPts;
        for (uint64_t b = 0; b < nB; ++b) {       // positions of the own range, batch by batch behind the upload
            uint64_t a = first + b * batchPts, np = std::min(batchPts, first + count - a);
            SKC_CUDA(cudaStreamWaitEvent(ctx->auxStream, skc_event(ctx, 1 + b), 0));
            k_positions<<<(unsigned)((np + 255) / 256), 256, 0, ctx->auxStream>>>(dAll + a, np, ctx->ipPosArea + a);
            SKC_LAUNCHED("k_positions");
        }
    } else {
        cudaEvent_t e0 = skc_event(ctx, 0);
        if (!e0) return 1;
        SKC_CUDA(cudaEventRecord(e0, ctx->stream));
        SKC_CUDA(cudaStreamWaitEvent(ctx->copyStream, e0, 0));
    }
    // the other ranks' ranges (multi-GPU): positions only are needed, but the records are 44 B AoS, so they go up whole
    const uint64_t chunk = 1ull << 22;
    size_t ev = 1 + nB;
    uint64_t ranges[2][2] = {{0, first}, {first + count, nTotal}};
    for (int r = 0; r < 2; ++r)
        for (uint64_t a = ranges[r][0]; a < ranges[r][1]; a += chunk) {
            uint64_t np = std::min(chunk, ranges[r][1] - a);
            cudaEvent_t e = skc_event(ctx, ev++);
            if (!e) return 1;
            SKC_CUDA(cudaMemcpyAsync(dAll + a, points + a, np * sizeof(skc_surface_point), cudaMemcpyHostToDevice, ctx->copyStream));
            SKC_CUDA(cudaEventRecord(e, ctx->copyStream));
            SKC_CUDA(cudaStreamWaitEvent(ctx->auxStream, e, 0));
            k_positions<<<(unsigned)((np + 255) / 256), 256, 0, ctx->auxStream>>>(dAll + a, np, ctx->ipPosArea + a);
            SKC_LAUNCHED("k_positions");
        }
    cudaEvent_t ready = skc_event(ctx, ev);
    if (!ready) return 1;
    SKC_CUDA(cudaEventRecord(ready, ctx->auxStream));
    int rc = topologyOnAux(ctx, ready);
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    if (rc) return rc;
    SKC_CUDA(e);
    if (count) return collectDeferred(ctx);
    return 0;
}

extern "C" int skc_irradiance(skc_ctx *ctx, const skc_surface_point *points, uint64_t nTotal,
                              uint64_t first, uint64_t count, uint32_t nTasks, float mix,
                              const float *albedo30, skc_irr_point *out) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (first + count > nTotal) return skc_fail(ctx, "skc_irradiance: bad range");
    if (!count) { if (skc_ensure_points(ctx, nTotal ? nTotal : 1)) return 1; if (ctx->ipCount < nTotal) ctx->ipCount = nTotal; return 0; }
    skc_surface_point *dPts = (skc_surface_point *)skc_scratch(ctx, 0, count * sizeof(skc_surface_point));
    if (!dPts) return 1;
    float *dAlb = nullptr;
    if (albedo30) {
        dAlb = (float *)skc_scratch(ctx, 2, count * SKC_NS * sizeof(float));
        if (!dAlb) return 1;
        SKC_CUDA(cudaMemcpyAsync(dAlb, albedo30, count * SKC_NS * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    }
    if (irradianceDev(ctx, dPts, nTotal, first, count, nTasks, mix, dAlb, nullptr, points + first)) return 1;
    if (out) {
        skc_irr_point *dOut = (skc_irr_point *)skc_scratch(ctx, 3, count * sizeof(skc_irr_point));
        if (!dOut) return 1;
        uint64_t threads = count * 32;
        k_export_points<<<(unsigned)((threads + 255) / 256), 256, 0, ctx->stream>>>(dPts, ctx->ipE, first, count, dOut);
        SKC_LAUNCHED("k_export_points");
        SKC_CUDA(cudaMemcpyAsync(out, dOut, count * sizeof(skc_irr_point), cudaMemcpyDeviceToHost, ctx->stream));
    }
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// ---- resident inputs (the reference's "pointsfile" cache, kept in HBM) ----------------------
extern "C" int skc_set_surface_points(skc_ctx *ctx, const skc_surface_point *points, uint64_t n) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (ctx->resPts) { cudaFree(ctx->resPts); ctx->resPts = nullptr; }
    ctx->resPtsCount = 0; ctx->resPtsCap = 0;
    if (!n) return 0;
    SKC_CUDA(cudaMalloc((void **)&ctx->resPts, n * sizeof(skc_surface_point)));
    ctx->resPtsCap = n;
    SKC_CUDA(cudaMemcpyAsync(ctx->resPts, points, n * sizeof(skc_surface_point), cudaMemcpyHostToDevice, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->resPtsCount = n;
    return 0;
}

extern "C" int skc_irradiance_resident(skc_ctx *ctx, uint64_t first, uint64_t count, uint32_t nTasks, float mix, uint64_t *counters3) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!ctx->resPts) return skc_fail(ctx, "skc_irradiance_resident: no surface points set");
    if (first + count > ctx->resPtsCount) return skc_fail(ctx, "skc_irradiance_resident: bad range");
    unsigned long long *dCnt = nullptr;
    if (counters3) {
        dCnt = (unsigned long long *)skc_scratch(ctx, 7, 4 * sizeof(unsigned long long));
        if (!dCnt) return 1;
        SKC_CUDA(cudaMemsetAsync(dCnt, 0, 4 * sizeof(unsigned long long), ctx->stream));
    }
    if (irradianceDev(ctx, ctx->resPts + first, ctx->resPtsCount, first, count, nTasks, mix, nullptr, dCnt)) return 1;
    if (counters3) {
        unsigned long long h[4];
        SKC_CUDA(cudaMemcpyAsync(h, dCnt, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
        SKC_CUDA(cudaStreamSynchronize(ctx->stream));
        counters3[0] = h[0]; counters3[1] = h[1]; counters3[2] = h[2];
    }
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

extern "C" void *skc_alloc_pinned(uint64_t bytes) {
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) return nullptr;
    return p;
}
extern "C" void skc_free_pinned(void *p) { if (p) cudaFreeHost(p); }

extern "C" int skc_set_irradiance_points(skc_ctx *ctx, const skc_irr_point *pts, uint64_t n) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    ctx->ipCount = 0;
    if (!n) return 0;
    if (skc_ensure_points(ctx, n)) return 1;
    skc_irr_point *d = (skc_irr_point *)skc_scratch(ctx, 3, n * sizeof(skc_irr_point));
    if (!d) return 1;
    SKC_CUDA(cudaMemcpyAsync(d, pts, n * sizeof(skc_irr_point), cudaMemcpyHostToDevice, ctx->stream));
    uint64_t threads = n * 32;
    k_import_points<<<(unsigned)((threads + 255) / 256), 256, 0, ctx->stream>>>(d, n, ctx->ipPosArea, ctx->ipE);
    SKC_LAUNCHED("k_import_points");
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->ipCount = n;
    return 0;
}

extern "C" int skc_irradiance_dev(skc_ctx *ctx, uint64_t nTotal, void **posArea_dev, void **E_dev) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (skc_ensure_points(ctx, nTotal)) return 1;
    if (ctx->ipCount < nTotal) ctx->ipCount = nTotal;
    if (posArea_dev) *posArea_dev = ctx->ipPosArea;
    if (E_dev) *E_dev = ctx->ipE;
    return 0;
}

// skc_rho.cu -- hemispherical-directional reflectance table of the skin's specular lobe on the device.
//
// Reference: ComputeRhoDataFromBxDF / RhoTask::Run (core/multipole.cpp:488-549), BxDF::rho with
// KahanSum (core/reflection.cpp:623-635, core/kahansum.h:35-62), StratifiedSample2D
// (core/montecarlo.cpp:158-168), Microfacet::Sample_f / f / G (core/reflection.cpp:228-239,391-396,
// core/reflection.h:430-437), Beckmann::D / Sample_f (core/reflection.h:516-523, reflection.cpp:548-569),
// FresnelDielectric::Evaluate / FrDiel (reflection.cpp:62-69,132-152), FixedFresnelDielectric
// (reflection.h:315-324), RNG (core/rng.cpp:43-110).
//
// Entry k of the table (k = 0..1024) is a 65 536-sample stratified Monte Carlo estimate with the jitter
// stream RNG(6428263 * k).  One CTA per entry: the MT19937 stream is regenerated 624 words at a time in
// shared memory (312 samples per block), every thread evaluates its samples in the reference's fp32
// arithmetic (libm's float functions restated bit for bit) and one thread adds them in the reference's order with
// its compensated fp32 sum, so the table equals the reference's bit for bit.  R = 1 and the Fresnel term is wavelength independent, so all 30
// spectral channels of an entry are equal.
#include "skc_internal.h"
#include "skc_bsdf.cuh"

#define MT_N 624
#define MT_M 397
#define RHO_SQRT_SAMPLES 256
#define RHO_ENTRIES 1025

__device__ __forceinline__ uint32_t rhoTwist(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t y = (a & 0x80000000u) | (b & 0x7fffffffu);
    return c ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
}
__device__ __forceinline__ float rhoRandomFloat(uint32_t y) {   // tempering + RNG::RandomFloat (rng.cpp:59-65,105-110)
    y ^= (y >> 11);
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= (y >> 18);
    return (float)(y & 0xffffffu) / 16777216.f;
}

// one term f * |cos(wi)| / pdf of BxDF::rho(wo, ...) for the Beckmann microfacet lobe.  Returns false where the reference adds
// nothing (`if (pdf > 0.)`, reflection.cpp:631); a term of 0 with pdf > 0 IS added (it folds the compensation into the sum).
__device__ bool rhoTerm(const RhoBxdf &b, float wox, float woz, float u1, float u2, float *term) {
    // Beckmann::Sample_f (reflection.cpp:548-569): libm's atanf / logf / cosf / sinf, restated bit for bit (skc_common.cuh)
    float theta = skc_atanf(sqrtf(-b.rms2 * skc_logf(1.f - u1)));
    float costheta = skc_cosf(theta), sintheta = skc_sinf(theta);
    float phi = u2 * 2.f * SKC_PI_F;
    float whx = sintheta * skc_cosf(phi), why = sintheta * skc_sinf(phi), whz = costheta;
    if (!(woz * whz > 0.f)) { whx = -whx; why = -why; whz = -whz; }
    float dotWoWh = wox * whx + 0.f * why + woz * whz;
    float k = 2.f * dotWoWh;
    float wix = -wox + k * whx, wiy = -0.f + k * why, wiz = -woz + k * whz;
    float pdf = beckmannD(b, whz) * costheta / (4.f * dotWoWh);
    if (dotWoWh <= 0.f || pdf < 1e-20f) pdf = 0.f;
    *term = 0.f;
    if (!(pdf > 0.)) return false;
    // Microfacet::Sample_f -> f (reflection.cpp:228-239,391-396)
    if (!(woz * wiz > 0.f)) return true;
    float cosThetaO = fabsf(woz), cosThetaI = fabsf(wiz);
    if (cosThetaI == 0.f || cosThetaO == 0.f) return true;
    float hx = wix + wox, hy = wiy + 0.f, hz = wiz + woz;
    if (hx == 0. && hy == 0. && hz == 0.) return true;
    float inv = 1.f / sqrtf(hx * hx + hy * hy + hz * hz);      // Normalize: v / Length = v * (1/len)
    hx *= inv; hy *= inv; hz *= inv;
    float cosThetaH = wix * hx + wiy * hy + wiz * hz;
    float F = fresnelDielectric(b, cosThetaH);
    float NdotWh = fabsf(hz), WOdotWh = fabsf(wox * hx + 0.f * hy + woz * hz);
    float G = fminf(1.f, fminf(2.f * NdotWh * cosThetaO / WOdotWh, 2.f * NdotWh * cosThetaI / WOdotWh));
    float f = 1.f * beckmannD(b, hz) * G * F / (4.f * cosThetaI * cosThetaO);
    *term = f * cosThetaI / pdf;
    return true;
}

// One CTA per table entry.  The reference sums the 65 536 terms of an entry with KahanSum<Spectrum> in sample order
// (reflection.cpp:625-633, kahansum.h:35-47): a compensated fp32 sum whose result depends on that order in its last bit.  So the
// terms of each 312-sample block of the MT19937 stream are computed by all threads into shared memory and then added by ONE
// thread, in order, with the reference's four operations per term.
__global__ void __launch_bounds__(256) k_rho_hd(RhoBxdf bxdf, float *__restrict__ hd30) {
    __shared__ uint32_t mt[MT_N];
    __shared__ float sTerm[MT_N / 2];
    __shared__ uint8_t sAdd[MT_N / 2];
    int tid = threadIdx.x;
    uint32_t id = blockIdx.x;
    if (tid == 0) {   // RNG::Seed(6428263 * id)
        uint32_t s = 6428263u * id;
        mt[0] = s;
        for (int i = 1; i < MT_N; ++i) { s = 1812433253u * (s ^ (s >> 30)) + (uint32_t)i; mt[i] = s; }
    }
    // multipole.cpp:531-536
    float costheta = (float)id / 1024;
    if (costheta == 0.f) costheta = 0.01f / 1024;
    float sinT = sqrtf(1 - costheta * costheta);
    float wox = sinT * 1.f, woz = costheta;               // SphericalDirection(sin, cos, 0): cos(0) = 1, sin(0) = 0
    const uint32_t nSamples = RHO_SQRT_SAMPLES * RHO_SQRT_SAMPLES;
    const float dx = 1.f / RHO_SQRT_SAMPLES;
    float sum = 0.f, comp = 0.f;                          // KahanSum state (thread 0)
    __syncthreads();
    for (uint32_t base = 0; base < 2 * nSamples; base += MT_N) {
        uint32_t v = 0;
        if (tid < 227) v = rhoTwist(mt[tid], mt[tid + 1], mt[tid + MT_M]);
        __syncthreads();
        if (tid < 227) mt[tid] = v;
        __syncthreads();
        if (tid < 227) v = rhoTwist(mt[tid + 227], mt[tid + 228], mt[tid]);
        __syncthreads();
        if (tid < 227) mt[tid + 227] = v;
        __syncthreads();
        if (tid < 169) v = rhoTwist(mt[tid + 454], mt[tid + 455], mt[tid + 227]);
        __syncthreads();
        if (tid < 169) mt[tid + 454] = v;
        __syncthreads();
        if (tid == 0) mt[623] = rhoTwist(mt[623], mt[0], mt[396]);
        __syncthreads();
        for (uint32_t k = tid; k < MT_N / 2; k += 256) {
            uint32_t s = base / 2 + k;
            bool add = false; float term = 0.f;
            if (s < nSamples) {
                uint32_t x = s % RHO_SQRT_SAMPLES, y = s / RHO_SQRT_SAMPLES;
                float jx = rhoRandomFloat(mt[2 * k]), jy = rhoRandomFloat(mt[2 * k + 1]);
                float u1 = fminf(((float)x + jx) * dx, 0.99999994f), u2 = fminf(((float)y + jy) * dx, 0.99999994f);
                add = rhoTerm(bxdf, wox, woz, u1, u2, &term);
            }
            sTerm[k] = term; sAdd[k] = add ? 1 : 0;
        }
        __syncthreads();
        if (tid == 0) {
            for (int k = 0; k < MT_N / 2; ++k)
                if (sAdd[k]) {                            // KahanSum::Add
                    float yv = __fsub_rn(sTerm[k], comp);
                    float t = __fadd_rn(sum, yv);
                    comp = __fsub_rn(__fsub_rn(t, sum), yv);
                    sum = t;
                }
        }
        __syncthreads();
    }
    __shared__ float sResult;
    if (tid == 0) sResult = sum / (float)nSamples;        // r.Value() / float(nSamples)
    __syncthreads();
    if (tid < 30) hd30[id * 30 + tid] = sResult;
}

extern "C" int skc_rho_table(skc_ctx *ctx, float roughness, float etaI, float etaT, int fixedFresnel, float *hd30) {
    if (!ctx) return 1;
    cudaSetDevice(ctx->device);
    if (!hd30) return skc_fail(ctx, "skc_rho_table: null output");
    RhoBxdf b;
    float rms = roughness < 1e-3f ? 1e-3f : roughness;     // Beckmann ctor, reflection.h:509-514
    b.rms2 = rms * rms;
    b.rcpRMS2 = 1 / b.rms2;
    b.etaI = etaI; b.etaT = etaT; b.fixedFresnel = fixedFresnel;
    float *d = (float *)skc_scratch(ctx, 2, RHO_ENTRIES * 30 * sizeof(float));
    if (!d) return 1;
    skc_stage_begin(ctx);
    k_rho_hd<<<RHO_ENTRIES, 256, 0, ctx->stream>>>(b, d);
    SKC_LAUNCHED("k_rho_hd");
    skc_stage_end(ctx, 7);
    SKC_CUDA(cudaMemcpyAsync(hd30, d, RHO_ENTRIES * 30 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    SKC_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

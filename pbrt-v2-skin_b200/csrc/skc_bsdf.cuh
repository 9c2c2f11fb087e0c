// skc_bsdf.cuh -- the LayeredSkin surface BSDF on the device (materials/layeredskin.cpp:139-167):
// Microfacet (reflection) + MicrofacetTransmission over a Beckmann distribution with a dielectric Fresnel term,
// and the Lambertian lobe of the default matte the emitter spheres carry.  All vectors are in the BSDF's local
// frame (z = shading normal).  Reference: core/reflection.h:430-437,507-529, core/reflection.cpp:62-69,132-152,
// 228-296,391-470,548-581; core/montecarlo.cpp:306-348,422-427.
#pragma once
#include "skc_common.cuh"

struct RhoBxdf {   // Beckmann(rms) + FresnelDielectric(etaI, etaT) [+ FixedFresnelDielectric]
    float rms2, rcpRMS2, etaI, etaT; int fixedFresnel; };

__device__ __forceinline__ float beckmannD(const RhoBxdf &b, float whz) {
    float costheta = fabsf(whz);
    float c2 = costheta * costheta;
    float d = c2 * c2 * SKC_PI_F;
    if (d == 0.f) return 0.f;
    float e = (c2 - 1) * b.rcpRMS2 / c2;
    return b.rcpRMS2 * skc_expf(e) / d;        // libm's expf, bit for bit (skc_common.cuh)
}

__device__ __forceinline__ float fresnelDielectric(const RhoBxdf &b, float cosi) {
    cosi = fminf(fmaxf(cosi, -1.f), 1.f);
    bool entering = cosi > 0.;
    float ei = entering ? b.etaI : b.etaT, et = entering ? b.etaT : b.etaI;
    float sint = ei / et * sqrtf(fmaxf(0.f, 1.f - cosi * cosi));
    float F;
    if (sint >= 1.) F = 1.f;
    else {
        float cost = sqrtf(fmaxf(0.f, 1.f - sint * sint));
        float ci = fabsf(cosi);
        float Rparl = ((et * ci) - (ei * cost)) / ((et * ci) + (ei * cost));
        float Rperp = ((ei * ci) - (et * cost)) / ((ei * ci) + (et * cost));
        F = (Rparl * Rparl + Rperp * Rperp) / 2.f;
    }
    if (b.fixedFresnel) F = F + F * (1.f - F) * (1.f - F);
    return F;
}


// ---- local-frame helpers (core/reflection.h:49-75) ---------------------------------------------------------------
__device__ __forceinline__ bool sameHemisphere(f3 w, f3 wp) { return w.z * wp.z > 0.f; }
// libm's float functions as glibc computes them (restated in skc_common.cuh; they are not correctly rounded, and the
// reference's results depend on their last bit)
__device__ __forceinline__ float cr_cosf(float x) { return skc_cosf(x); }     // libm's cosf / sinf bit for bit (skc_common.cuh)
__device__ __forceinline__ float cr_sinf(float x) { return skc_sinf(x); }
__device__ __forceinline__ float cr_logf(float x) { return skc_logf(x); }
__device__ __forceinline__ float cr_atanf(float x) { return skc_atanf(x); }

// Beckmann::Sample_f (core/reflection.cpp:548-569)
__device__ inline void beckmannSample(const RhoBxdf &b, f3 wo, float u1, float u2, f3 *wi, float *pdf) {
    float theta = cr_atanf(sqrtf(-b.rms2 * cr_logf(1.f - u1)));
    float costheta = cr_cosf(theta), sintheta = cr_sinf(theta);
    float phi = u2 * 2.f * SKC_PI_F;
    f3 wh = mk3(sintheta * cr_cosf(phi), sintheta * cr_sinf(phi), costheta);
    if (!sameHemisphere(wo, wh)) wh = mul3(wh, -1.f);
    float d = dot3(wo, wh);
    *wi = add3(mul3(wo, -1.f), mul3(wh, 2.f * d));
    float p = beckmannD(b, wh.z) * costheta / (4.f * d);
    if (d <= 0.f || p < 1e-20f) p = 0.f;
    *pdf = p;
}
// Beckmann::Pdf (core/reflection.cpp:572-579)
__device__ inline float beckmannPdf(const RhoBxdf &b, f3 wo, f3 wi) {
    f3 wh = normalize3(add3(wo, wi));
    float costheta = fabsf(wh.z);
    float d = dot3(wo, wh);
    float p = beckmannD(b, wh.z) * costheta / (4.f * d);
    if (d <= 0.f || p < 1e-20f) p = 0.f;
    return p;
}
// Microfacet::G (core/reflection.h:430-437)
__device__ __forceinline__ float microfacetG(f3 wo, f3 wi, f3 wh) {
    float NdotWh = fabsf(wh.z), NdotWo = fabsf(wo.z), NdotWi = fabsf(wi.z);
    float WOdotWh = fabsf(dot3(wo, wh));
    return fminf(1.f, fminf(2.f * NdotWh * NdotWo / WOdotWh, 2.f * NdotWh * NdotWi / WOdotWh));
}
// Microfacet::f (core/reflection.cpp:228-239).  The lobe's value in spectral channel c is ((R_c * D) * G) * F / den;
// comp = {D, G, F, den} carries the factors, the return value is the same expression with R = 1 (zero <=> black).
__device__ inline float microfacetF(const RhoBxdf &b, f3 wo, f3 wi, float4 *comp = nullptr) {
    float cosThetaO = fabsf(wo.z), cosThetaI = fabsf(wi.z);
    if (cosThetaI == 0.f || cosThetaO == 0.f) return 0.f;
    f3 wh = add3(wi, wo);
    if (wh.x == 0. && wh.y == 0. && wh.z == 0.) return 0.f;
    wh = normalize3(wh);
    float cosThetaH = dot3(wi, wh);
    float F = fresnelDielectric(b, cosThetaH);
    float D = beckmannD(b, wh.z), G = microfacetG(wo, wi, wh), den = 4.f * cosThetaI * cosThetaO;
    if (comp) *comp = make_float4(D, G, F, den);
    return 1.f * D * G * F / den;
}
__device__ __forceinline__ float microfacetPdf(const RhoBxdf &b, f3 wo, f3 wi) {   // reflection.cpp:399-402
    if (!sameHemisphere(wo, wi)) return 0.f;
    return beckmannPdf(b, wo, wi);
}
// MicrofacetTransmission::G (core/reflection.cpp:273-281)
__device__ __forceinline__ float transmissionG(f3 wo, f3 wi, f3 wh) {
    float NdotWh = fabsf(wh.z), NdotWo = fabsf(wo.z), NdotWi = fabsf(wi.z);
    float WOdotWh = fabsf(dot3(wo, wh)), WIdotWh = fabsf(dot3(wi, wh));
    return fminf(1.f, fminf(2.f * NdotWh * NdotWo / WOdotWh, 2.f * NdotWh * NdotWi / WIdotWh));
}
// MicrofacetTransmission::f (core/reflection.cpp:251-270).  Channel value (a * T_c) * (1 - F); comp = {a, 1 - F, -1, 1}.
__device__ inline float transmissionF(const RhoBxdf &b, float ior, f3 wo, f3 wi, float4 *comp = nullptr) {
    float cosThetaI = fabsf(wi.z), cosThetaO = fabsf(wo.z);
    if (cosThetaI == 0.f || cosThetaO == 0.f) return 0.f;
    bool entering = wo.z > 0.f;
    float et = entering ? ior : 1.f / ior;
    f3 wh = mul3(add3(wo, mul3(wi, et)), -1.f);
    float denominator = wh.x * wh.x + wh.y * wh.y + wh.z * wh.z;
    if (denominator == 0.f) return 0.f;
    wh = normalize3(wh);
    float cosThetaHI = dot3(wi, wh), cosThetaHO = dot3(wo, wh);
    if (cosThetaHI == 0.f || cosThetaHO == 0.f) return 0.f;
    float F = fresnelDielectric(b, entering ? fabsf(cosThetaHO) : -fabsf(cosThetaHO));
    float a = fabsf(cosThetaHI * cosThetaHO) * et * et * beckmannD(b, wh.z) * transmissionG(wo, wi, wh) / (cosThetaI * cosThetaO * denominator);
    if (comp) *comp = make_float4(a, 1.f - F, -1.f, 1.f);
    return (a * 1.f) * (1.f - F);
}
// MicrofacetTransmission::Pdf (core/reflection.cpp:444-463)
__device__ inline float transmissionPdf(const RhoBxdf &b, float ior, f3 wo, f3 wi) {
    if (sameHemisphere(wo, wi)) return 0.f;
    bool entering = wo.z > 0.f;
    float et = entering ? ior : 1.f / ior;
    f3 wh = mul3(add3(wo, mul3(wi, et)), -1.f);
    if (wh.x == 0.f && wh.y == 0.f && wh.z == 0.f) return 0.f;
    float denominator = wh.x * wh.x + wh.y * wh.y + wh.z * wh.z;
    wh = normalize3(wh);
    f3 wiR = add3(mul3(wo, -1.f), mul3(wh, 2.f * dot3(wo, wh)));
    float pdf = beckmannPdf(b, wo, wiR);
    float cosi = dot3(wo, wh);
    pdf *= 4 * cosi * cosi * et * et / denominator;
    return pdf;
}
// MicrofacetTransmission::Sample_f (core/reflection.cpp:405-441): direction and pdf only (BSDF::Sample_f re-evaluates f)
__device__ inline void transmissionSample(const RhoBxdf &b, float ior, f3 wo, float u1, float u2, f3 *wi, float *pdf) {
    beckmannSample(b, wo, u1, u2, wi, pdf);
    bool entering = wo.z > 0.f;
    float et = entering ? ior : 1.f / ior;
    f3 wh = add3(wo, *wi);
    if (wh.x == 0.f && wh.y == 0.f && wh.z == 0.f) return;      // returns black with the reflected wi and pdf left in place
    wh = normalize3(wh);
    float cosi = dot3(wo, wh);
    float sini2 = fmaxf(0.f, 1.f - cosi * cosi);
    float eta = 1.f / et;
    float sint2 = eta * eta * sini2;
    if (sint2 >= 1.) { *pdf = 0.f; return; }
    float cost = sqrtf(fmaxf(0.f, 1.f - sint2));
    float sintOverSini = eta;
    *wi = add3(mul3(wo, -sintOverSini), mul3(wh, sintOverSini * cosi - cost));
    f3 dv = add3(wo, mul3(*wi, et));
    float denominator = dv.x * dv.x + dv.y * dv.y + dv.z * dv.z;
    if (denominator == 0.f) { *pdf = 0.f; return; }
    *pdf *= 4 * cosi * cosi * et * et / denominator;
}
// ConcentricSampleDisk + CosineSampleHemisphere (core/montecarlo.cpp:306-348, montecarlo.h:232-237)
__device__ inline f3 cosineSampleHemisphere(float u1, float u2) {
    float r, theta, dx, dy;
    float sx = 2 * u1 - 1, sy = 2 * u2 - 1;
    if (sx == 0.0 && sy == 0.0) { dx = 0.f; dy = 0.f; }
    else {
        if (sx >= -sy) {
            if (sx > sy) { r = sx; theta = sy > 0.0 ? sy / r : 8.0f + sy / r; }
            else { r = sy; theta = 2.0f - sx / r; }
        } else {
            if (sx <= sy) { r = -sx; theta = 4.0f - sy / r; }
            else { r = -sy; theta = 6.0f + sx / r; }
        }
        theta *= SKC_PI_F / 4.f;
        dx = r * cr_cosf(theta); dy = r * cr_sinf(theta);
    }
    return mk3(dx, dy, sqrtf(fmaxf(0.f, 1.f - dx * dx - dy * dy)));
}

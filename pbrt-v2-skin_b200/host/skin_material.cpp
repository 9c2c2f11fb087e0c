// skin_material.cpp -- pigment model and profile-table construction of the LayeredSkin material.
// Arithmetic follows the reference's fp32 expressions term by term so that the layer coefficients
// (and therefore every table entry downstream) are identical; see the citations per function.
#include <algorithm>
#include <cmath>
#include "skin_material.h"
#include <cstring>
#include "skin_tables.h"

namespace skin {

namespace {

const int kPig = kPigmentSamples;                 // 61 samples, 400..700 nm / 5 nm (core/material.cpp:117-125)
struct Wld { float c[kPig]; };                    // WLDValue

inline float wavelength(int i) { return (float)(400 + 5 * i); }
inline float lerpf(float t, float a, float b) { return (1.f - t) * a + t * b; }   // core/pbrt.h Lerp

template <class F> Wld tabulate(F f) { Wld r; for (int i = 0; i < kPig; ++i) r.c[i] = f(wavelength(i)); return r; }
template <class F> Wld zip(const Wld &a, const Wld &b, F f) { Wld r; for (int i = 0; i < kPig; ++i) r.c[i] = f(a.c[i], b.c[i]); return r; }
inline Wld scaled(const Wld &a, float s) { Wld r; for (int i = 0; i < kPig; ++i) r.c[i] = a.c[i] * s; return r; }
inline Wld sum(const Wld &a, const Wld &b) { return zip(a, b, [](float x, float y) { return x + y; }); }
inline Wld fromTable(const float *v) { Wld r; std::copy(v, v + kPig, r.c); return r; }   // WLDValue::FromSampled on the same 61 abscissae

// AverageSpectrumSamples (core/spectrum.cpp:60-93) over the 61 pigment samples
float averageSamples(const float *vals, float lambdaStart, float lambdaEnd) {
    const int n = kPig;
    if (lambdaEnd <= wavelength(0)) return vals[0];
    if (lambdaStart >= wavelength(n - 1)) return vals[n - 1];
    float total = 0.f;
    if (lambdaStart < wavelength(0)) total += vals[0] * (wavelength(0) - lambdaStart);
    if (lambdaEnd > wavelength(n - 1)) total += vals[n - 1] * (lambdaEnd - wavelength(n - 1));
    int i = 0;
    while (lambdaStart > wavelength(i + 1)) ++i;
    auto at = [&](float w, int k) { return lerpf((w - wavelength(k)) / (wavelength(k + 1) - wavelength(k)), vals[k], vals[k + 1]); };
    for (; i + 1 < n && lambdaEnd >= wavelength(i); ++i) {
        float segStart = std::max(lambdaStart, wavelength(i)), segEnd = std::min(lambdaEnd, wavelength(i + 1));
        total += (0.5f * (at(segStart, i) + at(segEnd, i))) * (segEnd - segStart);
    }
    return total / (lambdaEnd - lambdaStart);
}

// SampledSpectrum::FromSampled (core/spectrum.h:302-321)
void toBins(const Wld &v, float out[kWavelengths]) {
    for (int i = 0; i < kWavelengths; ++i) {
        float l0 = lerpf((float)i / (float)kWavelengths, 400.f, 700.f);
        float l1 = lerpf((float)(i + 1) / (float)kWavelengths, 400.f, 700.f);
        out[i] = averageSamples(v.c, l0, l1);
    }
}

}  // namespace

SkinLayerSpectra ComputeSkinLayerSpectra(const SkinParams &p) {
    // materials/skincoeffs.h:46-135
    Wld baseline = tabulate([](float wl) { return 0.244f + 85.3f * expf(-(wl - 154.f) / 66.2f); });
    Wld eumel = tabulate([](float wl) { return 6.6e11f * powf(wl, -3.33f); });
    Wld pheomel = tabulate([](float wl) { return 2.9e15f * powf(wl, -4.75f); });
    Wld mie = tabulate([](float wl) { return 147.4f * powf(wl, (float)-0.22); });
    Wld rayleigh = tabulate([](float wl) { return 2e12f * powf(wl, -4.f); });
    float fEu = p.f_eu, fMel = p.f_mel, fOhg = p.f_ohg, fBlood = p.f_blood;
    Wld melanin = sum(scaled(eumel, fEu), scaled(pheomel, 1 - fEu));
    Wld muaEpi = sum(scaled(melanin, fMel), scaled(baseline, 1 - fMel));
    Wld muspEpi = sum(rayleigh, mie);
    const float ln10 = 2.303f, molarWeight = 64500.f, concentration = 150.f;
    Wld haem = sum(scaled(fromTable(kOxyHaemoglobin61), fOhg), scaled(fromTable(kDeoxyHaemoglobin61), 1.f - fOhg));
    Wld blood = scaled(haem, ln10 / molarWeight * concentration);
    Wld muaDerm = sum(scaled(blood, fBlood), scaled(baseline, 1 - fBlood));
    Wld muspDerm = scaled(muspEpi, 0.5f);

    // materials/layeredskin.cpp:49-62: per-cm coefficients -> scene units
    const float NM_PER_CM = 1e7f;
    float unit = p.nmperunit / NM_PER_CM;
    SkinLayerSpectra s;
    toBins(scaled(muaEpi, unit), s.mua[0]);
    toBins(scaled(muspEpi, unit), s.musp[0]);
    toBins(scaled(muaDerm, unit), s.mua[1]);
    toBins(scaled(muspDerm, unit), s.musp[1]);
    for (int l = 0; l < 2; ++l) {
        s.eta[l] = p.layerIor[l];
        s.thickness[l] = p.layerThickness[l] / p.nmperunit;
    }
    return s;
}

void BuildProfileJobs(const SkinLayerSpectra &s, int desiredLength, bool lerpOnThinSlab,
                      std::vector<skc_layer_spec> &specs, std::vector<uint32_t> &layersPerSet,
                      std::vector<skc_mpc_options> &opts) {
    const int layers = 2;
    for (int w = 0; w < kWavelengths; ++w) {
        float mfpTotal = 0.f;
        for (int l = 0; l < layers; ++l) mfpTotal += 1.f / (s.mua[l][w] + s.musp[l][w]);
        float mfp = mfpTotal / (float)layers;
        for (int l = 0; l < layers; ++l) specs.push_back({s.eta[l], s.thickness[l], s.mua[l][w], s.musp[l][w], 0.f});
        layersPerSet.push_back(layers);
        skc_mpc_options o;
        o.desiredLength = (uint32_t)desiredLength;
        o.lerpOnThinSlab = lerpOnThinSlab ? 1 : 0;
        o.desiredStepSize = 12.f * mfp / (float)o.desiredLength;
        opts.push_back(o);
    }
}

int ComputeMultipoleProfile(skc_ctx *ctx, const SkinLayerSpectra &s, int desiredLength, bool lerpOnThinSlab,
                            ProfileTable &out, skc_comm *comm) {
    std::vector<skc_layer_spec> specs;
    std::vector<uint32_t> lps;
    std::vector<skc_mpc_options> opts;
    BuildProfileJobs(s, desiredLength, lerpOnThinSlab, specs, lps, opts);
    // Several GPUs (first row): the wavelengths are independent, so rank r computes wavelengths r, r + world, ...
    // and the finished tables (+ the last d2 and the totals of each) are all-gathered; a rank's block holds `per` table slots.
    const int world = comm ? skc_comm_world(comm) : 1, rank = comm ? skc_comm_rank(comm) : 0;
    const int per = (kWavelengths + world - 1) / world;
    std::vector<skc_layer_spec> mySpecs;
    std::vector<uint32_t> myLps;
    std::vector<skc_mpc_options> myOpts;
    std::vector<int> mine;
    {
        size_t at = 0;
        for (int w = 0; w < kWavelengths; ++w) {
            if (w % world == rank) {
                mine.push_back(w);
                mySpecs.insert(mySpecs.end(), specs.begin() + at, specs.begin() + at + lps[w]);
                myLps.push_back(lps[w]); myOpts.push_back(opts[w]);
            }
            at += lps[w];
        }
    }
    uint32_t len = 0;      // the same for every wavelength: it depends on desiredLength only
    int rc = skc_profile_batch(ctx, specs.data(), lps.data(), opts.data(), 1, 1, &len, nullptr, nullptr, nullptr, nullptr);
    if (rc) return rc;
    const size_t slot = (size_t)len + 3;                              // table, last d2, total R, total T
    std::vector<float> blocks((size_t)world * per * slot, 0.f);
    float *myBlock = blocks.data() + (size_t)rank * per * slot;
    if (!mine.empty()) {
        const uint32_t n = (uint32_t)mine.size();
        std::vector<float> d2((size_t)n * len), rd((size_t)n * len), totals(2 * (size_t)n);
        rc = skc_profile_batch(ctx, mySpecs.data(), myLps.data(), myOpts.data(), n, 1, &len, d2.data(), rd.data(), nullptr, totals.data());
        if (rc) return rc;
        for (uint32_t k = 0; k < n; ++k) {
            float *dst = myBlock + (size_t)k * slot;
            memcpy(dst, rd.data() + (size_t)k * len, len * sizeof(float));
            dst[len] = d2[(size_t)k * len + len - 1]; dst[len + 1] = totals[2 * k]; dst[len + 2] = totals[2 * k + 1];
        }
    }
    if (comm && (rc = skc_comm_allgather_host(comm, blocks.data(), (uint64_t)per * slot * sizeof(float)))) return rc;
    out.len = len;
    out.rd.resize((size_t)kWavelengths * len);
    for (int w = 0; w < kWavelengths; ++w) {      // core/multipole.cpp:275-280
        const float *src = blocks.data() + ((size_t)(w % world) * per + (size_t)(w / world)) * slot;
        memcpy(out.rd.data() + (size_t)w * len, src, len * sizeof(float));
        const float last = src[len];
        out.dsqSpacing[w] = last / (float)(len - 1);
        out.rcpDsqSpacing[w] = (float)(len - 1) / last;
        out.totalReflectance[w] = src[len + 1];
        out.totalTransmittance[w] = src[len + 2];
    }
    return 0;
}

}  // namespace skin

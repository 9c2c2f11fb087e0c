// skc_texture.cuh -- the skin material's "texture albedo" on the device: MIPMap<RGBSpectrum>::Lookup (core/mipmap.h:253-365: bilinear
// `triangle`, the EWA filter over two pyramid levels), UVMapping2D, and Spectrum::FromRGB's reflectance branch
// (core/spectrum.cpp:103-145), in the reference's operation order.  TEXTURE_REPEAT, EWA (doTrilinear false), maxAnisotropy 8:
// CreateImageSpectrumTexture's defaults (textures/imagemap.cpp).
#pragma once
#include "skc_common.cuh"

__device__ __forceinline__ int texWrap(int a, int b) { int n = a / b; a -= n * b; if (a < 0) a += b; return a; }
__device__ __forceinline__ f3 texTexel(const AlbedoTexDev &T, int level, int s, int t) {
    const float *c = T.rgb + T.off[level] + 3 * ((size_t)texWrap(t, T.h[level]) * T.w[level] + texWrap(s, T.w[level]));
    return mk3(__ldg(c), __ldg(c + 1), __ldg(c + 2));
}
__device__ __forceinline__ f3 texAxpy(f3 acc, float w, f3 c) { return mk3(acc.x + c.x * w, acc.y + c.y * w, acc.z + c.z * w); }
__device__ inline f3 texBilinear(const AlbedoTexDev &T, int level, float s, float t) {       // MIPMap::triangle
    level = min(max(level, 0), T.nLevels - 1);
    s = s * (float)T.w[level] - 0.5f;
    t = t * (float)T.h[level] - 0.5f;
    const int s0 = (int)floorf(s), t0 = (int)floorf(t);
    const float ds = s - (float)s0, dt = t - (float)t0;
    const f3 a = texTexel(T, level, s0, t0), b = texTexel(T, level, s0, t0 + 1), c = texTexel(T, level, s0 + 1, t0), d = texTexel(T, level, s0 + 1, t0 + 1);
    const float wa = (1.f - ds) * (1.f - dt), wb = (1.f - ds) * dt, wc = ds * (1.f - dt), wd = ds * dt;
    return mk3(a.x * wa + b.x * wb + c.x * wc + d.x * wd, a.y * wa + b.y * wb + c.y * wc + d.y * wd, a.z * wa + b.z * wb + c.z * wc + d.z * wd);
}
__device__ inline f3 texEWA(const AlbedoTexDev &T, int level, float s, float t, float ds0, float dt0, float ds1, float dt1) {
    if (level >= T.nLevels) return texTexel(T, T.nLevels - 1, 0, 0);
    const float us = (float)T.w[level], vs = (float)T.h[level];
    s = s * us - 0.5f;
    t = t * vs - 0.5f;
    ds0 *= us; dt0 *= vs; ds1 *= us; dt1 *= vs;
    float A = dt0 * dt0 + dt1 * dt1 + 1;
    float B = -2.f * (ds0 * dt0 + ds1 * dt1);
    float C = ds0 * ds0 + ds1 * ds1 + 1;
    const float invF = 1.f / (A * C - B * B * 0.25f);
    A *= invF; B *= invF; C *= invF;
    const float det = -B * B + 4.f * A * C;
    const float invDet = 1.f / det;
    const float uSqrt = sqrtf(det * C), vSqrt = sqrtf(A * det);
    const int s0 = (int)ceilf(s - 2.f * invDet * uSqrt), s1 = (int)floorf(s + 2.f * invDet * uSqrt);
    const int t0 = (int)ceilf(t - 2.f * invDet * vSqrt), t1 = (int)floorf(t + 2.f * invDet * vSqrt);
    f3 sum = mk3(0.f, 0.f, 0.f);
    float sumWts = 0.f;
    for (int it = t0; it <= t1; ++it) {
        const float tt = (float)it - t;
        for (int is = s0; is <= s1; ++is) {
            const float ss = (float)is - s;
            const float r2 = A * ss * ss + B * ss * tt + C * tt * tt;
            if (r2 < 1.) {
                const float weight = __ldg(&T.lut[min((int)(r2 * 128), 127)]);
                sum = texAxpy(sum, weight, texTexel(T, level, is, it));
                sumWts += weight;
            }
        }
    }
    return mk3(sum.x / sumWts, sum.y / sumWts, sum.z / sumWts);
}
// MIPMap::Lookup(s, t, ds0, dt0, ds1, dt1), EWA
__device__ inline f3 texLookup(const AlbedoTexDev &T, float s, float t, float ds0, float dt0, float ds1, float dt1) {
    if (ds0 * ds0 + dt0 * dt0 < ds1 * ds1 + dt1 * dt1) { float x = ds0; ds0 = ds1; ds1 = x; x = dt0; dt0 = dt1; dt1 = x; }
    const float majorLength = sqrtf(ds0 * ds0 + dt0 * dt0);
    float minorLength = sqrtf(ds1 * ds1 + dt1 * dt1);
    if (minorLength * 8.f < majorLength && minorLength > 0.f) {
        const float scale = majorLength / (minorLength * 8.f);
        ds1 *= scale; dt1 *= scale; minorLength *= scale;
    }
    if (minorLength == 0.f) return texBilinear(T, 0, s, t);
    const float invLog2 = 1.f / skc_logf(2.f);
    const float lod = fmaxf(0.f, (float)T.nLevels - 1.f + skc_logf(minorLength) * invLog2);
    const int ilod = (int)floorf(lod);
    const float d = lod - (float)ilod;
    const f3 a = texEWA(T, ilod, s, t, ds0, dt0, ds1, dt1), b = texEWA(T, ilod + 1, s, t, ds0, dt0, ds1, dt1);
    return mk3(a.x * (1.f - d) + b.x * d, a.y * (1.f - d) + b.y * d, a.z * (1.f - d) + b.z * d);
}
// Spectrum::FromRGB(rgb, SPECTRUM_REFLECTANCE)[lane] (core/spectrum.cpp:103-145); refl = [7][32] white, cyan, magenta, yellow, red, green, blue
__device__ __forceinline__ float texReflectance(const float *__restrict__ refl, float r, float g, float b, int lane) {
    enum { W, C, M, Y, R, G, B };
    float a, d1, d2; int i1, i2;
    if (r <= g && r <= b) {
        a = r;
        if (g <= b) { d1 = g - r; i1 = C; d2 = b - g; i2 = B; } else { d1 = b - r; i1 = C; d2 = g - b; i2 = G; }
    } else if (g <= r && g <= b) {
        a = g;
        if (r <= b) { d1 = r - g; i1 = M; d2 = b - r; i2 = B; } else { d1 = b - g; i1 = M; d2 = r - b; i2 = R; }
    } else {
        a = b;
        if (r <= g) { d1 = r - b; i1 = Y; d2 = g - r; i2 = G; } else { d1 = g - b; i1 = Y; d2 = r - g; i2 = R; }
    }
    float v = 0.f;
    v += a * __ldg(&refl[W * 32 + lane]);
    v += d1 * __ldg(&refl[i1 * 32 + lane]);
    v += d2 * __ldg(&refl[i2 * 32 + lane]);
    v *= .94f;
    return fmaxf(v, 0.f);
}

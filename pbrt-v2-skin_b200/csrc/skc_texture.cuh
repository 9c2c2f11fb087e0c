// skc_texture.cuh -- the skin material's "texture albedo" on the device: MIPMap<RGBSpectrum>::Lookup (core/mipmap.h:253-365: bilinear
// `triangle`, the EWA filter over two pyramid levels), UVMapping2D, and Spectrum::FromRGB's reflectance branch
// (core/spectrum.cpp:103-145), in the reference's operation order.  TEXTURE_REPEAT, EWA (doTrilinear false), maxAnisotropy 8:
// CreateImageSpectrumTexture's defaults (textures/imagemap.cpp).
#pragma once
#include "skc_common.cuh"
#include "skc_bvh.cuh"

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

// Material::Bump of this fork (core/material.cpp:47-104: central differences, five evaluations of the displacement texture) on the shading
// geometry of a point of triangle vi with barycentrics b0, b1, b2 (Triangle::GetShadingGeometry, shapes/trianglemesh.inl:224-299, incl.
// dndu / dndv).  (dudx, dvdx, dudy, dvdy): the screen-space differentials of (u, v) -- zero for the tessellator's points, from
// ComputeDifferentials for camera hits.  Returns the bumped shading normal and dpdu (what LayeredSkin::GetBSDF builds its frame from).
// The reference negates du / dv for its backward samples and then divides by 2 * du with the negated value, and takes its backward u
// sample at v + dv: both kept.
__device__ inline void triBump(const MeshView &mv, int4 vi, float b0, float b1, float b2, const AlbedoTexDev &tex, float dudx, float dvdx, float dudy, float dvdy,
                               f3 *nnOut, f3 *dpduOut) {
    f3 dpduG, dpdvG, ng; float tu, tv;
    triDiffGeom(mv, vi, b0, b1, b2, &dpduG, &dpdvG, &tu, &tv, &ng);
    f3 ss = dpduG, ts = dpdvG, nns = ng, dndu = mk3(0.f, 0.f, 0.f), dndv = dndu;
    if (mv.N) {
        float uvs[3][2];
        if (mv.UV) {
            uvs[0][0] = mv.UV[2 * vi.x]; uvs[0][1] = mv.UV[2 * vi.x + 1]; uvs[1][0] = mv.UV[2 * vi.y]; uvs[1][1] = mv.UV[2 * vi.y + 1];
            uvs[2][0] = mv.UV[2 * vi.z]; uvs[2][1] = mv.UV[2 * vi.z + 1];
        } else { uvs[0][0] = 0.f; uvs[0][1] = 0.f; uvs[1][0] = 1.f; uvs[1][1] = 0.f; uvs[2][0] = 1.f; uvs[2][1] = 1.f; }
        float A00 = uvs[1][0] - uvs[0][0], A01 = uvs[2][0] - uvs[0][0], A10 = uvs[1][1] - uvs[0][1], A11 = uvs[2][1] - uvs[0][1];
        float C0 = tu - uvs[0][0], C1 = tv - uvs[0][1];
        float bb[3];
        float det = A00 * A11 - A01 * A10;
        bool ok = !(fabsf(det) < 1e-10f);
        if (ok) { bb[1] = (A11 * C0 - A01 * C1) / det; bb[2] = (A00 * C1 - A10 * C0) / det; if (isnan(bb[1]) || isnan(bb[2])) ok = false; }
        if (!ok) bb[0] = bb[1] = bb[2] = 1.f / 3.f; else bb[0] = 1.f - bb[1] - bb[2];
        const f3 n0 = ld3(mv.N, vi.x), n1 = ld3(mv.N, vi.y), n2 = ld3(mv.N, vi.z);
        const f3 ns = normalize3(xformNormal(mv.w2o, add3(add3(mul3(n0, bb[0]), mul3(n1, bb[1])), mul3(n2, bb[2]))));
        ss = normalize3(dpduG);
        ts = cross3(ss, ns);
        if (ts.x * ts.x + ts.y * ts.y + ts.z * ts.z > 0.f) { ts = normalize3(ts); ss = cross3(ts, ns); }
        else coordinateSystem(ns, &ss, &ts);
        const float du1 = uvs[0][0] - uvs[2][0], du2 = uvs[1][0] - uvs[2][0], dv1 = uvs[0][1] - uvs[2][1], dv2 = uvs[1][1] - uvs[2][1];
        const f3 dn1 = sub3(n0, n2), dn2 = sub3(n1, n2);
        const float determinant = du1 * dv2 - dv1 * du2;
        if (determinant != 0.f) {
            const float invdet = 1.f / determinant;
            dndu = xformNormal(mv.w2o, mul3(sub3(mul3(dn1, dv2), mul3(dn2, dv1)), invdet));
            dndv = xformNormal(mv.w2o, mul3(add3(mul3(dn1, -du2), mul3(dn2, du1)), invdet));
        }
        nns = normalize3(cross3(ss, ts));
        if (mv.flip) nns = mul3(nns, -1.f);
    }
    const float dsdx = tex.su * dudx, dtdx = tex.sv * dvdx, dsdy = tex.su * dudy, dtdy = tex.sv * dvdy;
    auto eval = [&](float u, float v) { return texLookup(tex, tex.su * u + tex.du, tex.sv * v + tex.dv, dsdx, dtdx, dsdy, dtdy).x; };
    float du = fabsf(dudx) + fabsf(dudy);
    if (du == 0.f) du = .01f;
    const float upDisplace = eval(tu + du, tv);
    float dv = fabsf(dvdx) + fabsf(dvdy);
    if (dv == 0.f) dv = .01f;
    const float vpDisplace = eval(tu, tv + dv);
    const float displace = eval(tu, tv);
    du = -du;
    const float unDisplace = eval(tu + du, tv + dv);               // the reference does not reset dgEval.v here: (u - du, v + dv)
    dv = -dv;
    const float vnDisplace = eval(tu, tv + dv);
    const f3 dpdu = add3(add3(ss, mul3(nns, (upDisplace - unDisplace) / (2 * du))), mul3(dndu, displace));
    const f3 dpdv = add3(add3(ts, mul3(nns, (vpDisplace - vnDisplace) / (2 * dv))), mul3(dndv, displace));
    f3 nn = normalize3(cross3(dpdu, dpdv));
    if (mv.flip) nn = mul3(nn, -1.f);
    if (dot3(nn, ng) < 0.f) nn = mul3(nn, -1.f);                    // Faceforward(dgBump->nn, dgGeom.nn)
    *nnOut = nn; *dpduOut = dpdu;
}

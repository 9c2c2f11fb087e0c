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
// The elliptical weighted average of one pyramid level (core/mipmap.h:312-365).  The footprint's two axes (ax, ay) and (bx, by), in
// texels of the level, give the quadric q(dx, dy) = qxx dx^2 + qxy dx dy + qyy dy^2 (normalised so that q = 1 is the ellipse); texels
// with q < 1 inside its bounding box are weighted by the 128-entry table of exp(-2 q) - exp(-2).  Same operations in the same order
// as the reference, so the filtered value is the reference's bit for bit.
struct EwaQuadric { float qxx, qxy, qyy; };
__device__ __forceinline__ EwaQuadric ewaQuadric(float ax, float ay, float bx, float by) {
    EwaQuadric q;
    q.qxx = ay * ay + by * by + 1;
    q.qxy = -2.f * (ax * ay + bx * by);
    q.qyy = ax * ax + bx * bx + 1;
    const float norm = 1.f / (q.qxx * q.qyy - q.qxy * q.qxy * 0.25f);
    q.qxx *= norm; q.qxy *= norm; q.qyy *= norm;
    return q;
}
__device__ inline f3 texEWA(const AlbedoTexDev &T, int level, float s, float t, float ax, float ay, float bx, float by) {
    if (level >= T.nLevels) return texTexel(T, T.nLevels - 1, 0, 0);
    const float wTex = (float)T.w[level], hTex = (float)T.h[level];
    const float cx = s * wTex - 0.5f, cy = t * hTex - 0.5f;                  // the lookup point in texel coordinates
    const EwaQuadric q = ewaQuadric(ax * wTex, ay * hTex, bx * wTex, by * hTex);
    const float disc = -q.qxy * q.qxy + 4.f * q.qxx * q.qyy, rcpDisc = 1.f / disc;
    const float halfX = 2.f * rcpDisc * sqrtf(disc * q.qyy), halfY = 2.f * rcpDisc * sqrtf(q.qxx * disc);
    const int xLo = (int)ceilf(cx - halfX), xHi = (int)floorf(cx + halfX), yLo = (int)ceilf(cy - halfY), yHi = (int)floorf(cy + halfY);
    f3 acc = mk3(0.f, 0.f, 0.f);
    float wSum = 0.f;
    for (int y = yLo; y <= yHi; ++y) {
        const float dy = (float)y - cy;
        for (int x = xLo; x <= xHi; ++x) {
            const float dx = (float)x - cx;
            const float r2 = q.qxx * dx * dx + q.qxy * dx * dy + q.qyy * dy * dy;
            if (!(r2 < 1.)) continue;
            const float wgt = __ldg(&T.lut[min((int)(r2 * 128), 127)]);
            acc = texAxpy(acc, wgt, texTexel(T, level, x, y));
            wSum += wgt;
        }
    }
    return mk3(acc.x / wSum, acc.y / wSum, acc.z / wSum);
}
// MIPMap::Lookup(s, t, ds0, dt0, ds1, dt1) with the EWA filter (core/mipmap.h:277-309): the longer of the two footprint vectors is the
// major axis; the minor one is lengthened until the anisotropy is at most 8; its length picks the two levels that are blended.
__device__ inline f3 texLookup(const AlbedoTexDev &T, float s, float t, float ax, float ay, float bx, float by) {
    if (ax * ax + ay * ay < bx * bx + by * by) { float x = ax; ax = bx; bx = x; x = ay; ay = by; by = x; }
    const float lenMajor = sqrtf(ax * ax + ay * ay);
    float lenMinor = sqrtf(bx * bx + by * by);
    if (lenMinor * 8.f < lenMajor && lenMinor > 0.f) {
        const float stretch = lenMajor / (lenMinor * 8.f);
        bx *= stretch; by *= stretch; lenMinor *= stretch;
    }
    if (lenMinor == 0.f) return texBilinear(T, 0, s, t);
    const float rcpLn2 = 1.f / skc_logf(2.f);                               // Log2(x) = logf(x) * (1.f / logf(2.f)), core/pbrt.h
    const float lod = fmaxf(0.f, (float)T.nLevels - 1.f + skc_logf(lenMinor) * rcpLn2);
    const int fine = (int)floorf(lod);
    const float blend = lod - (float)fine;
    const f3 lo = texEWA(T, fine, s, t, ax, ay, bx, by), hi = texEWA(T, fine + 1, s, t, ax, ay, bx, by);
    return mk3(lo.x * (1.f - blend) + hi.x * blend, lo.y * (1.f - blend) + hi.y * blend, lo.z * (1.f - blend) + hi.z * blend);
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

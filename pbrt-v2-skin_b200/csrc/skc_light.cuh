// skc_light.cuh -- device helpers shared by the irradiance pass (skc_irradiance.cu) and the direct-lighting
// pass (skc_frame.cu): MT19937 twist, (0,2)-sequence, sphere area-light sampling.
// Reference: core/rng.cpp:84-110, core/montecarlo.h:278-301, lights/diffuse.cpp:69-85,
// core/light.cpp:145-166, shapes/sphere.cpp:58-158,236-288, core/shape.cpp:96-109, core/light.h:87-92.
#pragma once
#include "skc_bvh.cuh"

#define MT_N 624
#define MT_M 397

__device__ __forceinline__ uint32_t mtTwist(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t y = (a & 0x80000000u) | (b & 0x7fffffffu);
    return c ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
}

// core/montecarlo.h:285-301
__device__ __forceinline__ float vanDerCorput(uint32_t n, uint32_t scramble) {
    n = __brev(n);
    n ^= scramble;
    return fminf(((n >> 8) & 0xffffff) / float(1 << 24), 0.99999994f /* OneMinusEpsilon */);
}
__device__ __forceinline__ float sobol2(uint32_t n, uint32_t scramble) {
    for (uint32_t v = 1u << 31; n != 0; n >>= 1, v ^= v >> 1)
        if (n & 0x1) scramble ^= v;
    return fminf(((scramble >> 8) & 0xffffff) / float(1 << 24), 0.99999994f);
}

__device__ inline void coordSys(f3 v1, f3 *v2, f3 *v3) {   // core/geometry.h:564-574
    if (fabsf(v1.x) > fabsf(v1.y)) {
        float invLen = 1.f / sqrtf(v1.x * v1.x + v1.z * v1.z);
        *v2 = mk3(-v1.z * invLen, 0.f, v1.x * invLen);
    } else {
        float invLen = 1.f / sqrtf(v1.y * v1.y + v1.z * v1.z);
        *v2 = mk3(0.f, v1.z * invLen, -v1.y * invLen);
    }
    *v3 = cross3(v1, *v2);
}

// the reference's libm sinf / cosf, bit for bit (skc_common.cuh)
__device__ __forceinline__ float cosf_cr(float x) { return skc_cosf(x); }
__device__ __forceinline__ float sinf_cr(float x) { return skc_sinf(x); }

// Sphere::Intersect's DifferentialGeometry normal (shapes/sphere.cpp:103-150 +
// core/diffgeom.cpp:40-55), needed by ShapeSet::Sample (core/light.cpp:156).
__device__ inline f3 sphereDgNormal(const skc_light &s, f3 phit /*object space, after phi fixups*/, float phi, f3 *dpduOut = nullptr) {
    (void)phi;
    float radius = s.radius;
    float theta = acosf(fminf(fmaxf(phit.z / radius, -1.f), 1.f));
    float zradius = sqrtf(phit.x * phit.x + phit.y * phit.y);
    float invzradius = 1.f / zradius;
    float cosphi = phit.x * invzradius, sinphi = phit.y * invzradius;
    f3 dpdu = mk3(-s.phiMax * phit.y, s.phiMax * phit.x, 0.f);
    float dth = s.thetaMax - s.thetaMin;
    f3 dpdv = mul3(mk3(phit.z * cosphi, phit.z * sinphi, -radius * sinf(theta)), dth);
    f3 wu = xformVector(s.o2w, dpdu), wv = xformVector(s.o2w, dpdv);
    if (dpduOut) *dpduOut = wu;
    return normalize3(cross3(wu, wv));   // orientation flip applied by the caller
}

// Sphere::Intersect restricted to what light sampling needs: thit and (optionally) dg.nn.
static __device__ __noinline__ bool sphereIntersectDg(const skc_light &s, const DRay &r, float *tHit, f3 *nn, f3 *dpduOut = nullptr) {
    DRay ray;
    ray.o = xformPoint(s.w2o, r.o);
    ray.d = xformVector(s.w2o, r.d);
    ray.mint = r.mint; ray.maxt = r.maxt;
    float radius = s.radius;
    float A = ray.d.x * ray.d.x + ray.d.y * ray.d.y + ray.d.z * ray.d.z;
    float B = 2 * (ray.d.x * ray.o.x + ray.d.y * ray.o.y + ray.d.z * ray.o.z);
    float C = ray.o.x * ray.o.x + ray.o.y * ray.o.y + ray.o.z * ray.o.z - radius * radius;
    float t0, t1;
    if (!quadratic(A, B, C, &t0, &t1)) return false;
    if (t0 > ray.maxt || t1 < ray.mint) return false;
    float thit = t0;
    if (t0 < ray.mint) { thit = t1; if (thit > ray.maxt) return false; }
    f3 phit = add3(ray.o, mul3(ray.d, thit));
    if (phit.x == 0.f && phit.y == 0.f) phit.x = 1e-5f * radius;
    float phi = atan2f(phit.y, phit.x);
    if (phi < 0.f) phi += 2.f * SKC_PI_F;
    float zmin = s.zmin, zmax = s.zmax, phiMax = s.phiMax;
    if ((zmin > -radius && phit.z < zmin) || (zmax < radius && phit.z > zmax) || phi > phiMax) {
        if (thit == t1) return false;
        if (t1 > ray.maxt) return false;
        thit = t1;
        phit = add3(ray.o, mul3(ray.d, thit));
        if (phit.x == 0.f && phit.y == 0.f) phit.x = 1e-5f * radius;
        phi = atan2f(phit.y, phit.x);
        if (phi < 0.f) phi += 2.f * SKC_PI_F;
        if ((zmin > -radius && phit.z < zmin) || (zmax < radius && phit.z > zmax) || phi > phiMax) return false;
    }
    *tHit = thit;
    if (nn) *nn = sphereDgNormal(s, phit, phi, dpduOut);
    return true;
}

// DiffuseAreaLight::Sample_L (lights/diffuse.cpp:69-81) for a one-sphere ShapeSet: wi, pdf and the
// shadow segment; the return value says whether the returned radiance L(ps, ns, -wi) is Lemit (true)
// or black (false).
__device__ inline bool sampleSphereLightL(const skc_light &L, f3 p, float pEps, float u1, float u2,
                                          f3 *wiOut, float *pdfOut, DRay *shadow) {
    f3 Pcenter = xformPoint(L.o2w, mk3(0.f, 0.f, 0.f));
    float radius = L.radius;
    bool reverse = L.reverseOrientation != 0.f;
    f3 pt, ns;
    float dc2 = dist2(p, Pcenter);
    bool inside = dc2 - radius * radius < 1e-4f;
    if (inside) {
        // Sphere::Sample(u1,u2,ns) (sphere.cpp:229-234): uniform over the sphere
        float z = 1.f - 2.f * u1;
        float rr = sqrtf(fmaxf(0.f, 1.f - z * z));
        float phi = 2.f * SKC_PI_F * u2;
        f3 po = mk3(radius * (rr * cosf_cr(phi)), radius * (rr * sinf_cr(phi)), radius * z);
        ns = normalize3(xformNormal(L.w2o, po));
        if (reverse) ns = mul3(ns, -1.f);
        pt = xformPoint(L.o2w, po);
    } else {
        f3 wc = normalize3(sub3(Pcenter, p));
        f3 wcX, wcY;
        coordSys(wc, &wcX, &wcY);
        float sinThetaMax2 = radius * radius / dc2;
        float cosThetaMax = sqrtf(fmaxf(0.f, 1.f - sinThetaMax2));
        // UniformSampleCone (core/montecarlo.cpp:413-420)
        float costheta = (1.f - u1) * cosThetaMax + u1 * 1.f;
        float sintheta = sqrtf(1.f - costheta * costheta);
        float phi = u2 * 2.f * SKC_PI_F;
        f3 dir = add3(add3(mul3(wcX, cosf_cr(phi) * sintheta), mul3(wcY, sinf_cr(phi) * sintheta)), mul3(wc, costheta));
        DRay r; r.o = p; r.d = dir; r.mint = 1e-3f; r.maxt = INFINITY;
        float thit;
        if (!sphereIntersectDg(L, r, &thit, nullptr)) thit = dot3(sub3(Pcenter, p), normalize3(r.d));
        pt = add3(r.o, mul3(r.d, thit));
        ns = normalize3(sub3(pt, Pcenter));
        if (reverse) ns = mul3(ns, -1.f);
    }
    // ShapeSet::Sample (core/light.cpp:145-158): re-intersect p -> pt with the shape set
    DRay r2; r2.o = p; r2.d = sub3(pt, p); r2.mint = 1e-3f; r2.maxt = INFINITY;
    float thit = 1.f, th; f3 nn;
    if (sphereIntersectDg(L, r2, &th, &nn)) { thit = th; if (reverse) nn = mul3(nn, -1.f); ns = nn; }
    f3 ps = add3(r2.o, mul3(r2.d, thit));
    f3 wi = normalize3(sub3(ps, p));
    // ShapeSet::Pdf (light.cpp:161-166) with one shape: area * pdf / sumArea
    float area = L.phiMax * radius * (L.zmax - L.zmin);
    float spdf;
    if (inside) {
        // Shape::Pdf(p, wi) (core/shape.cpp): intersect and convert the area measure
        DRay r3; r3.o = p; r3.d = wi; r3.mint = 1e-3f; r3.maxt = INFINITY;
        float t3; f3 n3;
        if (!sphereIntersectDg(L, r3, &t3, &n3)) spdf = 0.f;
        else {
            f3 ph = add3(r3.o, mul3(r3.d, t3));
            spdf = dist2(p, ph) / (fabsf(dot3(n3, mul3(wi, -1.f))) * area);
            if (isinf(spdf)) spdf = 0.f;
        }
    } else {
        float sinThetaMax2 = radius * radius / dc2;
        float cosThetaMax = sqrtf(fmaxf(0.f, 1.f - sinThetaMax2));
        spdf = 1.f / (2.f * SKC_PI_F * (1.f - cosThetaMax));   // UniformConePdf (core/montecarlo.cpp:400-402)
    }
    float pdf = 0.f + area * spdf;
    pdf = pdf / area;
    // VisibilityTester::SetSegment (core/light.h:87-92)
    float dist = len3(sub3(p, ps));
    float invd = 1.f / dist;
    shadow->o = p; shadow->d = mul3(sub3(ps, p), invd); shadow->mint = pEps; shadow->maxt = dist * (1.f - 1e-3f);
    // L(ps, ns, -wi): lights/diffuse.h:51-53
    bool lit = dot3(ns, mul3(wi, -1.f)) > 0.f;
    *wiOut = wi; *pdfOut = pdf;
    return lit;
}

// The irradiance loop's use of Sample_L (multipolesubsurface.cpp:119-126): false when the sample
// contributes nothing before the shadow test.
__device__ inline bool sampleSphereLight(const skc_light &L, f3 p, float pEps, f3 spn, float u1, float u2,
                                         f3 *wiOut, float *pdfOut, DRay *shadow) {
    bool lit = sampleSphereLightL(L, p, pEps, u1, u2, wiOut, pdfOut, shadow);
    if (dot3(*wiOut, spn) <= 0.f) return false;
    if (!lit || *pdfOut == 0.f) return false;
    return true;
}

// ---- LightSource "infinite" (SURVEY.md 8f N4): lights/infinite.cpp:114-119,195-234, core/montecarlo.h:78-97,151-168, core/mipmap.h:253-264 ----
// Record: radius = -4, o2w / w2o = LightToWorld / WorldToLight, zmin / zmax = the map's size; the map's level-0 texels and its
// Distribution2D are in EnvMapDev (skc_upload_envmap).  The radiance of a direction is an RGB value (a bilinear lookup), turned
// into a 30-sample spectrum per use by Spectrum(RGBSpectrum, SPECTRUM_ILLUMINANT) = FromRGB's illuminant branch
// (core/spectrum.cpp:146-186): the smallest component on the white basis, the two differences on a secondary and a primary one.
#define SKC_INV_PI_F 0.31830988618379067154f
#define SKC_INV_TWOPI_F 0.15915494309189533577f
// illum: [7][32] white, cyan, magenta, yellow, red, green, blue (host/skin_tables.h kRGBIllum2Spect30; EnvMapDev::illum)
__device__ __forceinline__ float envSpectrum(const float *__restrict__ illum, float r, float g, float b, int lane) {
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
    v += a * __ldg(&illum[W * 32 + lane]);
    v += d1 * __ldg(&illum[i1 * 32 + lane]);
    v += d2 * __ldg(&illum[i2 * 32 + lane]);
    v *= .86445f;
    return fmaxf(v, 0.f);          // Clamp()
}

// Distribution1D::SampleContinuous over cdf[0..n]: returns the sample in [0,1), its pdf and the segment
__device__ inline float envSample1D(const float *__restrict__ func, const float *__restrict__ cdf, float funcInt, int n, float u, float *pdf, int *off) {
    int lo = 0, hi = n + 1;                       // std::upper_bound: first element > u
    while (lo < hi) { int mid = (lo + hi) >> 1; if (!(u < __ldg(&cdf[mid]))) lo = mid + 1; else hi = mid; }
    const int offset = max(0, lo - 1);
    *off = offset;
    const float c0 = __ldg(&cdf[offset]), c1 = __ldg(&cdf[offset + 1]);
    const float du = (u - c0) / (c1 - c0);
    *pdf = __ldg(&func[offset]) / funcInt;
    return ((float)offset + du) / (float)n;
}
__device__ inline void envSample(const EnvMapDev &e, float u0, float u1, float *uv0, float *uv1, float *pdf) {      // Distribution2D::SampleContinuous
    float p0, p1; int v, dummy;
    *uv1 = envSample1D(e.margFunc, e.margCdf, e.margInt, e.h, u1, &p1, &v);
    *uv0 = envSample1D(e.condFunc + (size_t)v * e.w, e.condCdf + (size_t)v * (e.w + 1), __ldg(&e.condInt[v]), e.w, u0, &p0, &dummy);
    *pdf = p0 * p1;
}
__device__ inline float envDistPdf(const EnvMapDev &e, float u, float v) {                                            // Distribution2D::Pdf
    const int iu = min(max((int)(u * (float)e.w), 0), e.w - 1), iv = min(max((int)(v * (float)e.h), 0), e.h - 1);
    const float ci = __ldg(&e.condInt[iv]);
    if (ci * e.margInt == 0.f) return 0.f;
    return (__ldg(&e.condFunc[(size_t)iv * e.w + iu]) * __ldg(&e.margFunc[iv])) / (ci * e.margInt);
}
__device__ __forceinline__ int envWrap(int a, int b) { int n = a / b; a -= n * b; if (a < 0) a += b; return a; }
// MIPMap::Lookup(s, t) with width 0 = triangle(0, s, t), TEXTURE_REPEAT
__device__ inline f3 envLookup(const EnvMapDev &e, float s, float t) {
    s = s * (float)e.w - 0.5f;
    t = t * (float)e.h - 0.5f;
    const int s0 = (int)floorf(s), t0 = (int)floorf(t);
    const float ds = s - (float)s0, dt = t - (float)t0;
    const int sa = envWrap(s0, e.w), sb = envWrap(s0 + 1, e.w), ta = envWrap(t0, e.h), tb = envWrap(t0 + 1, e.h);
    const float w00 = (1.f - ds) * (1.f - dt), w01 = (1.f - ds) * dt, w10 = ds * (1.f - dt), w11 = ds * dt;
    const float *c00 = e.rgb + 3 * ((size_t)ta * e.w + sa), *c01 = e.rgb + 3 * ((size_t)tb * e.w + sa);
    const float *c10 = e.rgb + 3 * ((size_t)ta * e.w + sb), *c11 = e.rgb + 3 * ((size_t)tb * e.w + sb);
    f3 o;
    o.x = __ldg(c00) * w00 + __ldg(c01) * w01 + __ldg(c10) * w10 + __ldg(c11) * w11;
    o.y = __ldg(c00 + 1) * w00 + __ldg(c01 + 1) * w01 + __ldg(c10 + 1) * w10 + __ldg(c11 + 1) * w11;
    o.z = __ldg(c00 + 2) * w00 + __ldg(c01 + 2) * w01 + __ldg(c10 + 2) * w10 + __ldg(c11 + 2) * w11;
    return o;
}
__device__ __forceinline__ float sphericalTheta(f3 v) { return skc_acosf(fminf(fmaxf(v.z, -1.f), 1.f)); }
__device__ __forceinline__ float sphericalPhi(f3 v) { const float p = skc_atan2f(v.y, v.x); return p < 0.f ? p + 2.f * SKC_PI_F : p; }

// InfiniteAreaLight::Sample_L (lights/infinite.cpp:195-221): wi, pdf, the map's RGB in that direction, the VisibilityTester's ray
__device__ inline bool sampleInfiniteLight(const skc_light &L, const EnvMapDev &e, f3 p, float pEps, float u0, float u1, f3 *wi, float *pdf, f3 *rgb, DRay *shadow) {
    float uv0, uv1, mapPdf;
    envSample(e, u0, u1, &uv0, &uv1, &mapPdf);
    if (mapPdf == 0.f) return false;
    const float theta = uv1 * SKC_PI_F, phi = uv0 * 2.f * SKC_PI_F;
    const float costheta = skc_cosf(theta), sintheta = skc_sinf(theta);
    const float sinphi = skc_sinf(phi), cosphi = skc_cosf(phi);
    *wi = xformVector(L.o2w, mk3(sintheta * cosphi, sintheta * sinphi, costheta));
    *pdf = mapPdf / (2.f * SKC_PI_F * SKC_PI_F * sintheta);
    if (sintheta == 0.f) *pdf = 0.f;
    shadow->o = p; shadow->d = *wi; shadow->mint = pEps; shadow->maxt = INFINITY;      // VisibilityTester::SetRay (core/light.h:94-99)
    *rgb = envLookup(e, uv0, uv1);
    return true;
}
// InfiniteAreaLight::Pdf (lights/infinite.cpp:224-234)
__device__ inline float infiniteLightPdf(const skc_light &L, const EnvMapDev &e, f3 w) {
    const f3 wi = xformVector(L.w2o, w);
    const float theta = sphericalTheta(wi), phi = sphericalPhi(wi);
    const float sintheta = skc_sinf(theta);
    if (sintheta == 0.f) return 0.f;
    return envDistPdf(e, phi * SKC_INV_TWOPI_F, theta * SKC_INV_PI_F) / (2.f * SKC_PI_F * SKC_PI_F * sintheta);
}
// InfiniteAreaLight::Le (lights/infinite.cpp:114-119): the map's RGB for a ray that leaves the scene
__device__ inline f3 infiniteLightLe(const skc_light &L, const EnvMapDev &e, f3 d) {
    const f3 wh = normalize3(xformVector(L.w2o, d));
    return envLookup(e, sphericalPhi(wh) * SKC_INV_TWOPI_F, sphericalTheta(wh) * SKC_INV_PI_F);
}

// ---- delta lights: lights/point.cpp:49-56, lights/spot.cpp:52-72, lights/distant.cpp:49-56 -------------
// They reuse the 70-float record (include/skin_cuda.h): radius = -1 point, -2 spot, -3 distant; o2w / w2o = LightToWorld and
// its inverse; spot: zmin = cosTotalWidth, zmax = cosFalloffStart; distant: (zmin, zmax, phiMax) = lightDir; Lemit = the
// intensity I (point, spot) or the radiance L (distant); nSamples = 1.
__device__ __forceinline__ int lightKind(const skc_light &L) { return L.radius < 0.f ? (int)(-L.radius) : 0; }

__device__ inline float spotFalloff(const skc_light &L, f3 w) {          // SpotLight::Falloff (lights/spot.cpp:62-72)
    f3 wl = normalize3(xformVector(L.w2o, w));
    float costheta = wl.z;
    if (costheta < L.zmin) return 0.f;
    if (costheta > L.zmax) return 1.f;
    float delta = (costheta - L.zmin) / (L.zmax - L.zmin);
    return delta * delta * delta * delta;
}

// Sample_L of a delta light: wi, the shadow ray of its VisibilityTester (core/light.h:87-99 SetSegment / SetRay), and the two
// scalars that turn the stored spectrum into Li: Li = (Lemit * falloff) / d2 (d2 = 1 for a distant light, falloff = 1
// unless spot; x * 1 and x / 1 are exact, so one expression serves the three).  Returns false when Li is black.
__device__ inline bool sampleDeltaLight(const skc_light &L, int kind, f3 p, float pEps, f3 *wiOut, float *d2Out, float *falloffOut, DRay *shadow) {
    *d2Out = 1.f; *falloffOut = 1.f;
    if (kind == 3) {
        f3 wi = mk3(L.zmin, L.zmax, L.phiMax);
        *wiOut = wi;
        shadow->o = p; shadow->d = wi; shadow->mint = pEps; shadow->maxt = INFINITY;
        return true;
    }
    f3 lightPos = xformPoint(L.o2w, mk3(0.f, 0.f, 0.f));
    f3 wi = normalize3(sub3(lightPos, p));
    *wiOut = wi;
    *d2Out = dist2(lightPos, p);
    if (kind == 2) {
        float f = spotFalloff(L, mul3(wi, -1.f));
        *falloffOut = f;
        if (f == 0.f) return false;
    }
    // VisibilityTester::SetSegment(p, pEpsilon, lightPos, 0.) (core/light.h:87-92)
    float dist = len3(sub3(p, lightPos));
    float invd = 1.f / dist;
    shadow->o = p; shadow->d = mul3(sub3(lightPos, p), invd); shadow->mint = pEps; shadow->maxt = dist * (1.f - 0.f);
    return true;
}

// DiffuseAreaLight::Pdf(p, wi) -> ShapeSet::Pdf -> Sphere::Pdf (lights/diffuse.cpp:84-86, core/light.cpp:161-166,
// shapes/sphere.cpp:276-288, core/shape.cpp:96-109)
__device__ inline float sphereLightPdf(const skc_light &L, f3 p, f3 wi) {
    f3 Pcenter = xformPoint(L.o2w, mk3(0.f, 0.f, 0.f));
    float radius = L.radius;
    float area = L.phiMax * radius * (L.zmax - L.zmin);
    float dc2 = dist2(p, Pcenter);
    float spdf;
    if (dc2 - radius * radius < 1e-4f) {
        DRay r3; r3.o = p; r3.d = wi; r3.mint = 1e-3f; r3.maxt = INFINITY;
        float t3; f3 n3;
        if (!sphereIntersectDg(L, r3, &t3, &n3)) spdf = 0.f;
        else {
            f3 ph = add3(r3.o, mul3(r3.d, t3));
            spdf = dist2(p, ph) / (fabsf(dot3(n3, mul3(wi, -1.f))) * area);
            if (isinf(spdf)) spdf = 0.f;
        }
    } else {
        float sinThetaMax2 = radius * radius / dc2;
        float cosThetaMax = sqrtf(fmaxf(0.f, 1.f - sinThetaMax2));
        spdf = 1.f / (2.f * SKC_PI_F * (1.f - cosThetaMax));
    }
    float pdf = 0.f + area * spdf;
    return pdf / area;
}


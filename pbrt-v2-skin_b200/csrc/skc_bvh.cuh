// skc_bvh.cuh -- device-side BVH traversal and primitive tests.
//
// Follows accelerators/bvh.cpp:126-148 (slab test), :388-439 (Intersect), :442-488
// (IntersectP); shapes/trianglemesh.inl:55-136,140-210 (triangle); shapes/sphere.cpp:58-210
// (quadric sphere) of the reference, operation for operation, so that hit/no-hit, the hit
// primitive and t are bit-identical to the CPU code.
//
// Device layout: a node is two float4 (32 B, one 32 B sector):
//   n0 = {min.x, min.y, min.z, max.x}   n1 = {max.y, max.z, offset, nPrims | axis << 8}
// a primitive is three float4 (48 B): {p1.xyz, e1.x} {e1.yz, e2.xy} {e2.z, type, index, -}
// with e1 = p2 - p1, e2 = p3 - p1 precomputed (the same fp32 subtraction the CPU does).
#pragma once
#include "skc_common.cuh"

struct DRay { f3 o, d; float mint, maxt; };

struct BvhView {
    const float4 *nodes;
    const float4 *prims;
    const skc_light *spheres;
    const float4 *wide;        // 4-wide form of the same tree for any-hit rays (8 float4 per node, see bvhOccludedWide), or nullptr
    // alpha-tested triangles (shapes/trianglemesh.inl:117-126,174-207): the mesh's "texture alpha" as level-0 float texels with
    // its UVMapping2D, and the three vertices' (u, v) of every BVH primitive; alphaTex == nullptr: no alpha texture
    const float *alphaTex = nullptr, *alphaUV = nullptr;
    int alphaW = 0, alphaH = 0;
    float alphaSu = 1.f, alphaSv = 1.f, alphaDu = 0.f, alphaDv = 0.f;
};

// What closest-hit shading needs of the triangle mesh (shapes/trianglemesh.inl:224-299).
struct MeshView {
    const int4 *primInfo;      // per BVH primitive: vertex ids v0,v1,v2 and type
    const float *P, *N, *UV;   // world-space points, object-space normals, uv (N/UV may be null)
    float o2w[16], w2o[16];    // mesh ObjectToWorld matrix and inverse
    int flip;                  // ReverseOrientation ^ TransformSwapsHandedness
};

// Triangle::Intersect / IntersectP's alpha test: the texture at the hit's interpolated (u, v) -- UVMapping2D, then
// MIPMap<float>::Lookup with zero differentials = triangle(0, s, t) (textures/imagemap.cpp:92-101, core/mipmap.h:253-264,
// TEXTURE_REPEAT) -- and the hit is ignored where it evaluates to exactly 0.
__device__ __forceinline__ int alphaWrap(int a, int b) { int n = a / b; a -= n * b; if (a < 0) a += b; return a; }
__device__ inline bool alphaOpaque(const BvhView &bv, uint32_t prim, float b1, float b2) {
    const float *uv = bv.alphaUV + 6 * (size_t)prim;
    const float b0 = 1 - b1 - b2;
    const float tu = b0 * __ldg(uv) + b1 * __ldg(uv + 2) + b2 * __ldg(uv + 4);
    const float tv = b0 * __ldg(uv + 1) + b1 * __ldg(uv + 3) + b2 * __ldg(uv + 5);
    float s = bv.alphaSu * tu + bv.alphaDu, t = bv.alphaSv * tv + bv.alphaDv;
    s = s * (float)bv.alphaW - 0.5f;
    t = t * (float)bv.alphaH - 0.5f;
    const int s0 = (int)floorf(s), t0 = (int)floorf(t);
    const float ds = s - (float)s0, dt = t - (float)t0;
    const int sa = alphaWrap(s0, bv.alphaW), sb = alphaWrap(s0 + 1, bv.alphaW), ta = alphaWrap(t0, bv.alphaH), tb = alphaWrap(t0 + 1, bv.alphaH);
    const float v = (1.f - ds) * (1.f - dt) * __ldg(&bv.alphaTex[(size_t)ta * bv.alphaW + sa]) + (1.f - ds) * dt * __ldg(&bv.alphaTex[(size_t)tb * bv.alphaW + sa]) +
                    ds * (1.f - dt) * __ldg(&bv.alphaTex[(size_t)ta * bv.alphaW + sb]) + ds * dt * __ldg(&bv.alphaTex[(size_t)tb * bv.alphaW + sb]);
    return v != 0.f;
}

__device__ __forceinline__ f3 xformPoint(const float *m, f3 p) {
    float xp = m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3];
    float yp = m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7];
    float zp = m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11];
    float wp = m[12] * p.x + m[13] * p.y + m[14] * p.z + m[15];
    if (wp != 1.f) { float inv = 1.f / wp; xp *= inv; yp *= inv; zp *= inv; }
    return mk3(xp, yp, zp);
}
__device__ __forceinline__ f3 xformVector(const float *m, f3 v) {
    return mk3(m[0] * v.x + m[1] * v.y + m[2] * v.z,
               m[4] * v.x + m[5] * v.y + m[6] * v.z,
               m[8] * v.x + m[9] * v.y + m[10] * v.z);
}
// Transform::operator()(Normal): transpose of the inverse (core/transform.h:232-237)
__host__ __device__ __forceinline__ f3 xformNormal(const float *mInv, f3 n) {
    return mk3(mInv[0] * n.x + mInv[4] * n.y + mInv[8] * n.z,
               mInv[1] * n.x + mInv[5] * n.y + mInv[9] * n.z,
               mInv[2] * n.x + mInv[6] * n.y + mInv[10] * n.z);
}

// core/pbrt.h:354-368
__device__ __forceinline__ bool quadratic(float A, float B, float C, float *t0, float *t1) {
    float discrim = B * B - 4.f * A * C;
    if (discrim < 0.f) return false;
    float rootDiscrim = sqrtf(discrim);
    float q;
    if (B < 0) q = -.5f * (B - rootDiscrim);
    else       q = -.5f * (B + rootDiscrim);
    *t0 = q / A;
    *t1 = C / q;
    if (*t0 > *t1) { float t = *t0; *t0 = *t1; *t1 = t; }
    return true;
}

// shapes/sphere.cpp:58-101 / :160-210 -- returns true and thit (object == world parameter)
__device__ inline bool sphereIntersect(const skc_light &s, const DRay &r, float *tHit) {
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
    if (t0 < ray.mint) {
        thit = t1;
        if (thit > ray.maxt) return false;
    }
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
        if ((zmin > -radius && phit.z < zmin) || (zmax < radius && phit.z > zmax) || phi > phiMax)
            return false;
    }
    *tHit = thit;
    return true;
}

// shapes/trianglemesh.inl:55-90 (no alpha texture on this path).  b1/b2 returned for shading.
__device__ __forceinline__ bool triIntersect(f3 p1, f3 e1, f3 e2, const DRay &ray,
                                             float *tOut, float *b1Out, float *b2Out) {
    f3 s1 = cross3(ray.d, e2);
    float divisor = dot3(s1, e1);
    if (divisor == 0.f) return false;
    float invDivisor = 1.f / divisor;
    f3 d = sub3(ray.o, p1);
    float b1 = dot3(d, s1) * invDivisor;
    if (b1 < 0.f || b1 > 1.f) return false;
    f3 s2 = cross3(d, e1);
    float b2 = dot3(ray.d, s2) * invDivisor;
    if (b2 < 0.f || b1 + b2 > 1.f) return false;
    float t = dot3(e2, s2) * invDivisor;
    if (t < ray.mint || t > ray.maxt) return false;
    *tOut = t; *b1Out = b1; *b2Out = b2;
    return true;
}

// core/geometry.h:564-574
__host__ __device__ inline void coordinateSystem(f3 v1, f3 *v2, f3 *v3) {
    if (fabsf(v1.x) > fabsf(v1.y)) {
        float invLen = 1.f / sqrtf(v1.x * v1.x + v1.z * v1.z);
        *v2 = mk3(-v1.z * invLen, 0.f, v1.x * invLen);
    } else {
        float invLen = 1.f / sqrtf(v1.y * v1.y + v1.z * v1.z);
        *v2 = mk3(0.f, v1.z * invLen, -v1.y * invLen);
    }
    *v3 = cross3(v1, *v2);
}

__host__ __device__ __forceinline__ f3 ld3(const float *a, int i) { return mk3(a[3 * i], a[3 * i + 1], a[3 * i + 2]); }

// Shading point/normal of a triangle hit: trianglemesh.inl:92-133 (dg), :224-299 (shading
// geometry), core/diffgeom.cpp:40-55 (nn and the orientation flip).
// Shading normal and uv at barycentric (b0, b1, b2) of the triangle with vertex ids vi.xyz: GetDifferentialGeometries
// (shapes/trianglemesh.inl:303-345) + GetShadingGeometry (:224-299) + diffgeom.cpp:40-55
// ngOut / ssOut (optional): the geometric normal dgGeom.nn and the shading dpdu that BSDF::BSDF turns into its local frame
// (core/reflection.cpp:755-762)
__host__ __device__ inline void triShadingGeometry(const MeshView &mv, int4 vi, float b0, float b1, float b2, f3 *nOut, float *tuOut, float *tvOut,
                                          f3 *ngOut = nullptr, f3 *ssOut = nullptr) {
    f3 p1 = ld3(mv.P, vi.x), p2 = ld3(mv.P, vi.y), p3 = ld3(mv.P, vi.z);
    f3 e1 = sub3(p2, p1), e2 = sub3(p3, p1);
    float uvs[3][2];
    if (mv.UV) {
        uvs[0][0] = mv.UV[2 * vi.x]; uvs[0][1] = mv.UV[2 * vi.x + 1];
        uvs[1][0] = mv.UV[2 * vi.y]; uvs[1][1] = mv.UV[2 * vi.y + 1];
        uvs[2][0] = mv.UV[2 * vi.z]; uvs[2][1] = mv.UV[2 * vi.z + 1];
    } else {
        uvs[0][0] = 0.f; uvs[0][1] = 0.f; uvs[1][0] = 1.f; uvs[1][1] = 0.f; uvs[2][0] = 1.f; uvs[2][1] = 1.f;
    }
    float du1 = uvs[0][0] - uvs[2][0], du2 = uvs[1][0] - uvs[2][0];
    float dv1 = uvs[0][1] - uvs[2][1], dv2 = uvs[1][1] - uvs[2][1];
    f3 dp1 = sub3(p1, p3), dp2 = sub3(p2, p3);
    f3 dpdu, dpdv;
    float determinant = du1 * dv2 - dv1 * du2;
    if (determinant == 0.f) {
        coordinateSystem(normalize3(cross3(e2, e1)), &dpdu, &dpdv);
    } else {
        float invdet = 1.f / determinant;
        dpdu = mul3(sub3(mul3(dp1, dv2), mul3(dp2, dv1)), invdet);
        dpdv = mul3(add3(mul3(dp1, -du2), mul3(dp2, du1)), invdet);
    }
    float tu = b0 * uvs[0][0] + b1 * uvs[1][0] + b2 * uvs[2][0];
    float tv = b0 * uvs[0][1] + b1 * uvs[1][1] + b2 * uvs[2][1];
    *tuOut = tu; *tvOut = tv;
    f3 nn;
    if (ngOut) {
        f3 ng = normalize3(cross3(dpdu, dpdv));
        *ngOut = mv.flip ? mul3(ng, -1.f) : ng;
    }
    if (!mv.N) {
        nn = normalize3(cross3(dpdu, dpdv));
        if (ssOut) *ssOut = dpdu;
    } else {
        float A00 = uvs[1][0] - uvs[0][0], A01 = uvs[2][0] - uvs[0][0];
        float A10 = uvs[1][1] - uvs[0][1], A11 = uvs[2][1] - uvs[0][1];
        float C0 = tu - uvs[0][0], C1 = tv - uvs[0][1];
        float bb[3];
        float det = A00 * A11 - A01 * A10;     // core/transform.cpp:39-49
        bool ok = !(fabsf(det) < 1e-10f);
        if (ok) {
            bb[1] = (A11 * C0 - A01 * C1) / det;
            bb[2] = (A00 * C1 - A10 * C0) / det;
            if (isnan(bb[1]) || isnan(bb[2])) ok = false;
        }
        if (!ok) bb[0] = bb[1] = bb[2] = 1.f / 3.f;
        else bb[0] = 1.f - bb[1] - bb[2];
        f3 n0 = ld3(mv.N, vi.x), n1 = ld3(mv.N, vi.y), n2 = ld3(mv.N, vi.z);
        f3 nsum = add3(add3(mul3(n0, bb[0]), mul3(n1, bb[1])), mul3(n2, bb[2]));
        f3 ns = normalize3(xformNormal(mv.w2o, nsum));
        f3 ss = normalize3(dpdu);
        f3 ts = cross3(ss, ns);
        if (ts.x * ts.x + ts.y * ts.y + ts.z * ts.z > 0.f) {
            ts = normalize3(ts);
            ss = cross3(ts, ns);
        } else {
            coordinateSystem(ns, &ss, &ts);
        }
        nn = normalize3(cross3(ss, ts));
        if (ssOut) *ssOut = ss;
    }
    if (mv.flip) nn = mul3(nn, -1.f);
    *nOut = nn;
}


// The hit's DifferentialGeometry as Triangle::Intersect builds it (shapes/trianglemesh.inl:80-135): dpdu / dpdv from the uv parameterisation,
// the interpolated (u, v), and nn = Normalize(Cross(dpdu, dpdv)), flipped with the mesh's orientation (core/diffgeom.cpp:36-55)
__host__ __device__ inline void triDiffGeom(const MeshView &mv, int4 vi, float b0, float b1, float b2, f3 *dpduOut, f3 *dpdvOut, float *tuOut, float *tvOut, f3 *nnOut) {
    f3 p1 = ld3(mv.P, vi.x), p2 = ld3(mv.P, vi.y), p3 = ld3(mv.P, vi.z);
    f3 e1 = sub3(p2, p1), e2 = sub3(p3, p1);
    float uvs[3][2];
    if (mv.UV) {
        uvs[0][0] = mv.UV[2 * vi.x]; uvs[0][1] = mv.UV[2 * vi.x + 1];
        uvs[1][0] = mv.UV[2 * vi.y]; uvs[1][1] = mv.UV[2 * vi.y + 1];
        uvs[2][0] = mv.UV[2 * vi.z]; uvs[2][1] = mv.UV[2 * vi.z + 1];
    } else {
        uvs[0][0] = 0.f; uvs[0][1] = 0.f; uvs[1][0] = 1.f; uvs[1][1] = 0.f; uvs[2][0] = 1.f; uvs[2][1] = 1.f;
    }
    float du1 = uvs[0][0] - uvs[2][0], du2 = uvs[1][0] - uvs[2][0];
    float dv1 = uvs[0][1] - uvs[2][1], dv2 = uvs[1][1] - uvs[2][1];
    f3 dp1 = sub3(p1, p3), dp2 = sub3(p2, p3);
    f3 dpdu, dpdv;
    float determinant = du1 * dv2 - dv1 * du2;
    if (determinant == 0.f) {
        coordinateSystem(normalize3(cross3(e2, e1)), &dpdu, &dpdv);
    } else {
        float invdet = 1.f / determinant;
        dpdu = mul3(sub3(mul3(dp1, dv2), mul3(dp2, dv1)), invdet);
        dpdv = mul3(add3(mul3(dp1, -du2), mul3(dp2, du1)), invdet);
    }
    *tuOut = b0 * uvs[0][0] + b1 * uvs[1][0] + b2 * uvs[2][0];
    *tvOut = b0 * uvs[0][1] + b1 * uvs[1][1] + b2 * uvs[2][1];
    *dpduOut = dpdu; *dpdvOut = dpdv;
    f3 ng = normalize3(cross3(dpdu, dpdv));
    *nnOut = mv.flip ? mul3(ng, -1.f) : ng;
}

struct SlabSetup { f3 invDir; int neg[3]; int negMask; };   // negMask: the three signs as bits, for the child order by split axis (no dynamically indexed array -> no local memory)
__device__ __forceinline__ SlabSetup slabSetup(const DRay &r) {
    SlabSetup s;
    s.invDir = mk3(1.f / r.d.x, 1.f / r.d.y, 1.f / r.d.z);
    s.neg[0] = s.invDir.x < 0; s.neg[1] = s.invDir.y < 0; s.neg[2] = s.invDir.z < 0;
    s.negMask = s.neg[0] | (s.neg[1] << 1) | (s.neg[2] << 2);
    return s;
}
// accelerators/bvh.cpp:126-148
__device__ __forceinline__ bool slabTest(float4 n0, float4 n1, const DRay &ray, const SlabSetup &s) {
    float minx = n0.x, miny = n0.y, minz = n0.z, maxx = n0.w, maxy = n1.x, maxz = n1.y;
    float tmin = ((s.neg[0] ? maxx : minx) - ray.o.x) * s.invDir.x;
    float tmax = ((s.neg[0] ? minx : maxx) - ray.o.x) * s.invDir.x;
    float tymin = ((s.neg[1] ? maxy : miny) - ray.o.y) * s.invDir.y;
    float tymax = ((s.neg[1] ? miny : maxy) - ray.o.y) * s.invDir.y;
    if ((tmin > tymax) || (tymin > tmax)) return false;
    if (tymin > tmin) tmin = tymin;
    if (tymax < tmax) tmax = tymax;
    float tzmin = ((s.neg[2] ? maxz : minz) - ray.o.z) * s.invDir.z;
    float tzmax = ((s.neg[2] ? minz : maxz) - ray.o.z) * s.invDir.z;
    if ((tmin > tzmax) || (tzmin > tmax)) return false;
    if (tzmin > tmin) tmin = tzmin;
    if (tzmax < tmax) tmax = tzmax;
    return (tmin < ray.maxt) && (tmax > ray.mint);
}

#define SKC_BVH_STACK 64

// BVHAccel::IntersectP (bvh.cpp:442-488).  counters (optional): [0] slab tests, [1] prim tests.
template <bool COUNT>
__device__ inline bool bvhOccluded(const BvhView &bv, const DRay &ray, unsigned long long *cnt) {
    SlabSetup ss = slabSetup(ray);
    uint32_t todo[SKC_BVH_STACK];
    uint32_t todoOffset = 0, nodeNum = 0;
    uint32_t nNode = 0, nPrim = 0;
    bool occluded = false;
    while (true) {
        float4 n0 = __ldg(&bv.nodes[2 * nodeNum]);
        float4 n1 = __ldg(&bv.nodes[2 * nodeNum + 1]);
        if (COUNT) nNode++;
        if (slabTest(n0, n1, ray, ss)) {
            uint32_t offset = __float_as_uint(n1.z), meta = __float_as_uint(n1.w);
            uint32_t nPrims = meta & 0xff;
            if (nPrims > 0) {
                for (uint32_t i = 0; i < nPrims; ++i) {
                    const float4 *pp = &bv.prims[3 * (offset + i)];
                    float4 a = __ldg(pp), b = __ldg(pp + 1), c = __ldg(pp + 2);
                    if (COUNT) nPrim++;
                    float t, b1, b2;
                    bool h;
                    if (__float_as_int(c.y) == 0) {
                        h = triIntersect(mk3(a.x, a.y, a.z), mk3(a.w, b.x, b.y), mk3(b.z, b.w, c.x), ray, &t, &b1, &b2);
                        if (h && bv.alphaTex) h = alphaOpaque(bv, offset + i, b1, b2);
                    } else
                        h = sphereIntersect(bv.spheres[__float_as_int(c.z)], ray, &t);
                    if (h) { occluded = true; break; }
                }
                if (occluded) break;
                if (todoOffset == 0) break;
                nodeNum = todo[--todoOffset];
            } else {
                uint32_t axis = (meta >> 8) & 3;
                if ((ss.negMask >> axis) & 1) { todo[todoOffset++] = nodeNum + 1; nodeNum = offset; }
                else              { todo[todoOffset++] = offset; nodeNum = nodeNum + 1; }
            }
        } else {
            if (todoOffset == 0) break;
            nodeNum = todo[--todoOffset];
        }
    }
    if (COUNT && cnt) { atomicAdd(&cnt[0], (unsigned long long)nNode); atomicAdd(&cnt[1], (unsigned long long)nPrim); }
    return occluded;
}

// ---- any-hit over the 4-wide form of the same tree --------------------------------------------------------------------
// IntersectP's ANSWER does not depend on the order in which nodes are visited, only on which leaves are reached: a leaf is
// reached iff the ray passes the slab test of the leaf's box and of every ancestor's box.  A child's box lies inside its
// parent's, and the slab test is monotonic in the box (a larger box gives a wider [tmin, tmax] in the same fp32 operations), so
// a ray that passes a node's box passes all of its ancestors' boxes: testing the boxes of SOME ancestors only -- every second
// level -- reaches exactly the same leaves.  So the binary tree is collapsed at upload time (skc_bvh.cu buildWide): a wide node
// holds the boxes of up to four descendants of one binary node (its children, with the larger interior ones replaced by their
// own children), as one 128 B line:
//   q0..q5 = min.x[4] min.y[4] min.z[4] max.x[4] max.y[4] max.z[4]      q6 = child[4] (uint bits)      q7 = unused
//   child: 0xffffffff empty | 0x80000000 + (nPrims << 24) + first primitive (leaf) | index of a wide node
// The four slab tests of a node are independent (the binary walk is one dependent fetch-test-branch chain per box), leaves are
// tested as they are met, interior hits go to a short stack.  The slab arithmetic is the reference's (bvh.cpp:126-148), the
// boxes are the reference's, the primitive tests are the same functions: hit / no-hit is identical to bvhOccluded for every
// ray (tests/test_gpu_parity.py::test_wide_any_hit_equals_binary).  Visit counts differ, so the counting instances stay binary.
#define SKC_WIDE_EMPTY 0xffffffffu
#define SKC_WIDE_LEAF 0x80000000u
__device__ __forceinline__ bool slabTestSoA(float minx, float miny, float minz, float maxx, float maxy, float maxz, const DRay &ray, const SlabSetup &s) {
    float tmin = ((s.neg[0] ? maxx : minx) - ray.o.x) * s.invDir.x;
    float tmax = ((s.neg[0] ? minx : maxx) - ray.o.x) * s.invDir.x;
    float tymin = ((s.neg[1] ? maxy : miny) - ray.o.y) * s.invDir.y;
    float tymax = ((s.neg[1] ? miny : maxy) - ray.o.y) * s.invDir.y;
    if ((tmin > tymax) || (tymin > tmax)) return false;
    if (tymin > tmin) tmin = tymin;
    if (tymax < tmax) tmax = tymax;
    float tzmin = ((s.neg[2] ? maxz : minz) - ray.o.z) * s.invDir.z;
    float tzmax = ((s.neg[2] ? minz : maxz) - ray.o.z) * s.invDir.z;
    if ((tmin > tzmax) || (tzmin > tmax)) return false;
    if (tzmin > tmin) tmin = tzmin;
    if (tzmax < tmax) tmax = tzmax;
    return (tmin < ray.maxt) && (tmax > ray.mint);
}
__device__ __forceinline__ bool leafOccluded(const BvhView &bv, uint32_t first, uint32_t count, const DRay &ray) {
    for (uint32_t i = 0; i < count; ++i) {
        const float4 *pp = &bv.prims[3 * (first + i)];
        float4 a = __ldg(pp), b = __ldg(pp + 1), c = __ldg(pp + 2);
        float t, b1, b2;
        bool h;
        if (__float_as_int(c.y) == 0) {
            h = triIntersect(mk3(a.x, a.y, a.z), mk3(a.w, b.x, b.y), mk3(b.z, b.w, c.x), ray, &t, &b1, &b2);
            if (h && bv.alphaTex) h = alphaOpaque(bv, first + i, b1, b2);
        } else h = sphereIntersect(bv.spheres[__float_as_int(c.z)], ray, &t);
        if (h) return true;
    }
    return false;
}
#define SKC_WIDE_STACK 48
__device__ inline bool bvhOccludedWide(const BvhView &bv, const DRay &ray) {
    const SlabSetup ss = slabSetup(ray);
    uint32_t todo[SKC_WIDE_STACK];
    int top = 0;
    uint32_t node = 0;
    for (;;) {
        const float4 *q = bv.wide + 8 * (size_t)node;
        const float4 mnx = __ldg(q), mny = __ldg(q + 1), mnz = __ldg(q + 2), mxx = __ldg(q + 3), mxy = __ldg(q + 4), mxz = __ldg(q + 5), ch = __ldg(q + 6);
        const uint32_t c[4] = {__float_as_uint(ch.x), __float_as_uint(ch.y), __float_as_uint(ch.z), __float_as_uint(ch.w)};
        const bool h[4] = {c[0] != SKC_WIDE_EMPTY && slabTestSoA(mnx.x, mny.x, mnz.x, mxx.x, mxy.x, mxz.x, ray, ss),
                           c[1] != SKC_WIDE_EMPTY && slabTestSoA(mnx.y, mny.y, mnz.y, mxx.y, mxy.y, mxz.y, ray, ss),
                           c[2] != SKC_WIDE_EMPTY && slabTestSoA(mnx.z, mny.z, mnz.z, mxx.z, mxy.z, mxz.z, ray, ss),
                           c[3] != SKC_WIDE_EMPTY && slabTestSoA(mnx.w, mny.w, mnz.w, mxx.w, mxy.w, mxz.w, ray, ss)};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (!h[k]) continue;
            if (c[k] & SKC_WIDE_LEAF) { if (leafOccluded(bv, c[k] & 0x00ffffffu, (c[k] >> 24) & 0x7fu, ray)) return true; }
            else todo[top++] = c[k];
        }
        if (top == 0) return false;
        node = todo[--top];
    }
}
__device__ __forceinline__ bool bvhOccludedFast(const BvhView &bv, const DRay &ray) {
    return bv.wide ? bvhOccludedWide(bv, ray) : bvhOccluded<false>(bv, ray, nullptr);
}

// BVHAccel::Intersect (bvh.cpp:388-439): closest hit; ray.maxt shrinks (primitive.cpp:191).
template <bool COUNT>
__device__ inline int bvhClosest(const BvhView &bv, DRay &ray, float *b1Hit, float *b2Hit,
                                 unsigned long long *cnt) {
    SlabSetup ss = slabSetup(ray);
    uint32_t todo[SKC_BVH_STACK];
    uint32_t todoOffset = 0, nodeNum = 0;
    uint32_t nNode = 0, nPrim = 0;
    int hitPrim = -1;
    while (true) {
        float4 n0 = __ldg(&bv.nodes[2 * nodeNum]);
        float4 n1 = __ldg(&bv.nodes[2 * nodeNum + 1]);
        if (COUNT) nNode++;
        if (slabTest(n0, n1, ray, ss)) {
            uint32_t offset = __float_as_uint(n1.z), meta = __float_as_uint(n1.w);
            uint32_t nPrims = meta & 0xff;
            if (nPrims > 0) {
                for (uint32_t i = 0; i < nPrims; ++i) {
                    const float4 *pp = &bv.prims[3 * (offset + i)];
                    float4 a = __ldg(pp), b = __ldg(pp + 1), c = __ldg(pp + 2);
                    if (COUNT) nPrim++;
                    float t, b1 = 0.f, b2 = 0.f;
                    bool h;
                    if (__float_as_int(c.y) == 0) {
                        h = triIntersect(mk3(a.x, a.y, a.z), mk3(a.w, b.x, b.y), mk3(b.z, b.w, c.x), ray, &t, &b1, &b2);
                        if (h && bv.alphaTex) h = alphaOpaque(bv, offset + i, b1, b2);
                    } else
                        h = sphereIntersect(bv.spheres[__float_as_int(c.z)], ray, &t);
                    if (h) { ray.maxt = t; hitPrim = (int)(offset + i); *b1Hit = b1; *b2Hit = b2; }
                }
                if (todoOffset == 0) break;
                nodeNum = todo[--todoOffset];
            } else {
                uint32_t axis = (meta >> 8) & 3;
                if ((ss.negMask >> axis) & 1) { todo[todoOffset++] = nodeNum + 1; nodeNum = offset; }
                else              { todo[todoOffset++] = offset; nodeNum = nodeNum + 1; }
            }
        } else {
            if (todoOffset == 0) break;
            nodeNum = todo[--todoOffset];
        }
    }
    if (COUNT && cnt) { atomicAdd(&cnt[0], (unsigned long long)nNode); atomicAdd(&cnt[1], (unsigned long long)nPrim); }
    return hitPrim;
}

/* skin_oracle.c -- CPU restatement of pbrt-v2-skin's subsurface hot path, in plain C.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (pbrt-v2-skin_b200/, bench.py's own
 * arm) links, imports or executes this file; only tests/, __graft_entry__.smoke() and the
 * checker legs use it.  It restates the reference algorithm function by function (each cites
 * the reference file:line it follows) with the reference's own control flow -- recursive
 * pointer octree, recursive Mo with nested sums, scalar BVH loop -- so it is an independent
 * check of the CUDA path's re-designed data layout and traversal order.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks every function below against outputs of
 * the real reference built here from /root/reference (oracle/_ref, golden fixtures in
 * tests/golden/ (.npz)): scrambles+irradiance, BVH hits, octree root, Mo and Li bit-exact or
 * within fp32 round-off as stated per test.
 *
 * Build: gcc -O2 -ffp-contract=off -fPIC -shared (oracle/Makefile).  -ffp-contract=off keeps
 * the reference's x86-64 SSE2 arithmetic (no FMA contraction).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* core/pbrt.h:193-196: the reference redefines M_PI as a float literal */
#undef M_PI
#define M_PI 3.14159265358979323846f
#define NS 30
#define RHO_LEN 1025

typedef struct { float x, y, z; } v3;
static v3 V(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static v3 sub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static v3 add(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static v3 mul(v3 a, float s) { return V(a.x * s, a.y * s, a.z * s); }
static float dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
/* core/geometry.h:529-536 */
static v3 cross(v3 a, v3 b) {
    double ax = a.x, ay = a.y, az = a.z, bx = b.x, by = b.y, bz = b.z;
    return V((float)((ay * bz) - (az * by)), (float)((az * bx) - (ax * bz)), (float)((ax * by) - (ay * bx)));
}
static float len(v3 a) { return sqrtf(a.x * a.x + a.y * a.y + a.z * a.z); }
static v3 normalize(v3 a) { float inv = 1.f / len(a); return V(a.x * inv, a.y * inv, a.z * inv); }  /* geometry.h:102-106,562 */
static float dist2(v3 a, v3 b) { v3 d = sub(a, b); return d.x * d.x + d.y * d.y + d.z * d.z; }
static void coordinate_system(v3 v1, v3 *v2, v3 *v3o) {   /* geometry.h:564-574 */
    if (fabsf(v1.x) > fabsf(v1.y)) {
        float invLen = 1.f / sqrtf(v1.x * v1.x + v1.z * v1.z);
        *v2 = V(-v1.z * invLen, 0.f, v1.x * invLen);
    } else {
        float invLen = 1.f / sqrtf(v1.y * v1.y + v1.z * v1.z);
        *v2 = V(0.f, v1.z * invLen, -v1.y * invLen);
    }
    *v3o = cross(v1, *v2);
}

/* ------------------------------------------------------------------ MT19937, core/rng.cpp:43-110 */
typedef struct { uint32_t mt[624]; int mti; } rng_t;
static void rng_seed(rng_t *r, uint32_t seed) {
    r->mt[0] = seed;
    for (r->mti = 1; r->mti < 624; r->mti++)
        r->mt[r->mti] = 1812433253u * (r->mt[r->mti - 1] ^ (r->mt[r->mti - 1] >> 30)) + (uint32_t)r->mti;
}
static uint32_t rng_uint(rng_t *r) {
    static const uint32_t mag01[2] = {0u, 0x9908b0dfu};
    uint32_t y;
    if (r->mti >= 624) {
        int kk;
        for (kk = 0; kk < 624 - 397; kk++) { y = (r->mt[kk] & 0x80000000u) | (r->mt[kk + 1] & 0x7fffffffu); r->mt[kk] = r->mt[kk + 397] ^ (y >> 1) ^ mag01[y & 1u]; }
        for (; kk < 623; kk++) { y = (r->mt[kk] & 0x80000000u) | (r->mt[kk + 1] & 0x7fffffffu); r->mt[kk] = r->mt[kk + (397 - 624)] ^ (y >> 1) ^ mag01[y & 1u]; }
        y = (r->mt[623] & 0x80000000u) | (r->mt[0] & 0x7fffffffu);
        r->mt[623] = r->mt[396] ^ (y >> 1) ^ mag01[y & 1u];
        r->mti = 0;
    }
    y = r->mt[r->mti++];
    y ^= (y >> 11); y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= (y >> 18);
    return y;
}
/* core/montecarlo.h:285-301 */
static float van_der_corput(uint32_t n, uint32_t scramble) {
    n = (n << 16) | (n >> 16);
    n = ((n & 0x00ff00ff) << 8) | ((n & 0xff00ff00) >> 8);
    n = ((n & 0x0f0f0f0f) << 4) | ((n & 0xf0f0f0f0) >> 4);
    n = ((n & 0x33333333) << 2) | ((n & 0xcccccccc) >> 2);
    n = ((n & 0x55555555) << 1) | ((n & 0xaaaaaaaa) >> 1);
    n ^= scramble;
    float v = ((n >> 8) & 0xffffff) / (float)(1 << 24);
    return v < 0.99999994f ? v : 0.99999994f;
}
static float sobol2(uint32_t n, uint32_t scramble) {
    for (uint32_t v = 1u << 31; n != 0; n >>= 1, v ^= v >> 1)
        if (n & 0x1) scramble ^= v;
    float v = ((scramble >> 8) & 0xffffff) / (float)(1 << 24);
    return v < 0.99999994f ? v : 0.99999994f;
}

/* ------------------------------------------------------------------ records shared with include/skin_cuda.h */
typedef struct { float pmin[3], pmax[3]; uint32_t offset; uint8_t nPrimitives, axis, pad[2]; } bvh_node;
typedef struct { int32_t type, mesh, index, primitiveId; } prim_t;
typedef struct { float o2w[16], w2o[16], radius, zmin, zmax, phiMax, nSamples, reverse, thetaMin, thetaMax, Lemit[NS]; } light_t;
typedef struct { float o[3], d[3], mint, maxt; } ray_t;
typedef struct { int32_t hit, prim; float t, rayEpsilon, p[3], n[3], wo[3]; int32_t hasBssrdf; } hit_t;
typedef struct { float p[3], n[3], u, v; uint32_t materialId; float area, rayEpsilon; } spoint_t;
typedef struct { float p[3], n[3], E[NS], area, rayEpsilon; } ipoint_t;

typedef struct {
    const bvh_node *nodes; const prim_t *prims; const float *P; const float *N; const float *UV; const int32_t *idx;
    const light_t *spheres; const float *o2w; /* mesh m[16], mInv[16] */ int flip;
} scene_t;

static v3 xf_point(const float *m, v3 p) {   /* core/transform.h:192-212 */
    float xp = m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3];
    float yp = m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7];
    float zp = m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11];
    float wp = m[12] * p.x + m[13] * p.y + m[14] * p.z + m[15];
    if (wp == 1.) return V(xp, yp, zp);
    float inv = 1.f / wp;
    return V(xp * inv, yp * inv, zp * inv);
}
static v3 xf_vector(const float *m, v3 v) {
    return V(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[4] * v.x + m[5] * v.y + m[6] * v.z, m[8] * v.x + m[9] * v.y + m[10] * v.z);
}
static v3 xf_normal(const float *mi, v3 n) {   /* transform.h:232-237 */
    return V(mi[0] * n.x + mi[4] * n.y + mi[8] * n.z, mi[1] * n.x + mi[5] * n.y + mi[9] * n.z, mi[2] * n.x + mi[6] * n.y + mi[10] * n.z);
}

/* core/pbrt.h:354-368 */
static int quadratic(float A, float B, float C, float *t0, float *t1) {
    float discrim = B * B - 4.f * A * C;
    if (discrim < 0.) return 0;
    float rootDiscrim = sqrtf(discrim);
    float q;
    if (B < 0) q = -.5f * (B - rootDiscrim); else q = -.5f * (B + rootDiscrim);
    *t0 = q / A; *t1 = C / q;
    if (*t0 > *t1) { float t = *t0; *t0 = *t1; *t1 = t; }
    return 1;
}

/* shapes/sphere.cpp:58-158 (Intersect) / :160-210 (IntersectP); nn = dg.nn when asked */
static int sphere_intersect(const light_t *s, v3 ro, v3 rd, float mint, float maxt, float *tHit, v3 *nn) {
    v3 o = xf_point(s->w2o, ro), d = xf_vector(s->w2o, rd);
    float radius = s->radius;
    float A = d.x * d.x + d.y * d.y + d.z * d.z;
    float B = 2 * (d.x * o.x + d.y * o.y + d.z * o.z);
    float C = o.x * o.x + o.y * o.y + o.z * o.z - radius * radius;
    float t0, t1;
    if (!quadratic(A, B, C, &t0, &t1)) return 0;
    if (t0 > maxt || t1 < mint) return 0;
    float thit = t0;
    if (t0 < mint) { thit = t1; if (thit > maxt) return 0; }
    v3 phit = add(o, mul(d, thit));
    if (phit.x == 0.f && phit.y == 0.f) phit.x = 1e-5f * radius;
    float phi = atan2f(phit.y, phit.x);
    if (phi < 0.) phi += 2.f * M_PI;
    if ((s->zmin > -radius && phit.z < s->zmin) || (s->zmax < radius && phit.z > s->zmax) || phi > s->phiMax) {
        if (thit == t1) return 0;
        if (t1 > maxt) return 0;
        thit = t1;
        phit = add(o, mul(d, thit));
        if (phit.x == 0.f && phit.y == 0.f) phit.x = 1e-5f * radius;
        phi = atan2f(phit.y, phit.x);
        if (phi < 0.) phi += 2.f * M_PI;
        if ((s->zmin > -radius && phit.z < s->zmin) || (s->zmax < radius && phit.z > s->zmax) || phi > s->phiMax) return 0;
    }
    if (nn) {   /* sphere.cpp:103-150, core/diffgeom.cpp:40-55 */
        float theta = acosf(fminf(fmaxf(phit.z / radius, -1.f), 1.f));
        float zradius = sqrtf(phit.x * phit.x + phit.y * phit.y);
        float invzradius = 1.f / zradius;
        float cosphi = phit.x * invzradius, sinphi = phit.y * invzradius;
        v3 dpdu = V(-s->phiMax * phit.y, s->phiMax * phit.x, 0);
        v3 dpdv = mul(V(phit.z * cosphi, phit.z * sinphi, -radius * sinf(theta)), s->thetaMax - s->thetaMin);
        *nn = normalize(cross(xf_vector(s->o2w, dpdu), xf_vector(s->o2w, dpdv)));
        if (s->reverse != 0.f) *nn = mul(*nn, -1.f);
    }
    *tHit = thit;
    return 1;
}

/* shapes/trianglemesh.inl:55-90 / :140-168 */
static int tri_intersect(const scene_t *sc, int tri, v3 ro, v3 rd, float mint, float maxt, float *tOut, float *b1o, float *b2o) {
    const int32_t *v = sc->idx + 3 * tri;
    v3 p1 = V(sc->P[3 * v[0]], sc->P[3 * v[0] + 1], sc->P[3 * v[0] + 2]);
    v3 p2 = V(sc->P[3 * v[1]], sc->P[3 * v[1] + 1], sc->P[3 * v[1] + 2]);
    v3 p3 = V(sc->P[3 * v[2]], sc->P[3 * v[2] + 1], sc->P[3 * v[2] + 2]);
    v3 e1 = sub(p2, p1), e2 = sub(p3, p1);
    v3 s1 = cross(rd, e2);
    float divisor = dot(s1, e1);
    if (divisor == 0.) return 0;
    float invDivisor = 1.f / divisor;
    v3 d = sub(ro, p1);
    float b1 = dot(d, s1) * invDivisor;
    if (b1 < 0. || b1 > 1.) return 0;
    v3 s2 = cross(d, e1);
    float b2 = dot(rd, s2) * invDivisor;
    if (b2 < 0. || b1 + b2 > 1.) return 0;
    float t = dot(e2, s2) * invDivisor;
    if (t < mint || t > maxt) return 0;
    *tOut = t; *b1o = b1; *b2o = b2;
    return 1;
}

/* accelerators/bvh.cpp:126-148 */
static int slab(const bvh_node *n, v3 ro, float mint, float maxt, v3 invDir, const uint32_t neg[3]) {
    const float *b[2] = {n->pmin, n->pmax};
    float tmin = (b[neg[0]][0] - ro.x) * invDir.x, tmax = (b[1 - neg[0]][0] - ro.x) * invDir.x;
    float tymin = (b[neg[1]][1] - ro.y) * invDir.y, tymax = (b[1 - neg[1]][1] - ro.y) * invDir.y;
    if ((tmin > tymax) || (tymin > tmax)) return 0;
    if (tymin > tmin) tmin = tymin;
    if (tymax < tmax) tmax = tymax;
    float tzmin = (b[neg[2]][2] - ro.z) * invDir.z, tzmax = (b[1 - neg[2]][2] - ro.z) * invDir.z;
    if ((tmin > tzmax) || (tzmin > tmax)) return 0;
    if (tzmin > tmin) tmin = tzmin;
    if (tzmax < tmax) tmax = tzmax;
    return (tmin < maxt) && (tmax > mint);
}

/* BVHAccel::Intersect (bvh.cpp:388-439) when closest != 0, IntersectP (:442-488) otherwise.
 * counters (optional): [0] slab tests, [1] primitive tests -- the PBRT_BVH_* probe sites. */
static int bvh_trace(const scene_t *sc, v3 ro, v3 rd, float mint, float *maxt, int closest, float *b1h, float *b2h, uint64_t *cnt) {
    v3 invDir = V(1.f / rd.x, 1.f / rd.y, 1.f / rd.z);
    uint32_t neg[3] = {invDir.x < 0, invDir.y < 0, invDir.z < 0};
    uint32_t todo[64], todoOffset = 0, nodeNum = 0;
    int hitPrim = -1;
    while (1) {
        const bvh_node *node = &sc->nodes[nodeNum];
        if (cnt) cnt[0]++;
        if (slab(node, ro, mint, *maxt, invDir, neg)) {
            if (node->nPrimitives > 0) {
                for (uint32_t i = 0; i < node->nPrimitives; ++i) {
                    const prim_t *pr = &sc->prims[node->offset + i];
                    float t, b1 = 0, b2 = 0; int h;
                    if (cnt) cnt[1]++;
                    if (pr->type == 0) h = tri_intersect(sc, pr->index, ro, rd, mint, *maxt, &t, &b1, &b2);
                    else h = sphere_intersect(&sc->spheres[pr->index], ro, rd, mint, *maxt, &t, NULL);
                    if (h) {
                        hitPrim = (int)(node->offset + i);
                        if (!closest) return hitPrim;
                        *maxt = t; if (b1h) { *b1h = b1; *b2h = b2; }     /* core/primitive.cpp:191 */
                    }
                }
                if (todoOffset == 0) break;
                nodeNum = todo[--todoOffset];
            } else {
                if (neg[node->axis]) { todo[todoOffset++] = nodeNum + 1; nodeNum = node->offset; }
                else { todo[todoOffset++] = node->offset; nodeNum = nodeNum + 1; }
            }
        } else {
            if (todoOffset == 0) break;
            nodeNum = todo[--todoOffset];
        }
    }
    return hitPrim;
}

/* shading point / normal of a triangle hit: trianglemesh.inl:92-133,224-299; diffgeom.cpp:40-55 */
static void tri_shading(const scene_t *sc, int tri, v3 ro, v3 rd, float t, float b1, float b2, v3 *pOut, v3 *nOut) {
    const int32_t *v = sc->idx + 3 * tri;
    v3 p1 = V(sc->P[3 * v[0]], sc->P[3 * v[0] + 1], sc->P[3 * v[0] + 2]);
    v3 p2 = V(sc->P[3 * v[1]], sc->P[3 * v[1] + 1], sc->P[3 * v[1] + 2]);
    v3 p3 = V(sc->P[3 * v[2]], sc->P[3 * v[2] + 1], sc->P[3 * v[2] + 2]);
    v3 e1 = sub(p2, p1), e2 = sub(p3, p1);
    float uvs[3][2];
    if (sc->UV) for (int k = 0; k < 3; ++k) { uvs[k][0] = sc->UV[2 * v[k]]; uvs[k][1] = sc->UV[2 * v[k] + 1]; }
    else { uvs[0][0] = 0; uvs[0][1] = 0; uvs[1][0] = 1; uvs[1][1] = 0; uvs[2][0] = 1; uvs[2][1] = 1; }
    float du1 = uvs[0][0] - uvs[2][0], du2 = uvs[1][0] - uvs[2][0], dv1 = uvs[0][1] - uvs[2][1], dv2 = uvs[1][1] - uvs[2][1];
    v3 dp1 = sub(p1, p3), dp2 = sub(p2, p3), dpdu, dpdv;
    float determinant = du1 * dv2 - dv1 * du2;
    if (determinant == 0.f) coordinate_system(normalize(cross(e2, e1)), &dpdu, &dpdv);
    else {
        float invdet = 1.f / determinant;
        dpdu = mul(sub(mul(dp1, dv2), mul(dp2, dv1)), invdet);
        dpdv = mul(add(mul(dp1, -du2), mul(dp2, du1)), invdet);
    }
    float b0 = 1 - b1 - b2;
    float tu = b0 * uvs[0][0] + b1 * uvs[1][0] + b2 * uvs[2][0];
    float tv = b0 * uvs[0][1] + b1 * uvs[1][1] + b2 * uvs[2][1];
    *pOut = add(ro, mul(rd, t));
    v3 nn;
    if (!sc->N) nn = normalize(cross(dpdu, dpdv));
    else {
        float A[2][2] = {{uvs[1][0] - uvs[0][0], uvs[2][0] - uvs[0][0]}, {uvs[1][1] - uvs[0][1], uvs[2][1] - uvs[0][1]}};
        float C[2] = {tu - uvs[0][0], tv - uvs[0][1]}, b[3];
        float det = A[0][0] * A[1][1] - A[0][1] * A[1][0];   /* core/transform.cpp:39-49 */
        int ok = !(fabsf(det) < 1e-10f);
        if (ok) {
            b[1] = (A[1][1] * C[0] - A[0][1] * C[1]) / det;
            b[2] = (A[0][0] * C[1] - A[1][0] * C[0]) / det;
            if (isnan(b[1]) || isnan(b[2])) ok = 0;
        }
        if (!ok) b[0] = b[1] = b[2] = 1.f / 3.f; else b[0] = 1.f - b[1] - b[2];
        v3 n0 = V(sc->N[3 * v[0]], sc->N[3 * v[0] + 1], sc->N[3 * v[0] + 2]);
        v3 n1 = V(sc->N[3 * v[1]], sc->N[3 * v[1] + 1], sc->N[3 * v[1] + 2]);
        v3 n2 = V(sc->N[3 * v[2]], sc->N[3 * v[2] + 1], sc->N[3 * v[2] + 2]);
        v3 ns = normalize(xf_normal(sc->o2w + 16, add(add(mul(n0, b[0]), mul(n1, b[1])), mul(n2, b[2]))));
        v3 ss = normalize(dpdu), ts = cross(ss, ns);
        if (ts.x * ts.x + ts.y * ts.y + ts.z * ts.z > 0.f) { ts = normalize(ts); ss = cross(ts, ns); }
        else coordinate_system(ns, &ss, &ts);
        nn = normalize(cross(ss, ts));
    }
    if (sc->flip) nn = mul(nn, -1.f);
    *nOut = nn;
}

static scene_t make_scene(const void *nodes, const void *prims, const float *P, const float *N, const float *UV, const int32_t *idx,
                          const void *spheres, const float *o2w, int flip) {
    scene_t sc; sc.nodes = (const bvh_node *)nodes; sc.prims = (const prim_t *)prims; sc.P = P; sc.N = N; sc.UV = UV; sc.idx = idx;
    sc.spheres = (const light_t *)spheres; sc.o2w = o2w; sc.flip = flip;
    return sc;
}

/* ---- exported: BVH ------------------------------------------------------------------------ */
void orc_trace_closest(const void *nodes, const void *prims, const float *P, const float *N, const float *UV, const int32_t *idx,
                       const void *spheres, const float *o2w, int flip, const void *raysv, uint64_t n, void *hitsv, uint64_t *cnt) {
    scene_t sc = make_scene(nodes, prims, P, N, UV, idx, spheres, o2w, flip);
    const ray_t *rays = (const ray_t *)raysv; hit_t *hits = (hit_t *)hitsv;
    for (uint64_t i = 0; i < n; ++i) {
        v3 ro = V(rays[i].o[0], rays[i].o[1], rays[i].o[2]), rd = V(rays[i].d[0], rays[i].d[1], rays[i].d[2]);
        float maxt = rays[i].maxt, b1 = 0, b2 = 0;
        int prim = bvh_trace(&sc, ro, rd, rays[i].mint, &maxt, 1, &b1, &b2, cnt);
        hit_t h; memset(&h, 0, sizeof h); h.prim = prim; h.hit = prim >= 0;
        if (prim >= 0) {
            h.t = maxt; h.wo[0] = -rd.x; h.wo[1] = -rd.y; h.wo[2] = -rd.z;
            v3 p, nn = V(0, 0, 0);
            if (sc.prims[prim].type == 0) { tri_shading(&sc, sc.prims[prim].index, ro, rd, maxt, b1, b2, &p, &nn); h.rayEpsilon = 1e-3f * maxt; h.hasBssrdf = 1; }
            else { p = add(ro, mul(rd, maxt)); h.rayEpsilon = 5e-4f * maxt; }
            h.p[0] = p.x; h.p[1] = p.y; h.p[2] = p.z; h.n[0] = nn.x; h.n[1] = nn.y; h.n[2] = nn.z;
        }
        hits[i] = h;
    }
}
void orc_trace_any(const void *nodes, const void *prims, const float *P, const int32_t *idx, const void *spheres,
                   const void *raysv, uint64_t n, uint8_t *out, uint64_t *cnt) {
    scene_t sc = make_scene(nodes, prims, P, NULL, NULL, idx, spheres, NULL, 0);
    const ray_t *rays = (const ray_t *)raysv;
    for (uint64_t i = 0; i < n; ++i) {
        float maxt = rays[i].maxt;
        out[i] = bvh_trace(&sc, V(rays[i].o[0], rays[i].o[1], rays[i].o[2]), V(rays[i].d[0], rays[i].d[1], rays[i].d[2]), rays[i].mint, &maxt, 0, NULL, NULL, cnt) >= 0;
    }
}

/* ---- irradiance pass ---------------------------------------------------------------------- */
/* MultipoleBSSRDFData::rho (core/multipole.cpp:458-463), one channel */
static float rho_hd(const float *rho /*[1025][30]*/, float costheta, int c) {
    float fid = costheta * (RHO_LEN - 1);
    int id = (int)fid; if (id < 0) id = 0; if (id > RHO_LEN - 2) id = RHO_LEN - 2;
    float lerp = fid - (float)id;
    return (1.f - lerp) * rho[id * NS + c] + lerp * rho[(id + 1) * NS + c];
}

/* DiffuseAreaLight::Sample_L for one sphere (lights/diffuse.cpp:69-81, core/light.cpp:145-166,
 * shapes/sphere.cpp:236-273, core/light.h:87-92).  Returns 1 if the sample may contribute. */
static int sample_light(const light_t *L, v3 p, float pEps, v3 spn, float u1, float u2, v3 *wi, float *pdf, v3 *so, v3 *sd, float *smint, float *smaxt) {
    v3 Pcenter = xf_point(L->o2w, V(0, 0, 0));
    float radius = L->radius;
    int reverse = L->reverse != 0.f;
    v3 pt, ns;
    int inside = dist2(p, Pcenter) - radius * radius < 1e-4f;
    if (inside) {   /* Sphere::Sample(u1,u2) sphere.cpp:229-234 + UniformSampleSphere */
        float z = 1.f - 2.f * u1;
        float r = sqrtf(fmaxf(0.f, 1.f - z * z));
        float phi = 2.f * M_PI * u2;
        v3 po = V(radius * (r * cosf(phi)), radius * (r * sinf(phi)), radius * z);
        ns = normalize(xf_normal(L->w2o, po));
        if (reverse) ns = mul(ns, -1.f);
        pt = xf_point(L->o2w, po);
    } else {
        v3 wc = normalize(sub(Pcenter, p)), wcX, wcY;
        coordinate_system(wc, &wcX, &wcY);
        float sinThetaMax2 = radius * radius / dist2(p, Pcenter);
        float cosThetaMax = sqrtf(fmaxf(0.f, 1.f - sinThetaMax2));
        float costheta = (1.f - u1) * cosThetaMax + u1 * 1.f;      /* UniformSampleCone, montecarlo.cpp:413-420 */
        float sintheta = sqrtf(1.f - costheta * costheta);
        float phi = u2 * 2.f * M_PI;
        v3 dir = add(add(mul(wcX, cosf(phi) * sintheta), mul(wcY, sinf(phi) * sintheta)), mul(wc, costheta));
        float thit;
        if (!sphere_intersect(L, p, dir, 1e-3f, INFINITY, &thit, NULL)) thit = dot(sub(Pcenter, p), normalize(dir));
        pt = add(p, mul(dir, thit));
        ns = normalize(sub(pt, Pcenter));
        if (reverse) ns = mul(ns, -1.f);
    }
    v3 rdir = sub(pt, p);          /* ShapeSet::Sample re-intersection, light.cpp:150-158 */
    float thit = 1.f, th; v3 nn;
    if (sphere_intersect(L, p, rdir, 1e-3f, INFINITY, &th, &nn)) { thit = th; ns = nn; }
    v3 ps = add(p, mul(rdir, thit));
    *wi = normalize(sub(ps, p));
    float area = L->phiMax * radius * (L->zmax - L->zmin), spdf;
    if (inside) {   /* Shape::Pdf(p, wi), core/shape.cpp */
        float t3; v3 n3;
        if (!sphere_intersect(L, p, *wi, 1e-3f, INFINITY, &t3, &n3)) spdf = 0.f;
        else {
            v3 ph = add(p, mul(*wi, t3));
            spdf = dist2(p, ph) / (fabsf(dot(n3, mul(*wi, -1.f))) * area);
            if (isinf(spdf)) spdf = 0.f;
        }
    } else {
        float sinThetaMax2 = radius * radius / dist2(p, Pcenter);
        float cosThetaMax = sqrtf(fmaxf(0.f, 1.f - sinThetaMax2));
        spdf = 1.f / (2.f * M_PI * (1.f - cosThetaMax));           /* UniformConePdf */
    }
    float pd = 0.f; pd += area * spdf; *pdf = pd / area;           /* ShapeSet::Pdf, light.cpp:161-166 */
    float dist = len(sub(p, ps));
    float inv = 1.f / dist;
    *so = p; *sd = mul(sub(ps, p), inv); *smint = pEps; *smaxt = dist * (1.f - 1e-3f);
    int lit = dot(ns, mul(*wi, -1.f)) > 0.f;
    if (dot(*wi, spn) <= 0.) return 0;
    if (!lit || *pdf == 0.f) return 0;
    return 1;
}

/* IrradianceTask::Run (integrators/multipolesubsurface.cpp:72-152) for ALL tasks in order
 * (= the reference at --ncores 1).  albedo may be NULL (constant 1). */
void orc_irradiance(const void *nodes, const void *prims, const float *P, const int32_t *idx, const void *lightsv, uint32_t nLights,
                    const void *ptsv, uint64_t nPts, uint32_t nTasks, float mix, const float *rho, const float *albedo, void *outv, uint64_t *cnt) {
    scene_t sc = make_scene(nodes, prims, P, NULL, NULL, idx, lightsv, NULL, 0);
    const light_t *lights = (const light_t *)lightsv;
    const spoint_t *pts = (const spoint_t *)ptsv; ipoint_t *out = (ipoint_t *)outv;
    for (uint32_t task = 0; task < nTasks; ++task) {
        uint64_t idxStart = (uint64_t)(int)((uint64_t)task * nPts / nTasks), idxEnd = (uint64_t)(int)((uint64_t)(task + 1) * nPts / nTasks);
        if (idxStart == idxEnd) continue;
        rng_t rng; rng_seed(&rng, task * 47u); rng.mti = 624;
        for (uint64_t i = idxStart; i < idxEnd; ++i) {
            const spoint_t *sp = &pts[i];
            v3 p = V(sp->p[0], sp->p[1], sp->p[2]), n = V(sp->n[0], sp->n[1], sp->n[2]);
            float E[NS]; for (int c = 0; c < NS; ++c) E[c] = 0.f;
            for (uint32_t j = 0; j < nLights; ++j) {
                float El[NS]; for (int c = 0; c < NS; ++c) El[c] = 0.f;
                uint32_t nS = 1; while ((int)nS < (int)lights[j].nSamples) nS <<= 1;
                uint32_t s0 = rng_uint(&rng), s1 = rng_uint(&rng), sc2 = rng_uint(&rng);
                for (uint32_t s = 0; s < nS; ++s) {
                    float u0 = van_der_corput(s, s0), u1 = sobol2(s, s1), uc = van_der_corput(s, sc2); (void)uc;
                    v3 wi, so, sd; float pdf, mint, maxt;
                    if (!sample_light(&lights[j], p, sp->rayEpsilon, n, u0, u1, &wi, &pdf, &so, &sd, &mint, &maxt)) continue;
                    if (cnt) cnt[2]++;
                    if (bvh_trace(&sc, so, sd, mint, &maxt, 0, NULL, NULL, cnt) >= 0) continue;
                    float costheta = fminf(fabsf(dot(wi, n)), 1.f);
                    for (int c = 0; c < NS; ++c) {
                        float Ft = 1.f - rho_hd(rho, costheta, c);
                        El[c] += ((Ft * lights[j].Lemit[c]) * costheta) / pdf;
                    }
                }
                for (int c = 0; c < NS; ++c) E[c] += El[c] / (float)(int)nS;
            }
            if (albedo) for (int c = 0; c < NS; ++c) E[c] *= powf(albedo[i * NS + c], mix);
            memcpy(out[i].p, sp->p, 12); memcpy(out[i].n, sp->n, 12); memcpy(out[i].E, E, sizeof E);
            out[i].area = sp->area; out[i].rayEpsilon = sp->rayEpsilon;
        }
    }
}

/* scramble words only: [nPts][nLights][3] */
void orc_scrambles(uint64_t nPts, uint32_t nTasks, uint32_t nLights, uint32_t *out) {
    for (uint32_t task = 0; task < nTasks; ++task) {
        uint64_t idxStart = (uint64_t)(int)((uint64_t)task * nPts / nTasks), idxEnd = (uint64_t)(int)((uint64_t)(task + 1) * nPts / nTasks);
        if (idxStart == idxEnd) continue;
        rng_t rng; rng_seed(&rng, task * 47u); rng.mti = 624;
        for (uint64_t i = idxStart; i < idxEnd; ++i)
            for (uint32_t j = 0; j < nLights * 3; ++j) out[i * nLights * 3 + j] = rng_uint(&rng);
    }
}

/* ---- octree: integrators/diffusionutil.h:86-234 ----------------------------------------------- */
typedef struct onode {
    v3 p; int isLeaf; float Et[NS]; float sumArea;
    union { struct onode *children[8]; const ipoint_t *ips[8]; } u;
} onode;

typedef struct { onode **blocks; size_t nBlocks, used; } arena_t;
#define ARENA_BLOCK 65536
static onode *arena_alloc(arena_t *a) {
    if (a->nBlocks == 0 || a->used == ARENA_BLOCK) {
        a->blocks = (onode **)realloc(a->blocks, (a->nBlocks + 1) * sizeof(onode *));
        a->blocks[a->nBlocks++] = (onode *)malloc(ARENA_BLOCK * sizeof(onode));
        a->used = 0;
    }
    onode *n = &a->blocks[a->nBlocks - 1][a->used++];
    memset(n, 0, sizeof *n);
    n->isLeaf = 1;            /* SubsurfaceOctreeNode() :88-93 */
    return n;
}
static void arena_free(arena_t *a) { for (size_t i = 0; i < a->nBlocks; ++i) free(a->blocks[i]); free(a->blocks); }

typedef struct { v3 pMin, pMax; } bbox;
static bbox child_bound(int child, bbox b, v3 pMid) {   /* core/octree.h:87-97 */
    bbox c;
    c.pMin.x = (child & 4) ? pMid.x : b.pMin.x; c.pMax.x = (child & 4) ? b.pMax.x : pMid.x;
    c.pMin.y = (child & 2) ? pMid.y : b.pMin.y; c.pMax.y = (child & 2) ? b.pMax.y : pMid.y;
    c.pMin.z = (child & 1) ? pMid.z : b.pMin.z; c.pMax.z = (child & 1) ? b.pMax.z : pMid.z;
    return c;
}
static v3 mid_of(bbox b) { return V(.5f * b.pMin.x + .5f * b.pMax.x, .5f * b.pMin.y + .5f * b.pMax.y, .5f * b.pMin.z + .5f * b.pMax.z); }

static void insert(onode *node, bbox nb, const ipoint_t *ip, arena_t *arena) {   /* :94-132 */
    v3 pMid = mid_of(nb);
    if (node->isLeaf) {
        for (int i = 0; i < 8; ++i) if (!node->u.ips[i]) { node->u.ips[i] = ip; return; }
        node->isLeaf = 0;
        const ipoint_t *local[8];
        for (int i = 0; i < 8; ++i) { local[i] = node->u.ips[i]; node->u.children[i] = NULL; }
        for (int i = 0; i < 8; ++i) {
            const ipoint_t *q = local[i];
            int child = (q->p[0] > pMid.x ? 4 : 0) + (q->p[1] > pMid.y ? 2 : 0) + (q->p[2] > pMid.z ? 1 : 0);
            if (!node->u.children[child]) node->u.children[child] = arena_alloc(arena);
            insert(node->u.children[child], child_bound(child, nb, pMid), q, arena);
        }
    }
    int child = (ip->p[0] > pMid.x ? 4 : 0) + (ip->p[1] > pMid.y ? 2 : 0) + (ip->p[2] > pMid.z ? 1 : 0);
    if (!node->u.children[child]) node->u.children[child] = arena_alloc(arena);
    insert(node->u.children[child], child_bound(child, nb, pMid), ip, arena);
}

static const float *g_cieY;   /* SampledSpectrum::Y */
static float spectrum_y(const float *c) {   /* core/spectrum.h:387-393 */
    float yy = 0.f;
    for (int i = 0; i < NS; ++i) yy += g_cieY[i] * c[i];
    return yy * (float)(700 - 400) / (float)(106.856895f * NS);
}

static void init_hierarchy(onode *node) {   /* :133-173 (normals are not used by Mo while ModifiedDistanceSquared is #if 1) */
    float sumWt = 0.f;
    if (node->isLeaf) {
        for (int i = 0; i < 8; ++i) {
            if (!node->u.ips[i]) break;
            const ipoint_t *ip = node->u.ips[i];
            float ipEt[NS];
            for (int c = 0; c < NS; ++c) ipEt[c] = ip->E[c] * ip->area;
            float wt = spectrum_y(ipEt);
            for (int c = 0; c < NS; ++c) node->Et[c] += ipEt[c];
            node->p.x += wt * ip->p[0]; node->p.y += wt * ip->p[1]; node->p.z += wt * ip->p[2];
            sumWt += wt;
            node->sumArea += ip->area;
        }
    } else {
        for (int i = 0; i < 8; ++i) {
            onode *ch = node->u.children[i];
            if (!ch) continue;
            init_hierarchy(ch);
            float wt = spectrum_y(ch->Et);
            for (int c = 0; c < NS; ++c) node->Et[c] += ch->Et[c];
            node->p.x += wt * ch->p.x; node->p.y += wt * ch->p.y; node->p.z += wt * ch->p.z;
            sumWt += wt;
            node->sumArea += ch->sumArea;
        }
    }
    if (sumWt > 0.f) { float inv = 1.f / sumWt; node->p.x *= inv; node->p.y *= inv; node->p.z *= inv; }
}

typedef struct { const float *table; const float *rcp; uint32_t len; } profile_t;
/* sampleProfile (core/multipole.cpp:60-73) */
static float sample_profile(const profile_t *pr, int c, float d2) {
    double fSegId = d2 * pr->rcp[c];
    if (fSegId >= (double)(pr->len - 1)) return 0.f;
    uint32_t segId = (uint32_t)fSegId;
    float lerpAmount = (float)(fSegId - (double)segId);
    const float *data = pr->table + (size_t)c * pr->len;
    return (1.f - lerpAmount) * data[segId] + lerpAmount * data[segId + 1];
}

static int is_black(const float *c) { for (int i = 0; i < NS; ++i) if (c[i] != 0.) return 0; return 1; }

/* SubsurfaceOctreeNode::Mo (:175-210): recursive, children 0..7, nested sums.
 * counters: [0] nodes visited, [1] cluster contributions, [2] point contributions (probe sites) */
static void mo_rec(const onode *node, bbox nb, v3 pt, const profile_t *pr, float maxError, float *out, uint64_t *cnt) {
    for (int c = 0; c < NS; ++c) out[c] = 0.f;
    if (cnt) cnt[0]++;
    if (is_black(node->Et)) return;
    float d2n = dist2(pt, node->p);
    float dw = node->sumArea / d2n;
    int inside = pt.x >= nb.pMin.x && pt.x <= nb.pMax.x && pt.y >= nb.pMin.y && pt.y <= nb.pMax.y && pt.z >= nb.pMin.z && pt.z <= nb.pMax.z;
    if (dw < maxError && !inside) {
        if (cnt) cnt[1]++;
        for (int c = 0; c < NS; ++c) out[c] = sample_profile(pr, c, d2n) * node->Et[c];
        return;
    }
    if (node->isLeaf) {
        for (int i = 0; i < 8; ++i) {
            const ipoint_t *ip = node->u.ips[i];
            if (!ip) break;
            if (cnt) cnt[2]++;
            if (!is_black(ip->E)) {
                float d2 = dist2(pt, V(ip->p[0], ip->p[1], ip->p[2]));
                for (int c = 0; c < NS; ++c) out[c] += sample_profile(pr, c, d2) * ip->E[c] * ip->area;
            }
        }
    } else {
        v3 pMid = mid_of(nb);
        for (int child = 0; child < 8; ++child) {
            if (!node->u.children[child]) continue;
            float sub[NS];
            mo_rec(node->u.children[child], child_bound(child, nb, pMid), pt, pr, maxError, sub, cnt);
            for (int c = 0; c < NS; ++c) out[c] += sub[c];
        }
    }
}

typedef struct { arena_t arena; onode *root; bbox bounds; uint64_t nNodes; } octree_t;

/* Preprocess' build loop (integrators/multipolesubsurface.cpp:218-237) */
void *orc_octree_build(const void *ipsv, uint64_t n, const float *cieY30) {
    const ipoint_t *ips = (const ipoint_t *)ipsv;
    octree_t *t = (octree_t *)calloc(1, sizeof *t);
    g_cieY = cieY30;
    t->root = arena_alloc(&t->arena);
    t->bounds.pMin = V(INFINITY, INFINITY, INFINITY); t->bounds.pMax = V(-INFINITY, -INFINITY, -INFINITY);
    for (uint64_t i = 0; i < n; ++i) {
        t->bounds.pMin = V(fminf(t->bounds.pMin.x, ips[i].p[0]), fminf(t->bounds.pMin.y, ips[i].p[1]), fminf(t->bounds.pMin.z, ips[i].p[2]));
        t->bounds.pMax = V(fmaxf(t->bounds.pMax.x, ips[i].p[0]), fmaxf(t->bounds.pMax.y, ips[i].p[1]), fmaxf(t->bounds.pMax.z, ips[i].p[2]));
    }
    for (uint64_t i = 0; i < n; ++i) insert(t->root, t->bounds, &ips[i], &t->arena);
    init_hierarchy(t->root);
    t->nNodes = (t->arena.nBlocks - 1) * (uint64_t)ARENA_BLOCK + t->arena.used;
    return t;
}
void orc_octree_free(void *tv) { octree_t *t = (octree_t *)tv; arena_free(&t->arena); free(t); }
/* root34 = {p[3], sumArea, Et[30]}, bounds6 */
uint64_t orc_octree_info(const void *tv, float *root34, float *bounds6) {
    const octree_t *t = (const octree_t *)tv;
    root34[0] = t->root->p.x; root34[1] = t->root->p.y; root34[2] = t->root->p.z; root34[3] = t->root->sumArea;
    memcpy(root34 + 4, t->root->Et, sizeof t->root->Et);
    bounds6[0] = t->bounds.pMin.x; bounds6[1] = t->bounds.pMin.y; bounds6[2] = t->bounds.pMin.z;
    bounds6[3] = t->bounds.pMax.x; bounds6[4] = t->bounds.pMax.y; bounds6[5] = t->bounds.pMax.z;
    return t->nNodes;
}
void orc_mo(const void *tv, const float *p3, uint64_t n, const float *table, const float *rcp, uint32_t len, float maxError, float *Mo30, uint64_t *cnt) {
    const octree_t *t = (const octree_t *)tv;
    profile_t pr = {table, rcp, len};
    for (uint64_t i = 0; i < n; ++i)
        mo_rec(t->root, t->bounds, V(p3[3 * i], p3[3 * i + 1], p3[3 * i + 2]), &pr, maxError, Mo30 + NS * i, cnt);
}
/* Li subsurface term (integrators/multipolesubsurface.cpp:279-287); albedo NULL = 1 */
void orc_li_sss(const void *hitsv, uint64_t n, const float *Mo30, const float *rho, const float *albedo, float mix, float *L30) {
    const hit_t *hits = (const hit_t *)hitsv;
    for (uint64_t i = 0; i < n; ++i) {
        for (int c = 0; c < NS; ++c) L30[NS * i + c] = 0.f;
        if (!hits[i].hasBssrdf) continue;
        v3 nn = V(hits[i].n[0], hits[i].n[1], hits[i].n[2]), wo = V(hits[i].wo[0], hits[i].wo[1], hits[i].wo[2]);
        float costheta = fminf(fabsf(dot(wo, nn)), 1.f);
        for (int c = 0; c < NS; ++c) {
            float Ft = 1.f - rho_hd(rho, costheta, c);
            float a = albedo ? powf(albedo[NS * i + c], 1.f - mix) : 1.f;
            float v = ((0.31830988618379067154f * Ft) * Mo30[NS * i + c]) * a;
            L30[NS * i + c] = v > 0.f ? v : 0.f;
        }
    }
}

// skin_geom.cpp -- BVHAccel build and the surface-point tessellator (host mirror).
#include "skin_scene.h"
#include <functional>

namespace skin {

// ---------------------------------------------------------------------------------------------
// BVHAccel (accelerators/bvh.cpp:153-380): SAH with 12 buckets, leaves of <= maxPrimsInNode
// primitives, depth-first flattening.  std::partition / std::nth_element are the same library
// calls the reference makes on the same element order, so the tree is the reference's tree.
// ---------------------------------------------------------------------------------------------
struct BVHAccel::BuildNode {
    BBox bounds;
    BuildNode *children[2] = {nullptr, nullptr};
    uint32_t splitAxis = 0, firstPrimOffset = 0, nPrimitives = 0;
};

BVHAccel::BVHAccel(const std::vector<BBox> &primBounds, uint32_t mp, const std::string &sm) {
    maxPrimsInNode = std::min(255u, mp);
    splitMethod = sm == "middle" ? 0 : sm == "equal" ? 1 : 2;
    if (primBounds.empty()) return;
    std::vector<PrimInfo> buildData;
    buildData.reserve(primBounds.size());
    for (uint32_t i = 0; i < primBounds.size(); ++i) {
        PrimInfo pi; pi.primitiveNumber = i; pi.bounds = primBounds[i];
        pi.centroid = .5f * primBounds[i].pMin + .5f * primBounds[i].pMax;   // bvh.cpp:47
        buildData.push_back(pi);
    }
    uint32_t totalNodes = 0;
    orderedPrims.reserve(primBounds.size());
    BuildNode *root = recursiveBuild(buildData, 0, (uint32_t)primBounds.size(), &totalNodes);
    nodes.resize(totalNodes);
    memset(nodes.data(), 0, totalNodes * sizeof(skc_bvh_node));
    uint32_t offset = 0;
    flatten(root, &offset);
    for (BuildNode *n : arena) delete n;
    arena.clear();
}

BVHAccel::BuildNode *BVHAccel::recursiveBuild(std::vector<PrimInfo> &buildData, uint32_t start, uint32_t end, uint32_t *totalNodes) {
    (*totalNodes)++;
    BuildNode *node = new BuildNode();
    arena.push_back(node);
    BBox bbox;
    for (uint32_t i = start; i < end; ++i) bbox = Union(bbox, buildData[i].bounds);
    uint32_t nPrimitives = end - start;
    auto makeLeaf = [&]() {
        node->firstPrimOffset = (uint32_t)orderedPrims.size();
        for (uint32_t i = start; i < end; ++i) orderedPrims.push_back(buildData[i].primitiveNumber);
        node->nPrimitives = nPrimitives; node->bounds = bbox;
        return node;
    };
    if (nPrimitives == 1) return makeLeaf();
    BBox centroidBounds;
    for (uint32_t i = start; i < end; ++i) centroidBounds = Union(centroidBounds, buildData[i].centroid);
    int dim = centroidBounds.MaximumExtent();
    uint32_t mid = (start + end) / 2;
    if (centroidBounds.pMax[dim] == centroidBounds.pMin[dim]) return makeLeaf();
    auto byCentroid = [dim](const PrimInfo &a, const PrimInfo &b) { return a.centroid[dim] < b.centroid[dim]; };
    bool fallthroughEqual = false;
    if (splitMethod == 0) {
        float pmid = .5f * (centroidBounds.pMin[dim] + centroidBounds.pMax[dim]);
        PrimInfo *midPtr = std::partition(&buildData[start], &buildData[end - 1] + 1, [dim, pmid](const PrimInfo &a) { return a.centroid[dim] < pmid; });
        mid = (uint32_t)(midPtr - &buildData[0]);
        if (!(mid != start && mid != end)) fallthroughEqual = true;
    }
    if (splitMethod == 1 || fallthroughEqual) {
        mid = (start + end) / 2;
        std::nth_element(&buildData[start], &buildData[mid], &buildData[end - 1] + 1, byCentroid);
    } else if (splitMethod == 2) {
        if (nPrimitives <= 4) {
            mid = (start + end) / 2;
            std::nth_element(&buildData[start], &buildData[mid], &buildData[end - 1] + 1, byCentroid);
        } else {
            const int nBuckets = 12;
            struct BucketInfo { int count = 0; BBox bounds; };
            BucketInfo buckets[nBuckets];
            for (uint32_t i = start; i < end; ++i) {
                int b = nBuckets * ((buildData[i].centroid[dim] - centroidBounds.pMin[dim]) / (centroidBounds.pMax[dim] - centroidBounds.pMin[dim]));
                if (b == nBuckets) b = nBuckets - 1;
                buckets[b].count++;
                buckets[b].bounds = Union(buckets[b].bounds, buildData[i].bounds);
            }
            float cost[nBuckets - 1];
            for (int i = 0; i < nBuckets - 1; ++i) {
                BBox b0, b1;
                int count0 = 0, count1 = 0;
                for (int j = 0; j <= i; ++j) { b0 = Union(b0, buckets[j].bounds); count0 += buckets[j].count; }
                for (int j = i + 1; j < nBuckets; ++j) { b1 = Union(b1, buckets[j].bounds); count1 += buckets[j].count; }
                cost[i] = .125f + (count0 * b0.SurfaceArea() + count1 * b1.SurfaceArea()) / bbox.SurfaceArea();
            }
            float minCost = cost[0];
            uint32_t minCostSplit = 0;
            for (int i = 1; i < nBuckets - 1; ++i) if (cost[i] < minCost) { minCost = cost[i]; minCostSplit = i; }
            if (nPrimitives > maxPrimsInNode || minCost < nPrimitives) {
                BBox cb = centroidBounds;
                PrimInfo *pmid = std::partition(&buildData[start], &buildData[end - 1] + 1, [=](const PrimInfo &p) {
                    int b = nBuckets * ((p.centroid[dim] - cb.pMin[dim]) / (cb.pMax[dim] - cb.pMin[dim]));
                    if (b == nBuckets) b = nBuckets - 1;
                    return b <= (int)minCostSplit;
                });
                mid = (uint32_t)(pmid - &buildData[0]);
            } else
                return makeLeaf();
        }
    }
    node->splitAxis = dim; node->nPrimitives = 0; node->bounds = bbox;
    // bvh.cpp:352-356 passes both recursive calls as arguments of InitInterior; g++ evaluates
    // function arguments right to left, so the reference builds the SECOND child first and the
    // ordered primitive list fills from the far end of the tree.
    node->children[1] = recursiveBuild(buildData, mid, end, totalNodes);
    node->children[0] = recursiveBuild(buildData, start, mid, totalNodes);
    // InitInterior (bvh.cpp:73-79): bounds = union of the children
    node->bounds = Union(node->children[0]->bounds, node->children[1]->bounds);
    return node;
}

uint32_t BVHAccel::flatten(BuildNode *node, uint32_t *offset) {
    skc_bvh_node *ln = &nodes[*offset];
    ln->pmin[0] = node->bounds.pMin.x; ln->pmin[1] = node->bounds.pMin.y; ln->pmin[2] = node->bounds.pMin.z;
    ln->pmax[0] = node->bounds.pMax.x; ln->pmax[1] = node->bounds.pMax.y; ln->pmax[2] = node->bounds.pMax.z;
    uint32_t myOffset = (*offset)++;
    if (node->nPrimitives > 0) {
        ln->offset = node->firstPrimOffset; ln->nPrimitives = (uint8_t)node->nPrimitives;
    } else {
        ln->axis = (uint8_t)node->splitAxis; ln->nPrimitives = 0;
        flatten(node->children[0], offset);
        uint32_t second = flatten(node->children[1], offset);
        nodes[myOffset].offset = second;
    }
    return myOffset;
}

// ---------------------------------------------------------------------------------------------
// Tessellator (shapes/trianglemesh.cpp:187-351): DX11-style ring tessellation; one surface point
// per sub-triangle.
// ---------------------------------------------------------------------------------------------
struct BC {
    float b0, b1, b2;
    BC() : b0(0), b1(0), b2(0) {}
    BC(float a, float b, float c) : b0(a), b1(b), b2(c) {}
    V3 Evaluate(V3 v0, V3 v1, V3 v2) const { return b0 * v0 + b1 * v1 + b2 * v2; }
    BC Evaluate(const BC &a, const BC &b, const BC &c) const {
        return BC(b0 * a.b0 + b1 * b.b0 + b2 * c.b0, b0 * a.b1 + b1 * b.b1 + b2 * c.b1, b0 * a.b2 + b1 * b.b2 + b2 * c.b2);
    }
    static float L(float t, float v1, float v2) { return (1.f - t) * v1 + t * v2; }   // core/pbrt.h:250-252
    static BC Lerp(float v, const BC &a, const BC &b) { return BC(L(v, a.b0, b.b0), L(v, a.b1, b.b1), L(v, a.b2, b.b2)); }
};
static const BC kBaryCentric(1.f / 3.f, 1.f / 3.f, 1.f / 3.f);
static inline int Ceil2Int(float v) { return (int)ceilf(v); }

template <class F>
static void matching(BC b0Inner, BC b1Inner, int segsInner, BC b0Outer, BC b1Outer, int segsOuter, F &&domainShader) {   // :319-351
    int innerPos = 0, outerPos = 0;
    while (innerPos < segsInner || outerPos < segsOuter) {
        BC bInner = segsInner ? BC::Lerp((float)(innerPos) / segsInner, b0Inner, b1Inner) : b0Inner;
        BC bOuter = BC::Lerp((float)(outerPos) / segsOuter, b0Outer, b1Outer);
        float slopeInner = (innerPos < segsInner) ? fabsf((float)(innerPos + 1) + 1.f - (float)outerPos / segsOuter * (segsInner + 2)) : INFINITY;
        float slopeOuter = (outerPos < segsOuter) ? fabsf((float)innerPos + 1.f - (float)(outerPos + 1) / segsOuter * (segsInner + 2)) : INFINITY;
        if (slopeInner < slopeOuter) {
            BC bInnerNew = BC::Lerp((float)(innerPos + 1) / segsInner, b0Inner, b1Inner);
            domainShader(bInnerNew, bInner, bOuter);
            innerPos++;
        } else {
            BC bOuterNew = BC::Lerp((float)(outerPos + 1) / segsOuter, b0Outer, b1Outer);
            domainShader(bInner, bOuter, bOuterNew);
            outerPos++;
        }
    }
}

template <class F>
static void tessellator(float tfe0, float tfe1, float tfe2, float tfc, F &&domainShader) {   // :264-316
    int itfe0 = std::max(Ceil2Int(tfe0), 1), itfe1 = std::max(Ceil2Int(tfe1), 1), itfe2 = std::max(Ceil2Int(tfe2), 1), itfc = std::max(Ceil2Int(tfc), 1);
    if (itfe0 > 1 || itfe1 > 1 || itfe2 > 1) itfc = std::max(itfc, 2);
    BC b0(1.f, 0.f, 0.f), b1(0.f, 1.f, 0.f), b2(0.f, 0.f, 1.f), bc(kBaryCentric);
    int rings = (itfc + 1) / 2;
    for (int r = 0; r < rings - 1; r++) {
        int edgeInner = itfc - (rings - r) * 2;
        if (edgeInner >= 0) {
            int edgeOuter = edgeInner + 2;
            BC b0Inner = BC::Lerp((float)r / rings, bc, b0), b1Inner = BC::Lerp((float)r / rings, bc, b1), b2Inner = BC::Lerp((float)r / rings, bc, b2);
            BC b0Outer = BC::Lerp((float)(r + 1) / rings, bc, b0), b1Outer = BC::Lerp((float)(r + 1) / rings, bc, b1), b2Outer = BC::Lerp((float)(r + 1) / rings, bc, b2);
            matching(b0Inner, b1Inner, edgeInner, b0Outer, b1Outer, edgeOuter, domainShader);
            matching(b1Inner, b2Inner, edgeInner, b1Outer, b2Outer, edgeOuter, domainShader);
            matching(b2Inner, b0Inner, edgeInner, b2Outer, b0Outer, edgeOuter, domainShader);
        } else {
            BC b0Outer = BC::Lerp((float)(r + 1) / rings, bc, b0), b1Outer = BC::Lerp((float)(r + 1) / rings, bc, b1), b2Outer = BC::Lerp((float)(r + 1) / rings, bc, b2);
            domainShader(b0Outer, b1Outer, b2Outer);
        }
    }
    int edgeInner = itfc - 2;
    if (edgeInner >= 0) {
        BC b0Inner = BC::Lerp((float)(rings - 1) / rings, bc, b0), b1Inner = BC::Lerp((float)(rings - 1) / rings, bc, b1), b2Inner = BC::Lerp((float)(rings - 1) / rings, bc, b2);
        matching(b0Inner, b1Inner, edgeInner, b0, b1, itfe2, domainShader);
        matching(b1Inner, b2Inner, edgeInner, b1, b2, itfe0, domainShader);
        matching(b2Inner, b0Inner, edgeInner, b2, b0, itfe1, domainShader);
    } else {
        domainShader(b0, b1, b2);
    }
}

static void tessFactors(const TriangleMesh &m, int itri, float minDist, float tf[4]) {   // :201-214
    const V3 &v0 = m.p[m.vertexIndex[itri * 3]], &v1 = m.p[m.vertexIndex[itri * 3 + 1]], &v2 = m.p[m.vertexIndex[itri * 3 + 2]];
    float le0 = Length(v1 - v2), le1 = Length(v2 - v0), le2 = Length(v0 - v1);
    float tfe0 = le0 / minDist * 0.8f, tfe1 = le1 / minDist * 0.8f, tfe2 = le2 / minDist * 0.8f;
    tf[3] = floorf((tfe0 + tfe1 + tfe2) / 3.f + .5f);
    tf[0] = floorf(tfe0 + .5f); tf[1] = floorf(tfe1 + .5f); tf[2] = floorf(tfe2 + .5f);
}

uint64_t TriangleMesh::CountSurfacePoints(float minDist) const {
    uint64_t n = 0;
    for (int itri = 0; itri < ntris; ++itri) {
        float tf[4];
        tessFactors(*this, itri, minDist, tf);
        tessellator(tf[0], tf[1], tf[2], tf[3], [&](BC, BC, BC) { ++n; });
    }
    return n;
}

// Shading normal and uv at a barycentric point: GetDifferentialGeometries
// (shapes/trianglemesh.inl:303-345) + GetShadingGeometry (:224-299) + diffgeom.cpp:40-55
static void shadingAt(const TriangleMesh &m, int itri, const BC &bc, V3 *nn, float *tuOut, float *tvOut) {
    const int *v = &m.vertexIndex[3 * itri];
    const V3 &p1 = m.p[v[0]], &p2 = m.p[v[1]], &p3 = m.p[v[2]];
    V3 e1 = p2 - p1, e2 = p3 - p1;
    float uvs[3][2];
    if (!m.uvs.empty()) { for (int k = 0; k < 3; ++k) { uvs[k][0] = m.uvs[2 * v[k]]; uvs[k][1] = m.uvs[2 * v[k] + 1]; } }
    else { uvs[0][0] = 0.; uvs[0][1] = 0.; uvs[1][0] = 1.; uvs[1][1] = 0.; uvs[2][0] = 1.; uvs[2][1] = 1.; }
    float du1 = uvs[0][0] - uvs[2][0], du2 = uvs[1][0] - uvs[2][0], dv1 = uvs[0][1] - uvs[2][1], dv2 = uvs[1][1] - uvs[2][1];
    V3 dp1 = p1 - p3, dp2 = p2 - p3, dpdu, dpdv;
    float determinant = du1 * dv2 - dv1 * du2;
    if (determinant == 0.f) CoordinateSystem(Normalize(Cross(e2, e1)), &dpdu, &dpdv);
    else {
        float invdet = 1.f / determinant;
        dpdu = (dv2 * dp1 - dv1 * dp2) * invdet;
        dpdv = (-du2 * dp1 + du1 * dp2) * invdet;
    }
    float b0 = bc.b0, b1 = bc.b1, b2 = bc.b2;
    float tu = b0 * uvs[0][0] + b1 * uvs[1][0] + b2 * uvs[2][0];
    float tv = b0 * uvs[0][1] + b1 * uvs[1][1] + b2 * uvs[2][1];
    *tuOut = tu; *tvOut = tv;
    bool flip = m.ReverseOrientation ^ m.TransformSwapsHandedness;
    V3 out;
    if (m.n.empty()) out = Normalize(Cross(dpdu, dpdv));
    else {
        float A[2][2] = {{uvs[1][0] - uvs[0][0], uvs[2][0] - uvs[0][0]}, {uvs[1][1] - uvs[0][1], uvs[2][1] - uvs[0][1]}};
        float C[2] = {tu - uvs[0][0], tv - uvs[0][1]};
        float b[3];
        float det = A[0][0] * A[1][1] - A[0][1] * A[1][0];
        bool ok = !(fabsf(det) < 1e-10f);
        if (ok) {
            b[1] = (A[1][1] * C[0] - A[0][1] * C[1]) / det;
            b[2] = (A[0][0] * C[1] - A[1][0] * C[0]) / det;
            if (std::isnan(b[1]) || std::isnan(b[2])) ok = false;
        }
        if (!ok) b[0] = b[1] = b[2] = 1.f / 3.f;
        else b[0] = 1.f - b[1] - b[2];
        V3 ns = Normalize(m.ObjectToWorld.Normal(b[0] * m.n[v[0]] + b[1] * m.n[v[1]] + b[2] * m.n[v[2]]));
        V3 ss = Normalize(dpdu);
        V3 ts = Cross(ss, ns);
        if (LengthSquared(ts) > 0.f) { ts = Normalize(ts); ss = Cross(ts, ns); }
        else CoordinateSystem(ns, &ss, &ts);
        out = Normalize(Cross(ss, ts));
    }
    if (flip) out = out * -1.f;
    *nn = out;
}

void TriangleMesh::TessellateSurfacePoints(float minDist, uint32_t matId, std::vector<skc_surface_point> &points, bool incenter) const {
    for (int itri = 0; itri < ntris; ++itri) {
        const V3 &v0 = p[vertexIndex[itri * 3]], &v1 = p[vertexIndex[itri * 3 + 1]], &v2 = p[vertexIndex[itri * 3 + 2]];
        float tf[4];
        tessFactors(*this, itri, minDist, tf);
        tessellator(tf[0], tf[1], tf[2], tf[3], [&](BC bv0, BC bv1, BC bv2) {
            V3 sv0 = bv0.Evaluate(v0, v1, v2), sv1 = bv1.Evaluate(v0, v1, v2), sv2 = bv2.Evaluate(v0, v1, v2);
            BC bc;
            if (!incenter) bc = kBaryCentric.Evaluate(bv0, bv1, bv2);
            else {
                float l0 = Length(sv1 - sv2), l1 = Length(sv2 - sv0), l2 = Length(sv0 - sv1);
                BC bic(l0 / (l0 + l1 + l2), l1 / (l0 + l1 + l2), l2 / (l0 + l1 + l2));
                bc = bic.Evaluate(bv0, bv1, bv2);
            }
            skc_surface_point sp;
            V3 pp = bc.Evaluate(v0, v1, v2), nn;
            shadingAt(*this, itri, bc, &nn, &sp.u, &sp.v);
            sp.p[0] = pp.x; sp.p[1] = pp.y; sp.p[2] = pp.z;
            sp.n[0] = nn.x; sp.n[1] = nn.y; sp.n[2] = nn.z;
            sp.materialId = matId;
            sp.area = .5f * Length(Cross(sv1 - sv0, sv2 - sv0));
            sp.rayEpsilon = minDist / 10.f;
            points.push_back(sp);
        });
    }
}

}  // namespace skin

// skin_geom.cpp -- the host side of the Aggregate slot: a BVH builder whose OUTPUT is the reference's tree, and the mesh's
// surface-point methods.
//
// What must equal the reference (accelerators/bvh.cpp:153-380; tests/test_host_mirror.py compares every byte): the flat node
// array and the order of the primitives in the leaves -- the device traversal is bit-exact only on the same tree.  How the
// tree is produced is this file's own: primitives are never moved, the builder permutes an index array over column-wise
// centroid / bounds tables; ranges wait on an explicit work stack instead of recursing; the surface-area heuristic sweeps the
// twelve buckets once from each end instead of re-forming every candidate's two boxes; the flat array is numbered in a second
// pass over the recorded topology.  Two things are inherited because the result depends on them: the C++ library's
// std::partition / std::nth_element (their exact permutations decide every later split), called on the same sequence of
// comparisons; and the order in which ranges are finished -- the reference's compiler evaluates the two recursive calls of
// an interior node right to left, so a node's SECOND half is completed before its first and the leaf primitive list fills
// in that order.
#include "skin_scene.h"
#include <numeric>

namespace skin {

namespace {

constexpr int kBuckets = 12;                 // accelerators/bvh.cpp:286
constexpr float kTraversalCost = .125f;      // :313

struct PrimTable {                           // column-wise; indexed by primitive number, never permuted
    std::vector<float> centroid[3];
    const std::vector<BBox> *bounds = nullptr;
};

struct TopoNode {                            // one node of the tree as decided, in the order ranges were finished
    BBox box;
    uint32_t firstPrim = 0, primCount = 0;   // leaf: slice of the ordered primitive list
    int32_t child[2] = {-1, -1};             // interior: topology ids of the first / second half
    uint8_t axis = 0;
};

struct Pending { uint32_t lo, hi; int32_t node; };

inline int bucketOf(float c, float cmin, float cmax) {
    int b = (int)(kBuckets * ((c - cmin) / (cmax - cmin)));
    return b == kBuckets ? kBuckets - 1 : b;
}

// Where to cut ids[lo, hi) along `axis`, or hi for "keep together".  Reorders ids[lo, hi).
uint32_t chooseCut(std::vector<uint32_t> &ids, uint32_t lo, uint32_t hi, int axis, const BBox &box, float cmin, float cmax,
                   const PrimTable &T, int method, uint32_t maxLeaf) {
    const std::vector<float> &key = T.centroid[axis];
    const uint32_t count = hi - lo;
    auto first = ids.begin() + lo, last = ids.begin() + hi;
    auto medianCut = [&]() {
        const uint32_t mid = (lo + hi) / 2;
        std::nth_element(first, ids.begin() + mid, last, [&key](uint32_t a, uint32_t b) { return key[a] < key[b]; });
        return mid;
    };
    if (method == 1) return medianCut();                                       // "equal"
    if (method == 0) {                                                         // "middle": by the centroid box's midpoint, else median
        const float pivot = .5f * (cmin + cmax);
        const uint32_t cut = (uint32_t)(std::partition(first, last, [&key, pivot](uint32_t a) { return key[a] < pivot; }) - ids.begin());
        return (cut != lo && cut != hi) ? cut : medianCut();
    }
    if (count <= 4) return medianCut();                                        // "sah" on a handful of primitives
    // ---- surface-area heuristic over kBuckets slabs of the centroid range
    uint32_t inBucket[kBuckets] = {0};
    BBox bucketBox[kBuckets];
    for (uint32_t i = lo; i < hi; ++i) {
        const uint32_t id = ids[i];
        const int b = bucketOf(key[id], cmin, cmax);
        ++inBucket[b];
        bucketBox[b] = Union(bucketBox[b], (*T.bounds)[id]);
    }
    // boxes and counts of buckets [s+1, kBuckets) from the far end, then one sweep from the near end (min / max are exact, so
    // the order of the unions does not change a bit of the result)
    BBox above[kBuckets]; uint32_t nAbove[kBuckets];
    {
        BBox acc; uint32_t n = 0;
        for (int s = kBuckets - 1; s >= 0; --s) { above[s] = acc; nAbove[s] = n; acc = Union(acc, bucketBox[s]); n += inBucket[s]; }
    }
    const float wholeArea = box.SurfaceArea();
    int best = 0; float bestCost = 0.f;
    {
        BBox below; uint32_t nBelow = 0;
        for (int s = 0; s < kBuckets - 1; ++s) {
            below = Union(below, bucketBox[s]); nBelow += inBucket[s];
            const float cost = kTraversalCost + ((int)nBelow * below.SurfaceArea() + (int)nAbove[s] * above[s].SurfaceArea()) / wholeArea;
            if (s == 0 || cost < bestCost) { bestCost = cost; best = s; }     // first minimum wins
        }
    }
    if (count <= maxLeaf && !(bestCost < (float)count)) return hi;             // cheaper to keep the range as one leaf
    return (uint32_t)(std::partition(first, last, [&key, cmin, cmax, best](uint32_t a) { return bucketOf(key[a], cmin, cmax) <= best; }) - ids.begin());
}

}  // namespace

BVHAccel::BVHAccel(const std::vector<BBox> &primBounds, uint32_t mp, const std::string &sm) {
    maxPrimsInNode = std::min(255u, mp);
    splitMethod = sm == "middle" ? 0 : sm == "equal" ? 1 : 2;
    const uint32_t n = (uint32_t)primBounds.size();
    if (!n) return;
    PrimTable T;
    T.bounds = &primBounds;
    for (int a = 0; a < 3; ++a) T.centroid[a].resize(n);
    for (uint32_t i = 0; i < n; ++i) {
        const V3 c = .5f * primBounds[i].pMin + .5f * primBounds[i].pMax;     // bvh.cpp:47
        T.centroid[0][i] = c.x; T.centroid[1][i] = c.y; T.centroid[2][i] = c.z;
    }
    std::vector<uint32_t> ids(n);
    std::iota(ids.begin(), ids.end(), 0u);

    // ---- pass 1: decide every range; LIFO, second half on top so that it is finished first
    std::vector<TopoNode> topo;
    topo.reserve(2 * (size_t)n);
    orderedPrims.reserve(n);
    std::vector<Pending> work;
    topo.emplace_back();
    work.push_back({0u, n, 0});
    while (!work.empty()) {
        const Pending w = work.back();
        work.pop_back();
        BBox box, cbox;
        for (uint32_t i = w.lo; i < w.hi; ++i) box = Union(box, primBounds[ids[i]]);
        topo[w.node].box = box;
        uint32_t cut = w.hi;
        int axis = 0;
        if (w.hi - w.lo > 1) {
            for (uint32_t i = w.lo; i < w.hi; ++i) { const uint32_t id = ids[i]; cbox = Union(cbox, V3(T.centroid[0][id], T.centroid[1][id], T.centroid[2][id])); }
            axis = cbox.MaximumExtent();
            const float cmin = cbox.pMin[axis], cmax = cbox.pMax[axis];
            if (cmax != cmin) cut = chooseCut(ids, w.lo, w.hi, axis, box, cmin, cmax, T, splitMethod, maxPrimsInNode);
        }
        if (cut == w.hi) {                                                     // a leaf: its primitives join the ordered list now
            topo[w.node].firstPrim = (uint32_t)orderedPrims.size();
            topo[w.node].primCount = w.hi - w.lo;
            orderedPrims.insert(orderedPrims.end(), ids.begin() + w.lo, ids.begin() + w.hi);
            continue;
        }
        const int32_t a = (int32_t)topo.size(), b = a + 1;
        topo.emplace_back(); topo.emplace_back();
        topo[w.node].child[0] = a; topo[w.node].child[1] = b; topo[w.node].axis = (uint8_t)axis;
        work.push_back({w.lo, cut, a});
        work.push_back({cut, w.hi, b});
    }

    // ---- pass 2: depth-first numbering, first half directly after its parent (bvh.cpp:362-380); an interior node stores where
    // its second half landed
    nodes.assign(topo.size(), skc_bvh_node());
    memset(nodes.data(), 0, nodes.size() * sizeof(skc_bvh_node));
    struct Visit { int32_t node; int32_t parentSlot; };                        // parentSlot >= 0: this is that node's second half
    std::vector<Visit> order;
    order.push_back({0, -1});
    uint32_t next = 0;
    while (!order.empty()) {
        const Visit v = order.back();
        order.pop_back();
        const TopoNode &t = topo[v.node];
        const uint32_t slot = next++;
        if (v.parentSlot >= 0) nodes[v.parentSlot].offset = slot;
        skc_bvh_node &out = nodes[slot];
        out.pmin[0] = t.box.pMin.x; out.pmin[1] = t.box.pMin.y; out.pmin[2] = t.box.pMin.z;
        out.pmax[0] = t.box.pMax.x; out.pmax[1] = t.box.pMax.y; out.pmax[2] = t.box.pMax.z;
        if (t.child[0] < 0) { out.offset = t.firstPrim; out.nPrimitives = (uint8_t)t.primCount; continue; }
        out.axis = t.axis;
        order.push_back({t.child[1], (int32_t)slot});
        order.push_back({t.child[0], -1});
    }
}

// ---------------------------------------------------------------------------------------------
// Surface points (shapes/trianglemesh.cpp:187-351).  One tessellator serves host and device: the job functions of
// csrc/skc_tess.cu, evaluated here through skc_tessellate_host (a host loop in libskincuda, no GPU involved).
// ---------------------------------------------------------------------------------------------
static void meshMatrices(const TriangleMesh &m, float o2w32[32]) {
    memcpy(o2w32, m.ObjectToWorld.m.m, 16 * sizeof(float));
    memcpy(o2w32 + 16, m.ObjectToWorld.mInv.m, 16 * sizeof(float));
}

uint64_t TriangleMesh::CountSurfacePoints(float minDist) const {
    float o2w[32]; meshMatrices(*this, o2w);
    uint64_t count = 0;
    skc_tessellate_host(&p[0].x, n.empty() ? nullptr : &n[0].x, uvs.empty() ? nullptr : uvs.data(), vertexIndex.data(), (uint32_t)ntris, o2w,
                        (int)(ReverseOrientation ^ TransformSwapsHandedness), minDist, 0, 0, nullptr, 0, &count);
    return count;
}

void TriangleMesh::TessellateSurfacePoints(float minDist, uint32_t matId, std::vector<skc_surface_point> &points, bool incenter) const {
    static_assert(sizeof(V3) == 3 * sizeof(float), "V3 is three packed floats");
    float o2w[32]; meshMatrices(*this, o2w);
    const uint64_t count = CountSurfacePoints(minDist), base = points.size();
    points.resize(base + count);
    uint64_t written = 0;
    skc_tessellate_host(&p[0].x, n.empty() ? nullptr : &n[0].x, uvs.empty() ? nullptr : uvs.data(), vertexIndex.data(), (uint32_t)ntris, o2w,
                        (int)(ReverseOrientation ^ TransformSwapsHandedness), minDist, incenter ? 1 : 0, matId, points.data() + base, count, &written);
}

}  // namespace skin

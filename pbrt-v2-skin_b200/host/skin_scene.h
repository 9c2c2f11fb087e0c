// skin_scene.h -- host-side mirror of the reference's plugin classes on the subsurface path.
//
//   Scene description  -> what core/api.cpp builds from `pbrt scene.pbrt` for the skin scenes
//   TriangleMesh       -> shapes/trianglemesh.{h,cpp}: world-space vertices + the tessellator
//   BVHAccel           -> accelerators/bvh.{h,cpp}: SAH build, flat LinearBVHNode array
//   PerspectiveCamera  -> cameras/perspective.cpp
//   MultipoleSubsurfaceIntegrator -> integrators/multipolesubsurface.{h,cpp}: Preprocess / Li,
//                         both running on the device through include/skin_cuda.h
// Names, parameter names and defaults follow the reference so that its scene files map 1:1.
#pragma once
#include <string>
#include <vector>
#include <map>
#include "skin_math.h"
#include "../../include/skin_cuda.h"

namespace skin {

struct TriangleMesh {
    int ntris = 0, nverts = 0;
    std::vector<int> vertexIndex;
    std::vector<V3> p;       // WORLD space (shapes/trianglemesh.cpp:72-73)
    std::vector<V3> n;       // object space, may be empty
    std::vector<float> uvs;  // may be empty
    Transform ObjectToWorld;
    bool ReverseOrientation = false, TransformSwapsHandedness = false;
    uint32_t materialId = 0;
    // TriangleMesh::TessellateSurfacePoints (shapes/trianglemesh.cpp:187-257)
    void TessellateSurfacePoints(float minDist, uint32_t materialId, std::vector<skc_surface_point> &points, bool incenter) const;
    uint64_t CountSurfacePoints(float minDist) const;
};

struct SphereLight {      // Shape "sphere" + AreaLightSource "area" (shapes/sphere.cpp, lights/diffuse.cpp)
    Transform ObjectToWorld;
    float radius = 1.f, zmin = -1.f, zmax = 1.f, phiMaxDeg = 360.f;
    float Lrgb[3] = {1, 1, 1};
    int nSamples = 1;
    bool ReverseOrientation = false;
    skc_light Pack(const float *Lemit30) const;
    BBox WorldBound() const;
};

// LayeredSkin parameters (materials/layeredskin.cpp:222-262)
struct SkinParams {
    float roughness = 0.3f, nmperunit = 100e6f;
    float f_mel = 0.15f, f_eu = 1.f, f_blood = 0.002f, f_ohg = 0.3f;
    int desiredlength = 512;
    // "color Kr" / "color Kt" as the skin scenes give them (results/PigmentManips/Eum0.1_Deoxy0.5.pbrt:37-38); the scene-file
    // format issues these two like the reference driver does.  (The material's own defaults are Spectrum(1) for both.)
    float Kr[3] = {1, 1, 1}, Kt[3] = {0, 0, 0};
    bool lerponthinslab = true;
    float layerThickness[2] = {0.25e6f, 20e6f}, layerIor[2] = {1.4f, 1.4f};
};

struct SceneDesc {
    int xres = 320, yres = 180, spp = 4;
    float lookat[9] = {0, -5, 0, 0, 0, 0, 0, 0, 1};
    float fov = 30.f;
    // SurfaceIntegrator "multipolesubsurface" (multipolesubsurface.cpp:309-316)
    int maxdepth = 5;
    float maxerror = .05f, minsampledistance = .25f, mix = .5f;
    bool incenter = false;
    std::vector<SphereLight> lights;
    SkinParams skin;
    std::vector<TriangleMesh> meshes;
    std::string error;
    bool Load(const std::string &path);
};

// accelerators/bvh.{h,cpp}
class BVHAccel {
public:
    BVHAccel(const std::vector<BBox> &primBounds, uint32_t maxPrims = 4, const std::string &splitMethod = "sah");
    std::vector<skc_bvh_node> nodes;          // depth-first flat tree (bvh.cpp:362-380)
    std::vector<uint32_t> orderedPrims;       // primitive numbers in leaf order
private:
    struct BuildNode;
    struct PrimInfo { uint32_t primitiveNumber; V3 centroid; BBox bounds; };
    BuildNode *recursiveBuild(std::vector<PrimInfo> &buildData, uint32_t start, uint32_t end, uint32_t *totalNodes);
    uint32_t flatten(BuildNode *node, uint32_t *offset);
    uint32_t maxPrimsInNode;
    int splitMethod;
    std::vector<BuildNode *> arena;
};

// cameras/perspective.cpp + core/camera.cpp:83-102
class PerspectiveCamera {
public:
    PerspectiveCamera(const Transform &cam2world, float fov, int xres, int yres);
    void GenerateRay(float imageX, float imageY, skc_ray *ray) const;
    Transform CameraToWorld, RasterToCamera;
};

// Everything the device needs, built from a SceneDesc the way RenderOptions::MakeScene does
struct BuiltScene {
    std::vector<skc_prim> prims;              // BVH leaf order
    std::vector<skc_bvh_node> nodes;
    std::vector<skc_light> lights;
    const TriangleMesh *mesh = nullptr;       // single skin mesh supported on the device path
};
bool BuildScene(const SceneDesc &sd, const float *Lemit30PerLight, BuiltScene *out, std::string *err);

// RGB -> SampledSpectrum for an illuminant (core/spectrum.cpp FromRGB, SpectrumType illuminant)
void RGBIlluminantToSpectrum(const float rgb[3], float out30[30]);

// integrators/multipolesubsurface.{h,cpp}
class MultipoleSubsurfaceIntegrator {
public:
    MultipoleSubsurfaceIntegrator(int maxDepth, float maxError, float minDist, float mix, bool incenter)
        : maxSpecularDepth(maxDepth), maxError(maxError), minSampleDist(minDist), mix(mix), incenter(incenter) {}
    // Preprocess (multipolesubsurface.cpp:170-238): surface points -> device irradiance pass ->
    // device octree.  nCoresForTasks reproduces nTasks = RoundUpPow2(max(32*cores, n/4096)).
    bool Preprocess(skc_ctx *ctx, const SceneDesc &sd, int nCoresForTasks, std::string *err);
    // Li, subsurface term, for a batch of camera rays (multipolesubsurface.cpp:253-304)
    bool LiBatch(skc_ctx *ctx, const skc_ray *rays, uint64_t n, skc_hit *hits, float *L30, std::string *err) const;
    std::vector<skc_surface_point> surfacePoints;
    uint32_t nTasks = 0, octreeNodes = 0;
    int maxSpecularDepth;
    float maxError, minSampleDist, mix;
    bool incenter;
};

inline uint32_t RoundUpPow2(uint32_t v) { v--; v |= v >> 1; v |= v >> 2; v |= v >> 4; v |= v >> 8; v |= v >> 16; return v + 1; }

}  // namespace skin

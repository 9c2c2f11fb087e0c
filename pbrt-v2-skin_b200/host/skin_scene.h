// skin_scene.h -- host-side mirror of the reference's plugin classes on the subsurface path.
//
//   Scene description  -> what core/api.cpp builds from `pbrt scene.pbrt` for the skin scenes
//   TriangleMesh       -> shapes/trianglemesh.{h,cpp}: world-space vertices + the tessellator
//   BVHAccel           -> accelerators/bvh.{h,cpp}: the same flat LinearBVHNode array and leaf order, built by an iterative
//                         builder of this repo's own (host/skin_geom.cpp)
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

// The mesh's "texture alpha": a float image map (textures/imagemap.cpp CreateImageFloatTexture) as its MIPMap's level-0 texels and its
// UVMapping2D; the triangle tests ignore a hit whose texture value is exactly 0 (shapes/trianglemesh.inl:117-126,174-207)
struct AlphaTexture {
    bool present = false;
    int w = 0, h = 0;
    std::vector<float> texels;                  // [h][w]: scale * (powf(luminance, gamma) + shift)
    float su = 1.f, sv = 1.f, du = 0.f, dv = 0.f;
    bool operator==(const AlphaTexture &o) const { return present == o.present && w == o.w && h == o.h && texels == o.texels && su == o.su && sv == o.sv && du == o.du && dv == o.dv; }
};
// ImageTexture<float, float> of a PFM file (power-of-two sizes), textures/imagemap.cpp:56-86 + imagemap.h:94-97
bool LoadAlphaTexture(const std::string &file, float shift, float scale, float gamma, AlphaTexture *out, std::string *err);

struct TriangleMesh {
    int ntris = 0, nverts = 0;
    AlphaTexture alpha;
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

// A light of scene->lights, in scene order.  kind 0: Shape "sphere" + AreaLightSource "area" (shapes/sphere.cpp,
// lights/diffuse.cpp); kind 1 / 2 / 3: LightSource "point" / "spot" / "distant" (lights/point.cpp, spot.cpp, distant.cpp) --
// delta lights, no shape in the BVH; kind 4: LightSource "infinite" (lights/infinite.cpp), whose map and sampling distribution
// live in SceneDesc::env.  ObjectToWorld is the shape's / the light's LightToWorld.
struct SphereLight {
    int kind = 0;
    int envW = 0, envH = 0;                              // infinite: resolution of the radiance map
    float cosTotalWidth = 0.f, cosFalloffStart = 0.f;    // spot (lights/spot.cpp:42-50)
    V3 lightDir;                                        // distant: Normalize(LightToWorld(from - to)) (lights/distant.cpp:42-47)
    Transform ObjectToWorld;
    float radius = 1.f, zmin = -1.f, zmax = 1.f, phiMaxDeg = 360.f;
    float Lrgb[3] = {1, 1, 1};
    int nSamples = 1;
    bool ReverseOrientation = false;
    skc_light Pack(const float *Lemit30) const;
    BBox WorldBound() const;
};

// InfiniteAreaLight's radiance map and Distribution2D as flat arrays (host/skin_envmap.cpp; device: skc_upload_envmap)
struct EnvMap {
    bool present = false;
    int w = 0, h = 0;
    std::vector<float> rgb;                                  // [h][w][3]: level-0 texels (map * L.ToRGBSpectrum())
    std::vector<float> condFunc, condCdf, condInt;           // [h][w], [h][w + 1], [h]
    std::vector<float> margFunc, margCdf;                    // [h], [h + 1]
    float margInt = 0.f;
};

// LayeredSkin's "texture albedo": a "color" image map (textures/imagemap.cpp CreateImageSpectrumTexture) as its MIPMap's whole pyramid
// (the EWA filter reads every level), the filter's weight table and its UVMapping2D
struct AlbedoTexture {
    bool present = false;
    std::vector<int> w, h;                      // per level
    int channels = 3;                           // 3: a colour map; 1: a float map (the bump map's displacement)
    std::vector<float> rgb;                     // the levels one after the other, [h][w][channels] each
    float weightLut[128] = {0};                 // core/mipmap.h:186-193
    float su = 1.f, sv = 1.f, du = 0.f, dv = 0.f;
    bool operator==(const AlbedoTexture &o) const { return present == o.present && channels == o.channels && w == o.w && h == o.h && rgb == o.rgb && su == o.su && sv == o.sv && du == o.du && dv == o.dv; }
};
bool LoadAlbedoTexture(const std::string &file, float shift, float scale, float gamma, AlbedoTexture *out, std::string *err);
// ImageTexture<float, float> with its whole MIPMap<float> (channels = 1): the skin's "texture bumpmap"
bool LoadBumpTexture(const std::string &file, float shift, float scale, float gamma, AlbedoTexture *out, std::string *err);

// LayeredSkin parameters (materials/layeredskin.cpp:222-262)
struct SkinParams {
    float roughness = 0.4f, nmperunit = 100e6f;      // materials/layeredskin.cpp:234-235
    float f_mel = 0.15f, f_eu = 1.f, f_blood = 0.002f, f_ohg = 0.3f;
    int desiredlength = 512;
    // "color Kr" / "color Kt" as the skin scenes give them (results/PigmentManips/Eum0.1_Deoxy0.5.pbrt:37-38); the scene-file
    // format issues these two like the reference driver does.  (The material's own defaults are Spectrum(1) for both.)
    float Kr[3] = {1, 1, 1}, Kt[3] = {0, 0, 0};
    // "color albedo" (layeredskin.cpp:247; default Spectrum(1)): a constant; E is scaled by Pow(albedo, mix), Li's subsurface term by
    // Pow(albedo, 1 - mix) (multipolesubsurface.cpp:139-140,287)
    float albedo[3] = {1, 1, 1};
    bool hasAlbedo = false;
    AlbedoTexture albedoTex;                    // "texture albedo" (an image map); excludes hasAlbedo
    AlbedoTexture bumpTex;                      // "texture bumpmap" (a float image map, channels = 1)
    bool lerponthinslab = true;
    float layerThickness[2] = {0.25e6f, 20e6f}, layerIor[2] = {1.4f, 1.4f};
};

// Pixel filter of Film "image" (filters/{box,gaussian,mitchell,sinc,triangle}.cpp; api.cpp:648-672 default "box")
struct PixelFilter {
    std::string name = "box";
    float xwidth = .5f, ywidth = .5f;
    float alpha = 2.f, B = 1.f / 3.f, C = 1.f / 3.f, tau = 3.f;
    float Evaluate(float x, float y) const;
    void Table(float out256[256]) const;      // ImageFilm ctor, film/image.cpp:55-68
};

// Command-line options that reach the plugin factories (core/pbrt.h Options, main/pbrt.cpp:76-93)
struct LoadOptions {
    bool quickRender = false;       // --quick
    std::string imageFile;          // --outfile
};

struct PrimRef { int kind; int index; };      // kind 0 = meshes[index], 1 = lights[index]; order of renderOptions->primitives

struct PerspectiveCameraDesc { Transform CameraToWorld, RasterToCamera; };

struct SceneDesc {
    int xres = 320, yres = 180, spp = 4;
    float lookat[9] = {0, -5, 0, 0, 0, 0, 0, 0, 1};
    float fov = 30.f;
    // Camera "perspective": world-to-camera = the CTM at the Camera directive (api.cpp:858-864); screen window
    // (cameras/perspective.cpp:128-147).  The .scene line format sets it from `lookat`.
    Transform WorldToCamera;
    bool haveScreenWindow = false;
    float screenWindow[4] = {-1, 1, -1, 1};
    float frameAspect = 0.f;        // 0: xres / yres
    PixelFilter filter;
    std::string filmFilename;       // Film "image" "string filename" / --outfile; "" -> pbrt.tga (film/image.cpp:222-238 without OpenEXR)
    std::string splitMethod = "sah";  // Accelerator "bvh" (accelerators/bvh.cpp:549-550)
    int maxNodePrims = 4;
    std::vector<PrimRef> primOrder;   // empty: emitters first, then the meshes (the .scene format's order)
    std::vector<std::string> warnings;
    // SurfaceIntegrator "multipolesubsurface" (multipolesubsurface.cpp:309-316)
    int maxdepth = 5;
    float maxerror = .05f, minsampledistance = .25f, mix = .5f;
    bool incenter = false;
    int poissonMaxFails = 2000;           // surfacepoints.cpp:134-135: 2000, max(10, 2000 / 10) under --quick
    bool usePoissonPointFinder = false;   // "bool usepoissonpointfinder": FindPoissonPointDistribution instead of the tessellator (skc_poisson_resident)
    std::string pointsFile;           // "string pointsfile": raw SurfacePoint[44 B] records (multipolesubsurface.cpp:176-181, core/floatfile.h:45-63)
    // Renderer: "sampler" renders the frame; "tessellatesurfacepoints" only writes the surface points of the scene's
    // translucent meshes to "string filename" (renderers/surfacepoints.cpp:335-347,381-388) -- the file `pointsfile` reads
    std::string rendererName = "sampler";
    float rendererMinDist = .25f;
    std::string rendererFilename;
    bool rendererVersionedFile = false;   // "bool versioned": write the tagged points file (WritePointsFile) instead of the reference's raw dump
    std::vector<SphereLight> lights;
    EnvMap env;                       // of the scene's infinite light (at most one)
    SkinParams skin;
    std::vector<TriangleMesh> meshes;
    std::string error;
    // Load: `.pbrt` files go through the pbrt scene-description front end (skin_pbrt.cpp: the directives and
    // parameter lists of core/pbrtlex.ll + pbrtparse.yy + api.cpp), anything else is the line format of scenes/*.scene.
    bool Load(const std::string &path, const LoadOptions *opt = nullptr);
    bool LoadPbrt(const std::string &path, const LoadOptions *opt);
    bool LoadLines(const std::string &path);
    PerspectiveCameraDesc Camera() const;
};

// accelerators/bvh.{h,cpp}
class BVHAccel {
public:
    BVHAccel(const std::vector<BBox> &primBounds, uint32_t maxPrims = 4, const std::string &splitMethod = "sah");
    std::vector<skc_bvh_node> nodes;          // depth-first flat tree (bvh.cpp:362-380)
    std::vector<uint32_t> orderedPrims;       // primitive numbers in leaf order
private:
    uint32_t maxPrimsInNode;
    int splitMethod;                          // 0 middle, 1 equal counts, 2 surface-area heuristic
};

// cameras/perspective.cpp + core/camera.cpp:83-102
class PerspectiveCamera {
public:
    PerspectiveCamera(const Transform &cam2world, float fov, int xres, int yres, float frameAspect = 0.f, const float *screenWindow = nullptr);
    void GenerateRay(float imageX, float imageY, skc_ray *ray) const;
    Transform CameraToWorld, RasterToCamera;
};

// Everything the device needs, built from a SceneDesc the way RenderOptions::MakeScene does
struct BuiltScene {
    std::vector<skc_prim> prims;              // BVH leaf order
    std::vector<skc_bvh_node> nodes;
    std::vector<skc_light> lights;
    TriangleMesh merged;                      // several skin meshes joined into one (BuildScene)
    const TriangleMesh *mesh = nullptr;       // the one mesh the device holds: the scene's only mesh, or `merged`
};
// CreatePointLight / CreateSpotLight / CreateDistantLight (lights/point.cpp:63-70, spot.cpp:81-101, distant.cpp:67-75): l2w is
// the CTM at the LightSource directive
SphereLight MakePointLight(const Transform &l2w, V3 from, const float rgb[3]);
SphereLight MakeSpotLight(const Transform &l2w, V3 from, V3 to, const float rgb[3], float coneangle, float conedelta);
SphereLight MakeDistantLight(const Transform &l2w, V3 from, V3 to, const float rgb[3]);
// CreateInfiniteLight + InfiniteAreaLight's constructor (lights/infinite.cpp:67-103,185-193); scaleRgb may be null (Spectrum(1))
bool MakeInfiniteLight(const Transform &l2w, const float Lrgb[3], const float *scaleRgb, int nSamples, const std::string &mapname,
                       bool quickRender, SphereLight *light, EnvMap *env, std::string *err);
void SpectrumToRGB(const float c30[30], float rgb[3]);
bool BuildScene(const SceneDesc &sd, const float *Lemit30PerLight, BuiltScene *out, std::string *err);
// TriangleMesh from object-space arrays the way CreateTriangleMeshShape + the TriangleMesh ctor do (shapes/trianglemesh.cpp:55-80,360-420)
void InitMesh(TriangleMesh *m, const Transform &o2w, bool reverseOrientation, int nverts, const float *P, const float *N, const float *uv,
              int ntris, const int *indices);

// What api.cpp's MakeCamera / MakeFilm / MakeSampler / MakeRenderer derive from the scene for the device frame pass
void MakeFrameDesc(const SceneDesc &sd, int nCores, skc_frame_desc *d);
int BuildSkinMaterial(skc_ctx *ctx, const SkinParams &skin, skc_comm *comm = nullptr);
// the two albedo factors of the material for an integrator's `mix` -> skc_set_albedo_factors (cleared when the albedo is the default)
int InstallAlbedo(skc_ctx *ctx, const SkinParams &skin, float mix);
int UploadEnvMap(skc_ctx *ctx, const EnvMap &env);
int UploadAlpha(skc_ctx *ctx, const AlphaTexture &a);
// The `pointsfile` of MultipoleSubsurfaceIntegrator / Renderer "tessellatesurfacepoints" (multipolesubsurface.cpp:176-181,
// surfacepoints.cpp:335-347).  The reference's file is a raw dump of SurfacePoint[44 B] in host byte order with no
// header; it stays the default in both directions.  The tagged variant adds 32 bytes in front -- "SKINPTS\1", version, record size,
// count, a byte-order mark -- so that a reader can tell a points file from anything else, a truncated one from a whole one, and a
// foreign byte order from its own.  ReadPointsFile takes either (the tag decides); a raw file's trailing partial record is ignored
// as ReadBinaryFile does (core/floatfile.h:45-63).
bool ReadPointsFile(const std::string &path, std::vector<skc_surface_point> *pts, std::string *err);
bool WritePointsFile(const std::string &path, const skc_surface_point *pts, uint64_t n, bool versioned, std::string *err);
int UploadAlbedoTexture(skc_ctx *ctx, const AlbedoTexture &a);
int UploadBumpTexture(skc_ctx *ctx, const AlbedoTexture &a);     // skc_upload_bump_texture (or its removal)
   // skc_upload_albedo_texture (or its removal)    // skc_upload_alpha (or its removal)     // skc_upload_envmap (or its removal when the scene has no infinite light)   // comm: the wavelengths dealt over its ranks
int UploadScene(skc_ctx *ctx, const BuiltScene &b);

// SamplerRenderer::Render (renderers/samplerrenderer.cpp:191-225) on the device: scene + material upload, Preprocess,
// camera pass, film -> RGB [yres][xres][3].  nCores only fixes the task counts that seed the random streams.
struct RenderStats {
    double buildSec = 0, materialSec = 0, preprocessSec = 0, renderSec = 0;
    uint64_t nPoints = 0, nSamples = 0;
    uint32_t octreeNodes = 0, irradianceTasks = 0, renderTasks = 0;
};
bool RenderScene(skc_ctx *ctx, const SceneDesc &sd, int nCores, std::vector<float> *rgb, RenderStats *st, std::string *err);
// one rank's share of the frame (comm == nullptr: all of it); the image goes to rank 0 (host/skin_render.cpp)
bool RenderSceneRank(skc_ctx *ctx, skc_comm *comm, const SceneDesc &sd, const BuiltScene &built, int nCores, std::vector<float> *rgb, RenderStats *st, std::string *err);
// one process driving nGpus devices (`pbrt --gpus N`): a host thread and a context per device, NCCL between them
bool RenderSceneMultiGPU(int firstDevice, int nGpus, const SceneDesc &sd, int nCores, std::vector<float> *rgb, RenderStats *st, std::string *err);
// Renderer "tessellatesurfacepoints" (renderers/surfacepoints.cpp:300-347): device tessellation -> raw SurfacePoint file
bool TessellateSurfacePointsToFile(skc_ctx *ctx, const SceneDesc &sd, uint64_t *nPoints, std::string *err);
// ::WriteImage (core/imageio.cpp:78-104): .pfm / .tga by suffix (no OpenEXR in this build)
bool WriteImage(const std::string &name, const float *rgb, int w, int h, std::string *err);

// RGB -> SampledSpectrum for an illuminant (core/spectrum.cpp FromRGB, SpectrumType illuminant)
void RGBIlluminantToSpectrum(const float rgb[3], float out30[30]);
// Spectrum::FromRGB(rgb, SPECTRUM_REFLECTANCE) (core/spectrum.cpp:103-145): what ParamSet turns every "color" parameter into
void RGBToSpectrum(const float rgb[3], float out30[30]);

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

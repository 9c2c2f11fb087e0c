/* skin_cuda.h -- C ABI of the B200 device layer for pbrt-v2-skin's subsurface hot path.
 *
 * This is the thin `extern "C"` boundary BASELINE.json's north_star asks for (the
 * reference tree would hold it as src/cuda/skin_cuda.h).  Plain pointers and sizes only;
 * every array is caller-owned HOST memory unless the name ends in `_dev`.  All calls are
 * stream-ordered on the context's stream and return 0 on success; on failure they return
 * non-zero and skc_last_error() describes why (the plugin maps that to pbrt's Severe()).
 * There is no CPU fallback anywhere behind this header.
 *
 * Each entry point names the reference code (Thomi6York/pbrt-v2-skin, paths under src/)
 * whose work it replaces.
 */
#ifndef SKIN_CUDA_H
#define SKIN_CUDA_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SKC_NSPECTRAL 30 /* core/spectrum.h:46 nSpectralSamples; Spectrum == SampledSpectrum (core/pbrt.h:160) */
#define SKC_RHO_LEN 1025 /* core/multipole.cpp:521-549 rho_hd table length */

typedef struct skc_ctx skc_ctx;

/* accelerators/bvh.cpp:113-123 LinearBVHNode, byte-compatible (32 B). */
typedef struct {
    float pmin[3], pmax[3];
    uint32_t offset;      /* leaf: primitivesOffset, interior: secondChildOffset */
    uint8_t nPrimitives;  /* 0 -> interior */
    uint8_t axis;
    uint8_t pad[2];
} skc_bvh_node;

/* One entry of BVHAccel::primitives (accelerators/bvh.h:66): what the leaf's virtual
 * Primitive::Intersect resolves to.  type 0 = triangle `index` of the (world-space) mesh,
 * type 1 = quadric sphere `index` (the emitter shapes: SURVEY.md appendix A1). */
typedef struct {
    int32_t type, mesh, index, primitiveId;
} skc_prim;

/* One light of scene->lights as a 70-float record.
 *   radius > 0: shapes/sphere.h Sphere + lights/diffuse.h DiffuseAreaLight (the emitter sphere is also a BVH primitive).
 *   radius < 0: a delta light without a shape -- radius = -1 PointLight (lights/point.cpp), -2 SpotLight
 *     (lights/spot.cpp), -3 DistantLight (lights/distant.cpp); o2w / w2o = LightToWorld / WorldToLight; spot: zmin =
 *     cosTotalWidth, zmax = cosFalloffStart; distant: (zmin, zmax, phiMax) = lightDir (world space, normalised); Lemit =
 *     Intensity (point, spot) or L (distant); nSamples = 1; the other fields 0.  radius = -4: InfiniteAreaLight (lights/infinite.cpp),
 *     not a delta light: zmin / zmax = width / height of its radiance map (skc_upload_envmap), nSamples; Lemit unused. */
typedef struct {
    float o2w[16], w2o[16];            /* ObjectToWorld matrix and its inverse, row-major */
    float radius, zmin, zmax, phiMax;
    float nSamples, reverseOrientation, thetaMin, thetaMax;
    float Lemit[SKC_NSPECTRAL];
} skc_light;

/* core/geometry.h RayBase without time/depth (32 B). */
typedef struct {
    float o[3], d[3], mint, maxt;
} skc_ray;

/* Closest-hit record: what Intersection carries that the path consumes (56 B).
 * p,n = bsdf->dgShading.{p,nn} (integrators/multipolesubsurface.cpp:263-265), wo = -ray.d. */
typedef struct {
    int32_t hit, prim;   /* prim = index into the uploaded skc_prim array, -1 on miss */
    float t, rayEpsilon;
    float p[3], n[3], wo[3];
    int32_t hasBssrdf;   /* hit primitive is a subsurface (LayeredSkin) triangle */
} skc_hit;

/* renderers/surfacepoints.h:45-55 SurfacePoint, byte-compatible (44 B). */
typedef struct {
    float p[3], n[3];
    float u, v;
    uint32_t materialId;
    float area, rayEpsilon;
} skc_surface_point;

/* integrators/irradiancepoint.h:36-45 IrradiancePoint, byte-compatible (152 B). */
typedef struct {
    float p[3], n[3];
    float E[SKC_NSPECTRAL];
    float area, rayEpsilon;
} skc_irr_point;

/* multipole/MultipoleProfileCalculator/MultipoleProfileCalculator.h:34-67 PODs. */
typedef struct { float ior, thickness, mua, musp, g_HG; } skc_layer_spec;     /* MPC_LayerSpec */
typedef struct { float desiredStepSize; uint32_t desiredLength; int32_t lerpOnThinSlab; } skc_mpc_options; /* MPC_Options */

/* ---- context ----------------------------------------------------------------------- */
/* One context per process/GPU.  device < 0 selects the current CUDA device. */
skc_ctx *skc_create(int device);
void skc_destroy(skc_ctx *ctx);
const char *skc_last_error(const skc_ctx *ctx);   /* ctx may be NULL: last create error */
int skc_synchronize(skc_ctx *ctx);
/* Counters of kernels launched by this context since creation (bench.py gpu_launches). */
uint64_t skc_launch_count(const skc_ctx *ctx);
/* Streaming multiprocessors of the context's device (launch shapes are multiples of it). */
int skc_sm_count(const skc_ctx *ctx);
/* Device time of the last call of each stage in ms (CUDA events on the ctx stream).
 * stage: 0 scramble, 1 irradiance, 2 octree build, 3 closest hit, 4 Mo, 5 any-hit, 6 profile, 7 rho table, 8 tessellation,
 * 9 sampler stream + camera rays, 10 direct lighting, 11 spectral sum + film */
float skc_stage_ms(const skc_ctx *ctx, int stage);

/* CUDA-event timer on the context's stream spanning any sequence of calls (bench.py). */
int skc_timer_begin(skc_ctx *ctx);
float skc_timer_end_ms(skc_ctx *ctx);

/* Test hook: the device's sinf / cosf on n host floats.  The reference calls libm's sinf/cosf in UniformSampleCone
 * (core/montecarlo.cpp) and Beckmann::Sample_f (core/reflection.cpp); the device restates glibc's algorithm so that the results
 * agree bit for bit (csrc/skc_common.cuh). */
int skc_test_sincosf(skc_ctx *ctx, const float *x, uint64_t n, float *sinOut, float *cosOut);
/* ... and any one of them: fn 0 sinf, 1 cosf, 2 logf, 3 expf, 4 atanf (Beckmann::Sample_f / D, core/reflection.cpp:548-569). */
int skc_test_libmf(skc_ctx *ctx, int fn, const float *x, uint64_t n, float *out);

/* ---- BVHAccel (accelerators/bvh.cpp) ------------------------------------------------ */
/* Upload the flat tree BVHAccel::flattenBVHTree (bvh.cpp:362-380) produced, the ordered
 * primitive list, the world-space triangle meshes (shapes/trianglemesh.cpp:72-73) and the
 * quadric spheres.  Vertex normals `N` (object space) and `o2w` feed GetShadingGeometry
 * (shapes/trianglemesh.inl:224-299); N/uv may be NULL. */
int skc_upload_bvh(skc_ctx *ctx, const skc_bvh_node *nodes, uint32_t nNodes,
                   const skc_prim *prims, uint32_t nPrims,
                   const float *P, const float *N, const float *uv, uint32_t nVerts,
                   const int32_t *indices, uint32_t nTris, const float *meshO2W /*[32] m,minv*/,
                   int reverseOrientation, int transformSwapsHandedness,
                   const skc_light *spheres, uint32_t nSpheres);
/* BVHAccel::IntersectP (bvh.cpp:442-488): out[i] = 1 if occluded. */
int skc_trace_any(skc_ctx *ctx, const skc_ray *rays, uint64_t n, uint8_t *out);
/* BVHAccel::Intersect (bvh.cpp:388-439) + GeometricPrimitive::Intersect
 * (core/primitive.cpp:180-193) + GetShadingGeometry for the shading point/normal. */
int skc_trace_closest(skc_ctx *ctx, const skc_ray *rays, uint64_t n, skc_hit *out);

/* ---- material tables (core/multipole.cpp) ------------------------------------------- */
/* Rd table of MultipoleProfileData (core/multipole.cpp:95-113): table[lambda][len] plus the
 * per-wavelength rcpDsqSpacing (core/multipole.cpp:274-275), and the rho_hd table
 * rhoHd[1025][30] (core/multipole.cpp:458-463). */
int skc_upload_profile(skc_ctx *ctx, const float *table, const float *rcpDsqSpacing,
                       uint32_t len, const float *rhoHd);

/* ---- MultipoleSubsurfaceIntegrator::Preprocess (integrators/multipolesubsurface.cpp) - */
/* Lights sampled by the irradiance loop (lights/diffuse.cpp:69-81). */
int skc_upload_lights(skc_ctx *ctx, const skc_light *lights, uint32_t nLights);
/* IrradianceTask::Run (multipolesubsurface.cpp:72-152) for points [first, first+count) of
 * the nTotal-point set cut into nTasks slices (:73-74, RNG(taskNum*47) :81).  albedo30 is
 * the material's albedo texture value per point ([count][30]) or NULL for constant 1.
 * Results stay on the device (appended to the context's irradiance-point set at index
 * `first`); `out` may be NULL, else receives count records. */
int skc_irradiance(skc_ctx *ctx, const skc_surface_point *points, uint64_t nTotal,
                   uint64_t first, uint64_t count, uint32_t nTasks, float mix,
                   const float *albedo30, skc_irr_point *out);
/* Replace the device irradiance-point set (e.g. the reference's exported points, or the
 * all-gathered shards of the other GPUs). */
int skc_set_irradiance_points(skc_ctx *ctx, const skc_irr_point *pts, uint64_t n);
/* Device-resident packed record (p, area, E[30]) range access for the multi-GPU exchange:
 * returns device pointers to the float4 {p,area} array and the [n][32] E array. */
int skc_irradiance_dev(skc_ctx *ctx, uint64_t nTotal, void **posArea_dev, void **E_dev);
/* Rows first, first+stride, ... (count of them) of the device irradiance-point set back to the host: posArea4 [count][4]
 * = IrradiancePoint::p, area and E30 [count][30] = IrradiancePoint::E (integrators/irradiancepoint.h:36-45); either may be
 * NULL.  For parity checks of point sets too large to copy out whole. */
int skc_get_irradiance_rows(skc_ctx *ctx, uint64_t first, uint64_t count, uint64_t stride, float *posArea4, float *E30);
/* ... and the other way: overwrite E30 [count][30] (and posArea4 [count][4] unless NULL) of device points first .. first+count-1. */
int skc_set_irradiance_rows(skc_ctx *ctx, uint64_t first, uint64_t count, const float *posArea4, const float *E30);
/* After an exchange into arrays sized for padded shards: the number of valid points. */
int skc_set_point_count(skc_ctx *ctx, uint64_t n);
/* SubsurfaceOctreeNode::Insert + InitHierarchy (integrators/diffusionutil.h:94-173) over
 * the device irradiance-point set; the flat tree stays on the device.  Optional outputs:
 * node count, root {p[3], sumArea, Et[30]} (34 floats), bounds (6 floats). */
int skc_build_octree(skc_ctx *ctx, uint32_t *nNodesOut, float *root34, float *bounds6);
/* Surface points on the device: GetSurfacePointsThroughTessellation (renderers/surfacepoints.cpp:362-369)
 * -> TriangleMesh::TessellateSurfacePoints (shapes/trianglemesh.cpp:187-351) for the mesh given to skc_upload_bvh, same
 * points in the same order as the reference (centroid or incentre of every sub-triangle of the ring tessellation with
 * edge factors round(0.8 * edge / minSampleDist)).  The result becomes the resident surface-point set (as after
 * skc_set_surface_points); skc_get_surface_points copies a range of it to the host. */
int skc_tessellate_resident(skc_ctx *ctx, float minSampleDist, int incenter, uint32_t materialId, uint64_t *countOut);
/* The same for one shard of the set (several GPUs): only the points [first, first+count) -- rounded out to whole
 * triangles -- are written; *countOut is still the size of the whole set.  The rest of the resident array is left as it is. */
int skc_tessellate_range(skc_ctx *ctx, float minSampleDist, int incenter, uint32_t materialId, uint64_t first, uint64_t count,
                         uint64_t *countOut);
int skc_get_surface_points(skc_ctx *ctx, uint64_t first, uint64_t count, skc_surface_point *out);
/* The tessellator's job functions evaluated on the HOST (no context, no GPU): what the host mirror's
 * TriangleMesh::TessellateSurfacePoints / CountSurfacePoints call, so host and device share ONE implementation (and the CPU
 * tests that pin the host mirror to the reference's points pin the device kernel's logic too).  Not a fallback: the render
 * path tessellates with the kernel.  P world-space points, N object-space normals or NULL, UV or NULL, idx [nTris][3],
 * o2w32 = {ObjectToWorld.m, .mInv} row-major (NULL = identity), flip = ReverseOrientation ^ TransformSwapsHandedness.
 * out == NULL counts only; returns non-zero when cap is too small or an argument is missing. */
int skc_tessellate_host(const float *P, const float *N, const float *UV, const int32_t *idx, uint32_t nTris, const float *o2w32, int flip,
                        float minDist, int incenter, uint32_t materialId, skc_surface_point *out, uint64_t cap, uint64_t *countOut);
/* The Poisson-disk surface-point finder, MultipoleSubsurfaceIntegrator's "bool usepoissonpointfinder" (renderers/surfacepoints.cpp:103-284,
 * FindPoissonPointDistribution :354-359), as the reference runs it with ONE SurfacePointTask (--ncores 1: RNG(0); with more tasks the
 * reference's result depends on thread timing): rounds of 20 000 random walks of 30 bounces from origin3 (the camera position) through
 * the uploaded scene and its bounding sphere on the device, the accept / reject chain (no accepted point within minSampleDist;
 * stop after maxFails rejections in a row: 2000, 200 under --quick) on the host.  The accepted points become the resident
 * surface-point set, as after skc_tessellate_resident (skc_get_surface_points reads them).  stats4 (optional): {points, paths
 * traced, rays traced, rounds}. */
int skc_poisson_resident(skc_ctx *ctx, const float *origin3, float minSampleDist, uint32_t materialId, int maxFails, uint64_t *countOut, uint64_t *stats4);

/* MultipoleSubsurfaceIntegrator::Preprocess (multipolesubsurface.cpp:170-238) as ONE call: the irradiance pass over
 * points [first, first+count) and, concurrently on a second stream, the octree topology of ALL nTotal points (it
 * needs only their positions).  Follow with skc_build_octree_finalize -- directly on one GPU; after the all-gather of
 * the other ranks' irradiance rows (skc_irradiance_dev) on several.  skc_preprocess takes host points (uploaded
 * batch by batch behind the kernels), skc_preprocess_resident the set given to skc_set_surface_points. */
int skc_preprocess(skc_ctx *ctx, const skc_surface_point *points, uint64_t nTotal, uint64_t first, uint64_t count,
                   uint32_t nTasks, float mix);
int skc_preprocess_resident(skc_ctx *ctx, uint64_t first, uint64_t count, uint32_t nTasks, float mix);
/* The two halves of skc_build_octree, for callers that can overlap them with an exchange: the topology pass reads
 * only the point positions (bounds, splitting by `p > pMid`, leaf order = insertion order; diffusionutil.h:94-132,
 * multipolesubsurface.cpp:218-237); finalize gathers the records into leaf order and runs InitHierarchy
 * (diffusionutil.h:133-173) and needs the irradiance rows. */
int skc_build_octree_topology(skc_ctx *ctx);
int skc_build_octree_finalize(skc_ctx *ctx, uint32_t *nNodes, float *root34, float *bounds6);

/* ---- MultipoleSubsurfaceIntegrator::Li, subsurface term ----------------------------- */
/* SubsurfaceOctreeNode::Mo (diffusionutil.h:175-210) for n points: Mo30[n][30]. */
int skc_mo(skc_ctx *ctx, const float *p3, uint64_t n, float maxError, float *Mo30);
/* Li's subsurface term (multipolesubsurface.cpp:279-287) for closest-hit records:
 * L[i] = clamp0(INV_PI * (1-rho_hd(cos)) * Mo * albedo^(1-mix)); albedo30 NULL = 1.
 * Mo30 and/or L30 may be NULL. */
int skc_li_sss(skc_ctx *ctx, const skc_hit *hits, uint64_t n, float maxError, float mix,
               const float *albedo30, float *Mo30, float *L30);
/* Whole camera pass on device-resident data: closest hit -> Mo -> subsurface radiance.
 * Only L30 (and optionally hits) crosses back to the host. */
int skc_camera_pass(skc_ctx *ctx, const skc_ray *rays, uint64_t n, float maxError, float mix,
                    skc_hit *hitsOut, float *L30);

/* The Mo kernel keeps a compact form of the Rd tables in shared memory (csrc/skc_mo_table.h: one abscissa for all
 * wavelengths, cubic interpolation, per-wavelength windows verified against sampleProfile at upload time; outside
 * a window the lookup is the reference's exact one).  SKC_MO_TABLE=exact in the environment keeps the exact tables
 * only.  stats8 = {rows, shared-memory bytes, coverage, largest verified relative error inside the windows, and
 * from the last counted skc_camera_pass_resident: lookups in range, from the shared-memory table heads, from the
 * global tables, 0}. */
/* LayeredSkin's "texture albedo" as an image map (materials/layeredskin.cpp:183,247; textures/imagemap.cpp): the nLevels levels of its
 * MIPMap<RGBSpectrum> (w[l] x h[l] texels each, rgb = the levels one after the other, [h][w][3]; values as ImageTexture holds them),
 * the EWA weight table (core/mipmap.h:186-193) and the UVMapping2D.  E of an irradiance point is scaled by Pow(albedo(u, v), mix)
 * (multipolesubsurface.cpp:139-140; zero differentials: bilinear lookup at level 0), the subsurface term of a camera hit by
 * Pow(albedo, 1 - mix) (:287) with the EWA lookup under the camera ray's differentials (skc_frame_* path only).  Overrides
 * skc_set_albedo_factors.  nLevels == 0 removes it. */
int skc_upload_albedo_texture(skc_ctx *ctx, uint32_t nLevels, const int *w, const int *h, const float *rgb, const float *weightLut128,
                              float su, float sv, float du, float dv);
/* LayeredSkin's "texture bumpmap" as a float image map (materials/layeredskin.cpp:139,299-304; core/material.cpp:47-104 Material::Bump,
 * this fork's central-difference variant): the nLevels levels of its MIPMap<float> (displacement [h][w] per level, finest first), the EWA
 * weight table and the UVMapping2D.  skc_tessellate_resident / _range then give every irradiance point the bumped normal
 * (shapes/trianglemesh.cpp:236-242: zero differentials, du = dv = .01, bilinear lookups), and skc_frame_render builds the BSDF of every
 * skin hit on the bumped shading normal and dpdu (differentials from the camera ray).  nLevels == 0 removes it. */
int skc_upload_bump_texture(skc_ctx *ctx, uint32_t nLevels, const int *w, const int *h, const float *displacement, const float *weightLut128,
                            float su, float sv, float du, float dv);
/* Alpha-tested triangles (shapes/trianglemesh.inl:117-126,174-207): the mesh's "texture alpha", a float image map.  texels [h][w] are
 * the values ImageTexture<float, float> holds at MIPMap level 0 -- scale * (powf(luminance, gamma) + shift), made by the caller
 * (textures/imagemap.h:94-97); su, sv, du, dv its UVMapping2D.  Every ray-triangle hit, closest or any, whose texture value at the
 * hit's (u, v) is exactly 0 is ignored.  Call after skc_upload_bvh (and again after a new mesh); w == 0 removes it. */
int skc_upload_alpha(skc_ctx *ctx, uint32_t w, uint32_t h, const float *texels, float su, float sv, float du, float dv);
/* LightSource "infinite" (lights/infinite.cpp): its light record has radius -4, o2w / w2o = LightToWorld / WorldToLight, zmin / zmax =
 * the radiance map's width / height, nSamples; the map and its Distribution2D (core/montecarlo.h) come through this call, as
 * InfiniteAreaLight's constructor prepares them (the host mirror does: host/skin_envmap.cpp): rgb [h][w][3] level-0 texels,
 * condFunc [h][w], condCdf [h][w + 1], condInt [h] (the rows' Distribution1D), margFunc [h], margCdf [h + 1], margInt (the
 * marginal).  One infinite light per scene; w == 0 removes the map. */
int skc_upload_envmap(skc_ctx *ctx, uint32_t w, uint32_t h, const float *rgb, const float *condFunc, const float *condCdf,
                      const float *condInt, const float *margFunc, const float *margCdf, float margInt);
/* LayeredSkin's constant "color albedo" (materials/layeredskin.cpp:183,247; multipolesubsurface.cpp:139-140,287): the
 * irradiance pass scales E by Pow(albedo, mix), Li scales the subsurface term by Pow(albedo, 1 - mix).  The caller passes the two
 * 30-entry factors, made with the host's powf (the reference's own values); they apply to every later skc_irradiance* and
 * skc_camera_pass* / skc_frame_render / skc_li_sss call that is not given per-point albedos.  NULL, NULL = Spectrum(1). */
int skc_set_albedo_factors(skc_ctx *ctx, const float *powMix30, const float *powOneMinusMix30);
int skc_mo_table_stats(skc_ctx *ctx, double *stats8);
/* Contributions of the last counted skc_camera_pass_resident by the number of hits of their packet that accepted the same
 * cluster / leaf point together: clusters {<8, 8-15, 16-27, 28-32}, then leaf points {same four bins}. */
int skc_mo_group_histogram(skc_ctx *ctx, uint64_t *hist8);
/* Host-only (no GPU, no context): the builder and evaluator behind skc_upload_profile, for the error-bound test
 * (tests/test_mo_table.py).  skc_moc_build returns NULL when no wavelength gets a window. */
void *skc_moc_build(const float *table, const float *rcpDsqSpacing, uint32_t len, float tol, int K, int M, uint32_t headCapFloats);
double skc_moc_info(void *handle, float *info4x30 /* {fLo, fHi, headLen, maxErr} */, uint32_t *rowsOut, uint32_t *headFloatsOut);
void skc_moc_eval(void *handle, int lambda, const float *d2, uint64_t n, float *out, float *exactOut);
void skc_moc_free(void *handle);

/* skc_mo for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same kernels and results as
 * skc_mo, but all scratch (ticket counter, staged points, padded rows, output staging) lives in one file-scope
 * `static __device__` array of csrc/skc_mo.cu: 4 260 096 bytes, held by every GPU that loads the library until exit.  The
 * call makes no cudaMalloc / cudaFree and uses no allocating library routine; it needs the octree and the profile tables
 * already set up in ctx.  p3 [n][3] and Mo30 [n][30] may be host or device memory.  Limit: n <= SKC_MO_STATIC_MAX_HITS
 * points per call; a larger n is rejected with an error (split the call, or use skc_mo).  All calls share the array and it
 * keeps the last call's contents (what must start at zero is cleared per call); calls that overlap on the GPU -- two
 * contexts or two host threads on one device -- race on it. */
#define SKC_MO_STATIC_MAX_HITS 16384u
int skc_mo_static(skc_ctx *ctx, const float *p3, uint64_t n, float maxError, float *Mo30);

/* ---- resident inputs --------------------------------------------------------------------- */
/* The reference caches the surface-point set in a file ("string pointsfile",
 * multipolesubsurface.cpp:176-181); here the cache is HBM.  The *_resident calls run the same
 * kernels as skc_irradiance / skc_camera_pass on inputs uploaded once.  counters (optional,
 * NULL = none; costs atomics) report the work the kernels did, for the roofline arithmetic:
 *   skc_irradiance_resident   counters3 = {BVH slab tests, primitive tests, shadow rays}
 *   skc_camera_pass_resident  counters7 = {slab tests, primitive tests, octree nodes
 *                             classified, leaf points tested, cluster contributions,
 *                             point contributions, profile-table reads (one per contribution and
 *                             wavelength inside its table's range)} */
int skc_set_surface_points(skc_ctx *ctx, const skc_surface_point *points, uint64_t n);
int skc_irradiance_resident(skc_ctx *ctx, uint64_t first, uint64_t count, uint32_t nTasks, float mix,
                            uint64_t *counters3);
int skc_set_camera_rays(skc_ctx *ctx, const skc_ray *rays, uint64_t n);
int skc_camera_pass_resident(skc_ctx *ctx, uint64_t first, uint64_t count, float maxError, float mix,
                             float *L30, uint64_t *counters7);
/* Scene::Intersect (anyHit = 0; hitsOut = count skc_hit records or NULL) / Scene::IntersectP (anyHit = 1; hitsOut = count
 * bytes or NULL) over the resident rays [first, first+count): core/scene.h:51-68 -> accelerators/bvh.cpp:388-488.
 * counters2 (optional) = {slab tests, primitive tests}. */
int skc_trace_resident(skc_ctx *ctx, uint64_t first, uint64_t count, int anyHit, void *hitsOut, uint64_t *counters2);
/* Pinned host memory for callers that want full-rate PCIe copies through the host-buffer calls. */
void *skc_alloc_pinned(uint64_t bytes);
void skc_free_pinned(void *p);

/* ---- MPC_* profile generator (multipole/MultipoleProfileCalculator) ------------------ */
/* MPC_ComputeDiffusionProfile (+ optional MPC_ResampleForUniformDistanceSquaredDistribution)
 * for nSets layer stacks in one batch.  specs holds layersPerSet[s] records per set, packed.
 * Outputs: length (same for all sets with equal desiredLength), and per set
 * d2[len], R[len], T[len], totals[2].  Call once with out arrays NULL to get *lenOut. */
int skc_profile_batch(skc_ctx *ctx, const skc_layer_spec *specs, const uint32_t *layersPerSet,
                      const skc_mpc_options *opts, uint32_t nSets, int resample,
                      uint32_t *lenOut, float *d2, float *R, float *T, float *totals);

/* `resample` of the profile calculator (MultipoleProfileCalculator.cpp:355-402) at arbitrary squared distances: what
 * MPC_ResampleForUniformDistanceSquaredDistribution (:404-426) with any target length and MPC_ResampleDistribution (:428-448)
 * both reduce to.  d2 / R / T [len] host arrays of one profile (d2 ascending), queryDsq [nQueries]; Rout / Tout [nQueries]. */
int skc_profile_resample(skc_ctx *ctx, const float *d2, const float *R, const float *T, uint32_t len,
                         const float *queryDsq, uint32_t nQueries, float *Rout, float *Tout);

/* ---- rho_hd table ------------------------------------------------
 * Replaces ComputeRhoDataFromBxDF (core/multipole.cpp:518-549) for the Microfacet(Beckmann(roughness),
 * FresnelDielectric(etaI, etaT)) lobe LayeredSkin builds (materials/layeredskin.cpp:103-109):
 * 1025 entries, entry k = BxDF::rho(wo(cos = k/1024), 65536 stratified samples jittered by RNG(6428263*k)).
 * fixedFresnel != 0 selects FixedFresnelDielectric ("doublerefsslf").  hd30 receives [1025][30] floats
 * (RhoData::hd as Spectrum rows), the layout skc_upload_profile takes. */
int skc_rho_table(skc_ctx *ctx, float roughness, float etaI, float etaT, int fixedFresnel, float *hd30);

/* ---- whole frame on the device ------------------------------------------
 * The rest of SamplerRenderer::Render (renderers/samplerrenderer.cpp:59-167,191-225) around the subsurface term:
 *   LDSampler stream           samplers/lowdiscrepancy.cpp:58-79, core/montecarlo.cpp:200-252, core/montecarlo.h:183-331
 *   PerspectiveCamera rays     cameras/perspective.cpp:79-113, core/camera.cpp:83-102
 *   Li                         integrators/multipolesubsurface.cpp:253-304: Le + subsurface term + UniformSampleAllLights
 *                              (core/integrator.cpp:47-175) with the LayeredSkin BSDF (materials/layeredskin.cpp:139-167:
 *                              Microfacet + MicrofacetTransmission over Beckmann, core/reflection.cpp:228-296,391-470,548-581,
 *                              675-780) and multiple importance sampling; emitter spheres carry the default matte
 *   ImageFilm::AddSample       film/image.cpp:77-137 (filter table, XYZ accumulation)
 * The frame description is what api.cpp / the plugin constructors derive from the scene file. */
typedef struct {
    int32_t xres, yres;            /* Film "image" resolution, full crop window */
    int32_t spp;                   /* LDSampler pixelsamples, already a power of two (lowdiscrepancy.cpp:41-50) */
    int32_t nTasks;                /* RoundUpPow2(max(32*cores, nPixels/256)), samplerrenderer.cpp:206-208 */
    int32_t extra1D;               /* 1-sample 1-D arrays requested after the surface integrator's (EmissionIntegrator: 2) */
    float cameraToWorld[16];       /* row-major Transform::m */
    float rasterToCamera[16];
    float filterXWidth, filterYWidth;
    float filterTable[256];        /* ImageFilm::filterTable (16 x 16) */
    float cieXYZ[3][SKC_NSPECTRAL];/* SampledSpectrum::X, Y, Z (core/spectrum.cpp Init) */
    float xyzScale;                /* (sampledLambdaEnd - sampledLambdaStart) / (CIE_Y_integral * nSpectralSamples) */
    float roughness, ior;          /* LayeredSkin: roughness, layers[0].ior */
    float Kr[SKC_NSPECTRAL], Kt[SKC_NSPECTRAL];  /* LayeredSkin "color Kr" / "color Kt" as spectra; a black one drops its lobe */
    float matteKd;                 /* MatteMaterial Kd of the emitter spheres (default 0.5) */
    int32_t taskRank, taskWorld;   /* several GPUs: this context renders the tasks whose emission index % taskWorld == taskRank
                                    * (round-robin tiles); taskWorld <= 1 = all tasks */
} skc_frame_desc;
/* Allocates the film and the sample stream for the frame; *nSamplesOut = samples of the whole sample extent
 * (Film::GetSampleExtent, film/image.cpp:158-167) x spp, in SamplerRendererTask order (task nTasks-1 first). */
int skc_frame_setup(skc_ctx *ctx, const skc_frame_desc *desc, uint64_t *nSamplesOut);
/* Generates the LDSampler stream of every task (RNG(taskNum) regenerated per CTA) and the camera rays; they become
 * the resident camera rays (as after skc_set_camera_rays). */
int skc_frame_samples(skc_ctx *ctx);
/* The part of the frame that does not depend on Preprocess -- skc_frame_samples, Scene::Intersect of every camera ray and
 * EstimateDirect's terms (core/integrator.cpp:108-155) -- enqueued on the context's second stream; returns without
 * waiting.  Call it before the Preprocess calls; skc_frame_render(ctx, 0, nSamples, ...) then picks the results up on
 * the device.  Replaces skc_frame_samples in that flow; results are bit-identical to the staged calls.  One difference a
 * caller of skc_frame_get_samples can see: on this path the integrator's sample arrays are generated only for pixels with at
 * least one camera ray that hits something (nothing reads the others; image samples and rays are complete). */
int skc_frame_prefetch(skc_ctx *ctx);
/* Test access: image positions [count][2], the 1-D/2-D arrays of each sample [count][floatsPerSample]
 * (every 1-D array, then every 2-D array, as Sample::oneD / twoD) and the rays.  Any output may be NULL. */
int skc_frame_get_samples(skc_ctx *ctx, uint64_t first, uint64_t count, float *xy, float *arrays, skc_ray *rays);
uint32_t skc_frame_sample_floats(skc_ctx *ctx);
/* Li for samples [first, first+count) of the RENDER ORDER and ImageFilm::AddSample of each into the device film.
 * The render order visits the pixels of every task window along a Z-curve (compact Mo packets); skc_frame_get_order
 * gives, for render-order positions [first, first+count), the index of that sample in the sampler stream.  L30
 * (optional, host) receives the per-sample radiance [count][30] in render order. */
int skc_frame_get_order(skc_ctx *ctx, uint64_t first, uint64_t count, uint64_t *streamIndex);
int skc_frame_render(skc_ctx *ctx, uint64_t first, uint64_t count, float maxError, float mix, float *L30);
/* Film pixels [yres][xres]{Lxyz[3], weightSum}: clear, device pointer (for the reduce to rank 0), copy out. */
int skc_frame_film_clear(skc_ctx *ctx);
void *skc_frame_film_dev(skc_ctx *ctx);
int skc_frame_film_get(skc_ctx *ctx, float *pixels4);
/* The image ImageFilm::WriteImage hands to ::WriteImage (film/image.cpp:178-213): XYZ -> RGB, divided by the weight sum,
 * clamped at 0; rgb3 = [yres][xres][3] on the host. */
int skc_frame_film_rgb(skc_ctx *ctx, float *rgb3);

/* ---- several GPUs (no counterpart in the single-process reference, core/parallel.cpp:722-748) ----------------
 * The path shards in two places and each has one exchange step, both NCCL collectives over NVLink inside this library:
 *   irradiance points  every rank computes a contiguous range (skc_tessellate_range + skc_irradiance_resident(first, count)) of
 *                      arrays sized for world x shard points; ncclAllGather in place; every rank then builds the same octree
 *   camera tiles       every rank renders its render tasks (skc_frame_desc.taskRank / taskWorld) into its own film; ncclReduce
 *                      (sum) to the root, because the reconstruction filter spreads samples across tile borders
 * One communicator per context.  Bootstrap: one process per GPU -- rank 0 calls skc_comm_unique_id, the 128 bytes reach the
 * other ranks by whatever the launcher offers, every rank calls skc_comm_create; or one process with several contexts --
 * skc_comm_create_local, then one host thread per context.  All calls are enqueued on the context's streams. */
typedef struct skc_comm skc_comm;
int skc_comm_unique_id(uint8_t *id128);
skc_comm *skc_comm_create(skc_ctx *ctx, int rank, int world, const uint8_t *id128);
int skc_comm_create_local(skc_ctx **ctxs, int n, skc_comm **comms);
void skc_comm_destroy(skc_comm *comm);
int skc_comm_rank(const skc_comm *comm);
int skc_comm_world(const skc_comm *comm);
/* {p, area} of all shards, in place, on the main stream (the octree topology pass that follows needs only these). */
int skc_comm_allgather_positions(skc_comm *comm, uint64_t shardPoints);
/* Irradiance rows of all shards (8x the bytes), in place, on the copy stream behind the work enqueued so far; returns at once.
 * skc_comm_allgather_rows_wait makes the main stream wait for them on the device -- call it before skc_build_octree_finalize. */
int skc_comm_allgather_rows_begin(skc_comm *comm, uint64_t shardPoints);
int skc_comm_allgather_rows_wait(skc_comm *comm);
/* Sum of all ranks' films into the root's, on the main stream after the rank's skc_frame_render. */
int skc_comm_reduce_film(skc_comm *comm, int root);
/* max over ranks of a host float (device timings are reported as the max over ranks); a barrier of the main streams. */
int skc_comm_max_f32(skc_comm *comm, float *value);
int skc_comm_barrier(skc_comm *comm);
/* bytesPerRank host bytes from every rank to every rank (buf holds world x bytesPerRank; rank r's block at r x bytesPerRank):
 * material tables computed wavelength by wavelength across the ranks. */
int skc_comm_allgather_host(skc_comm *comm, void *buf, uint64_t bytesPerRank);

#ifdef __cplusplus
}
#endif
#endif /* SKIN_CUDA_H */

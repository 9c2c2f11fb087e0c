// skin_integrator.cpp -- MultipoleSubsurfaceIntegrator on the device + the C interface the
// Python tests/bench bind (skh_*).  Mirrors integrators/multipolesubsurface.cpp:170-304.
#include "skin_scene.h"
#include "skin_material.h"
#include "skin_cie.h"
#include "skin_tables.h"
#include <cstdio>
#include <cstring>

namespace skin {

bool MultipoleSubsurfaceIntegrator::Preprocess(skc_ctx *ctx, const SceneDesc &sd, int nCoresForTasks, std::string *err) {
    if (sd.lights.empty()) return true;                       // multipolesubsurface.cpp:173
    // GetSurfacePointsThroughTessellation (renderers/surfacepoints.cpp:362-369)
    surfacePoints.clear();
    for (const TriangleMesh &m : sd.meshes) m.TessellateSurfacePoints(minSampleDist, m.materialId, surfacePoints, incenter);
    int nPoints = (int)surfacePoints.size();
    int nT = std::max(32 * nCoresForTasks, nPoints / 4096);   // :194-199
    nTasks = RoundUpPow2((uint32_t)nT);
    if (skc_irradiance(ctx, surfacePoints.data(), surfacePoints.size(), 0, surfacePoints.size(), nTasks, mix, nullptr, nullptr)) {
        *err = skc_last_error(ctx); return false;
    }
    if (skc_build_octree(ctx, &octreeNodes, nullptr, nullptr)) { *err = skc_last_error(ctx); return false; }
    return true;
}

bool MultipoleSubsurfaceIntegrator::LiBatch(skc_ctx *ctx, const skc_ray *rays, uint64_t n, skc_hit *hits, float *L30, std::string *err) const {
    if (skc_camera_pass(ctx, rays, n, maxError, mix, hits, L30)) { *err = skc_last_error(ctx); return false; }
    return true;
}

}  // namespace skin

// ---------------------------------------------------------------------------------------------
// C interface for tests / bench (plain pointers; the handle owns everything it returns)
// ---------------------------------------------------------------------------------------------
using namespace skin;

struct skh_scene {
    SceneDesc sd;
    BuiltScene built;
    std::vector<skc_surface_point> points;
    std::vector<skc_ray> rays;
    std::vector<float> xy;
    std::string err;
};

extern "C" {

skh_scene *skh_load(const char *path, char *errbuf, int errlen) {
    skh_scene *s = new skh_scene();
    std::string err;
    if (!s->sd.Load(path) || !BuildScene(s->sd, nullptr, &s->built, &(err = s->sd.error))) {
        if (err.empty()) err = s->sd.error;
        if (errbuf) snprintf(errbuf, errlen, "%s", err.c_str());
        delete s;
        return nullptr;
    }
    return s;
}
// The same with the command-line options that reach the plugin factories (--quick, --outfile): main/pbrt.cpp:76-93
skh_scene *skh_load_opts(const char *path, int quick, const char *outfile, char *errbuf, int errlen) {
    skh_scene *s = new skh_scene();
    LoadOptions opt;
    opt.quickRender = quick != 0;
    if (outfile) opt.imageFile = outfile;
    std::string err;
    if (!s->sd.Load(path, &opt) || !BuildScene(s->sd, nullptr, &s->built, &(err = s->sd.error))) {
        if (err.empty()) err = s->sd.error;
        if (errbuf) snprintf(errbuf, errlen, "%s", err.c_str());
        delete s;
        return nullptr;
    }
    return s;
}
void skh_free(skh_scene *s) { delete s; }
int skh_warning_count(const skh_scene *s) { return (int)s->sd.warnings.size(); }
const char *skh_warning(const skh_scene *s, int i) { return s->sd.warnings[i].c_str(); }
const char *skh_film_filename(const skh_scene *s) { return s->sd.filmFilename.c_str(); }
// Pixel filter: out = {xwidth, ywidth}, table256 = ImageFilm::filterTable
void skh_filter(const skh_scene *s, float *out2, float *table256) { out2[0] = s->sd.filter.xwidth; out2[1] = s->sd.filter.ywidth; s->sd.filter.Table(table256); }
// BVH build parameters: returns maxnodeprims, split method name copied into buf
int skh_accelerator(const skh_scene *s, char *buf, int len) { snprintf(buf, len, "%s", s->sd.splitMethod.c_str()); return s->sd.maxNodePrims; }
void skh_light_samples(const skh_scene *s, int *out) { for (size_t i = 0; i < s->sd.lights.size(); ++i) out[i] = s->sd.lights[i].nSamples; }

// sizes: [0] nodes [1] prims [2] verts [3] tris [4] lights [5] hasN [6] hasUV [7] xres [8] yres [9] spp
void skh_sizes(const skh_scene *s, int64_t *out) {
    const TriangleMesh &m = *s->built.mesh;
    out[0] = (int64_t)s->built.nodes.size(); out[1] = (int64_t)s->built.prims.size(); out[2] = m.nverts; out[3] = m.ntris;
    out[4] = (int64_t)s->built.lights.size(); out[5] = !m.n.empty(); out[6] = !m.uvs.empty();
    out[7] = s->sd.xres; out[8] = s->sd.yres; out[9] = s->sd.spp;
}
const void *skh_nodes(const skh_scene *s) { return s->built.nodes.data(); }
const void *skh_prims(const skh_scene *s) { return s->built.prims.data(); }
const void *skh_lights(const skh_scene *s) { return s->built.lights.data(); }
const void *skh_mesh_P(const skh_scene *s) { return s->built.mesh->p.data(); }
const void *skh_mesh_N(const skh_scene *s) { return s->built.mesh->n.empty() ? nullptr : s->built.mesh->n.data(); }
const void *skh_mesh_uv(const skh_scene *s) { return s->built.mesh->uvs.empty() ? nullptr : s->built.mesh->uvs.data(); }
const void *skh_mesh_idx(const skh_scene *s) { return s->built.mesh->vertexIndex.data(); }
void skh_mesh_o2w(const skh_scene *s, float *out32, int *reverse, int *swaps) {
    memcpy(out32, s->built.mesh->ObjectToWorld.m.m, 64); memcpy(out32 + 16, s->built.mesh->ObjectToWorld.mInv.m, 64);
    *reverse = s->built.mesh->ReverseOrientation; *swaps = s->built.mesh->TransformSwapsHandedness;
}
// integrator params: maxerror, minsampledistance, mix, maxdepth
void skh_integrator_params(const skh_scene *s, float *out4) {
    out4[0] = s->sd.maxerror; out4[1] = s->sd.minsampledistance; out4[2] = s->sd.mix; out4[3] = (float)s->sd.maxdepth;
}
// "bool usepoissonpointfinder": {use, maxFails} and the finder's origin, the camera position CameraToWorld(Point(0, 0, 0))
void skh_poisson_params(const skh_scene *s, int *use2, float *origin3) {
    use2[0] = s->sd.usePoissonPointFinder ? 1 : 0; use2[1] = s->sd.poissonMaxFails;
    const PerspectiveCameraDesc cam = s->sd.Camera();
    const float *m = cam.CameraToWorld.m.m[0];
    origin3[0] = m[3]; origin3[1] = m[7]; origin3[2] = m[11];
}
// skin params: roughness nmperunit f_mel f_eu f_blood f_ohg desiredlength lerp t0 ior0 t1 ior1
void skh_skin_params(const skh_scene *s, float *out12) {
    const SkinParams &k = s->sd.skin;
    float v[12] = {k.roughness, k.nmperunit, k.f_mel, k.f_eu, k.f_blood, k.f_ohg, (float)k.desiredlength, (float)k.lerponthinslab,
                   k.layerThickness[0], k.layerIor[0], k.layerThickness[1], k.layerIor[1]};
    memcpy(out12, v, sizeof v);
}
// The infinite light's map and distribution (SceneDesc::env): sizes {w, h} (0, 0 = none), then the arrays by index
// 0 rgb, 1 condFunc, 2 condCdf, 3 condInt, 4 margFunc, 5 margCdf; margInt through the return value of skh_env_sizes' float
float skh_env_sizes(const skh_scene *s, int *wh2) { wh2[0] = s->sd.env.present ? s->sd.env.w : 0; wh2[1] = s->sd.env.present ? s->sd.env.h : 0; return s->sd.env.margInt; }
const float *skh_env_array(const skh_scene *s, int which) {
    const EnvMap &e = s->sd.env;
    const std::vector<float> *v[6] = {&e.rgb, &e.condFunc, &e.condCdf, &e.condInt, &e.margFunc, &e.margCdf};
    return which >= 0 && which < 6 ? v[which]->data() : nullptr;
}
// The material's albedo texture: number of levels (0 = none), per-level sizes, mapping; then the texels and the weight table
int skh_albedo_tex_info(const skh_scene *s, int *w12, int *h12, float *map4) {
    const AlbedoTexture &a = s->sd.skin.albedoTex;
    if (!a.present) return 0;
    for (size_t i = 0; i < a.w.size() && i < 12; ++i) { w12[i] = a.w[i]; h12[i] = a.h[i]; }
    map4[0] = a.su; map4[1] = a.sv; map4[2] = a.du; map4[3] = a.dv;
    return (int)a.w.size();
}
// The material's bump map, same layout with one float per texel
int skh_bump_tex_info(const skh_scene *s, int *w12, int *h12, float *map4) {
    const AlbedoTexture &a = s->sd.skin.bumpTex;
    if (!a.present) return 0;
    for (size_t i = 0; i < a.w.size() && i < 12; ++i) { w12[i] = a.w[i]; h12[i] = a.h[i]; }
    map4[0] = a.su; map4[1] = a.sv; map4[2] = a.du; map4[3] = a.dv;
    return (int)a.w.size();
}
const float *skh_bump_tex_texels(const skh_scene *s) { return s->sd.skin.bumpTex.rgb.data(); }
const float *skh_bump_tex_lut(const skh_scene *s) { return s->sd.skin.bumpTex.weightLut; }
const float *skh_albedo_tex_rgb(const skh_scene *s) { return s->sd.skin.albedoTex.rgb.data(); }
const float *skh_albedo_tex_lut(const skh_scene *s) { return s->sd.skin.albedoTex.weightLut; }
// The device mesh's alpha texture: sizes {w, h} (0 = none) and mapping {su, sv, du, dv}; the texels
static const AlphaTexture &alphaOf(const skh_scene *s) { return s->built.mesh ? s->built.mesh->alpha : s->sd.meshes[0].alpha; }
void skh_alpha_info(const skh_scene *s, int *wh2, float *map4) {
    const AlphaTexture &a = alphaOf(s);
    wh2[0] = a.present ? a.w : 0; wh2[1] = a.present ? a.h : 0;
    map4[0] = a.su; map4[1] = a.sv; map4[2] = a.du; map4[3] = a.dv;
}
const float *skh_alpha_texels(const skh_scene *s) { return alphaOf(s).texels.data(); }
// "color albedo" of the material: rgb, and whether it was given at all (Spectrum(1) otherwise)
int skh_skin_albedo(const skh_scene *s, float *rgb3) { memcpy(rgb3, s->sd.skin.albedo, 12); return s->sd.skin.hasAlbedo ? 1 : 0; }
// Tessellation (Preprocess' point source).  Returns the count; pointer valid until the next call.
// The points file in both its forms (skin_scene.h ReadPointsFile / WritePointsFile); out may be NULL to ask for the count only
int skh_points_write(const char *path, const void *pts, uint64_t n, int versioned, char *errbuf, int errlen) {
    std::string err;
    if (WritePointsFile(path, (const skc_surface_point *)pts, n, versioned != 0, &err)) return 0;
    if (errbuf && errlen > 0) snprintf(errbuf, errlen, "%s", err.c_str());
    return 1;
}
int skh_points_read(const char *path, void *out, uint64_t cap, uint64_t *n, char *errbuf, int errlen) {
    std::vector<skc_surface_point> pts; std::string err;
    if (!ReadPointsFile(path, &pts, &err)) { if (errbuf && errlen > 0) snprintf(errbuf, errlen, "%s", err.c_str()); return 1; }
    *n = pts.size();
    if (out && cap >= pts.size() && !pts.empty()) memcpy(out, pts.data(), pts.size() * sizeof(skc_surface_point));
    return 0;
}
uint64_t skh_tessellate(skh_scene *s, float minDist, int incenter, const void **pts) {
    s->points.clear();
    for (const TriangleMesh &m : s->sd.meshes) m.TessellateSurfacePoints(minDist, m.materialId, s->points, incenter != 0);
    *pts = s->points.data();
    return s->points.size();
}
uint64_t skh_count_points(const skh_scene *s, float minDist) {
    uint64_t n = 0;
    for (const TriangleMesh &m : s->sd.meshes) n += m.CountSurfacePoints(minDist);
    return n;
}
// Camera rays for given raster positions (cameras/perspective.cpp)
void skh_camera_rays(skh_scene *s, const float *xy, uint64_t n, skc_ray *out) {
    PerspectiveCamera cam(InverseT(s->sd.WorldToCamera), s->sd.fov, s->sd.xres, s->sd.yres, s->sd.frameAspect,
                          s->sd.haveScreenWindow ? s->sd.screenWindow : nullptr);   // api.cpp:862
    for (uint64_t i = 0; i < n; ++i) cam.GenerateRay(xy[2 * i], xy[2 * i + 1], &out[i]);
}
// A deterministic stratified sample pattern for benches: spp samples per pixel, pixels visited
// tile by tile (16x16, the granularity of the reference's SamplerRendererTask windows,
// renderers/samplerrenderer.cpp:206-208) and row-major inside a tile.  NOT the reference's
// LDSampler stream; parity tests use the exported camera_xy instead.
uint64_t skh_bench_rays(skh_scene *s, const void **rays) {
    int W = s->sd.xres, H = s->sd.yres, spp = std::max(1, s->sd.spp);
    int sx = 1; while (sx * sx < spp) ++sx;
    s->xy.resize((size_t)W * H * spp * 2);
    size_t k = 0;
    const int T = 16;
    for (int ty = 0; ty < H; ty += T) for (int tx = 0; tx < W; tx += T)
        for (int m = 0; m < T * T; ++m) {
            int lx, ly;
            if (getenv("SKH_MORTON")) { lx = (m & 1) | ((m >> 1) & 2) | ((m >> 2) & 4) | ((m >> 3) & 8); ly = ((m >> 1) & 1) | ((m >> 2) & 2) | ((m >> 3) & 4) | ((m >> 4) & 8); }
            else { lx = m % T; ly = m / T; }
            int x = tx + lx, y = ty + ly;
            if (x >= W || y >= H) continue;
            for (int q = 0; q < spp; ++q) { s->xy[k++] = x + ((q % sx) + .5f) / sx; s->xy[k++] = y + ((q / sx) + .5f) / sx; }
        }
    s->rays.resize((size_t)W * H * spp);
    skh_camera_rays(s, s->xy.data(), s->rays.size(), s->rays.data());
    *rays = s->rays.data();
    return s->rays.size();
}
// nTasks of IrradianceTask (multipolesubsurface.cpp:194-199)
uint32_t skh_irradiance_tasks(uint64_t nPoints, int nCores) {
    return RoundUpPow2((uint32_t)std::max(32 * nCores, (int)nPoints / 4096));
}

// LayeredSkin layer coefficients (a11): out124 = [layer][mua30 | musp30], eta[2], thickness[2]
void skh_skin_coefficients(const skh_scene *s, float *out124) {
    SkinLayerSpectra sp = ComputeSkinLayerSpectra(s->sd.skin);
    float *o = out124;
    for (int l = 0; l < 2; ++l) { memcpy(o, sp.mua[l], 120); o += 30; memcpy(o, sp.musp[l], 120); o += 30; }
    o[0] = sp.eta[0]; o[1] = sp.eta[1]; o[2] = sp.thickness[0]; o[3] = sp.thickness[1];
}
// The same without a scene: params9 = f_mel f_eu f_blood f_ohg nmperunit thickness0 ior0 thickness1 ior1
void skh_skin_spectra(const float *params9, float *out124) {
    SkinParams k;
    k.f_mel = params9[0]; k.f_eu = params9[1]; k.f_blood = params9[2]; k.f_ohg = params9[3]; k.nmperunit = params9[4];
    k.layerThickness[0] = params9[5]; k.layerIor[0] = params9[6]; k.layerThickness[1] = params9[7]; k.layerIor[1] = params9[8];
    SkinLayerSpectra sp = ComputeSkinLayerSpectra(k);
    float *o = out124;
    for (int l = 0; l < 2; ++l) { memcpy(o, sp.mua[l], 120); o += 30; memcpy(o, sp.musp[l], 120); o += 30; }
    o[0] = sp.eta[0]; o[1] = sp.eta[1]; o[2] = sp.thickness[0]; o[3] = sp.thickness[1];
}
// ComputeMultipoleProfile on the device for the scene's skin material and upload of the result together
// with the caller's rho_hd table, or the device-generated one when rhoHd is NULL (skc_upload_profile).  hdr120 = [30][dsqSpacing, rcpDsqSpacing, totalR, totalT];
// rdOut (optional) receives [30][len].  Two-phase like skc_profile_batch: rdOut == NULL only reports len.
int skh_build_profile(skh_scene *s, skc_ctx *ctx, const float *rhoHd, uint32_t *len, float *hdr120, float *rdOut, int upload) {
    SkinLayerSpectra sp = ComputeSkinLayerSpectra(s->sd.skin);
    if (!hdr120 && !rdOut && !upload) {
        std::vector<skc_layer_spec> specs; std::vector<uint32_t> lps; std::vector<skc_mpc_options> opts;
        BuildProfileJobs(sp, s->sd.skin.desiredlength, s->sd.skin.lerponthinslab, specs, lps, opts);
        return skc_profile_batch(ctx, specs.data(), lps.data(), opts.data(), kWavelengths, 1, len, nullptr, nullptr, nullptr, nullptr);
    }
    ProfileTable t;
    int rc = ComputeMultipoleProfile(ctx, sp, s->sd.skin.desiredlength, s->sd.skin.lerponthinslab, t);
    if (rc) return rc;
    if (len) *len = t.len;
    if (hdr120) for (int w = 0; w < kWavelengths; ++w) {
        hdr120[4 * w] = t.dsqSpacing[w]; hdr120[4 * w + 1] = t.rcpDsqSpacing[w];
        hdr120[4 * w + 2] = t.totalReflectance[w]; hdr120[4 * w + 3] = t.totalTransmittance[w];
    }
    if (rdOut) memcpy(rdOut, t.rd.data(), t.rd.size() * sizeof(float));
    if (upload) {
        std::vector<float> rho;
        if (!rhoHd) {   // ComputeRhoDataFromBxDF for Microfacet(Beckmann(roughness), FresnelDielectric(1, ior0)), layeredskin.cpp:103-109
            rho.resize(1025 * 30);
            if ((rc = skc_rho_table(ctx, s->sd.skin.roughness, 1.f, s->sd.skin.layerIor[0], 0, rho.data()))) return rc;
            rhoHd = rho.data();
        }
        if ((rc = skc_upload_profile(ctx, t.rd.data(), t.rcpDsqSpacing, t.len, rhoHd))) return rc;
        return InstallAlbedo(ctx, s->sd.skin, s->sd.mix);
    }
    return 0;
}

// ---- whole frame --------------------------------------------------------------------------------
// What api.cpp's MakeCamera / MakeFilm / MakeSampler / MakeRenderer derive from the scene file for the device frame
// pass: perspective camera matrices (api.cpp:862, cameras/perspective.cpp:124-155), Film "image" with the default
// box filter of half-width 0.5 (film/image.cpp:55-66, filters/box.cpp), LDSampler pixelsamples rounded up to a power
// of two (lowdiscrepancy.cpp:41-50), the render task count (samplerrenderer.cpp:206-208), EmissionIntegrator's two
// 1-sample arrays (integrators/emission.cpp:42-45), the XYZ matching tables.
void skh_frame_desc(const skh_scene *s, int nCores, skc_frame_desc *d) { MakeFrameDesc(s->sd, nCores, d); }
// ImageFilm::WriteImage (film/image.cpp:178-213): XYZ -> RGB, divide by the weight sum, clamp at 0.  rgb = [h][w][3].
void skh_film_to_rgb(const float *pixels4, int w, int h, float *rgb) {
    for (int i = 0; i < w * h; ++i) {
        const float *xyz = pixels4 + 4 * i;
        float *o = rgb + 3 * i;
        o[0] = 3.240479f * xyz[0] - 1.537150f * xyz[1] - 0.498535f * xyz[2];      // XYZToRGB (core/spectrum.h:51-55)
        o[1] = -0.969256f * xyz[0] + 1.875991f * xyz[1] + 0.041556f * xyz[2];
        o[2] = 0.055648f * xyz[0] - 0.204043f * xyz[1] + 1.057311f * xyz[2];
        float weightSum = xyz[3];
        if (weightSum != 0.f) {
            float invWt = 1.f / weightSum;
            o[0] = std::max(0.f, o[0] * invWt); o[1] = std::max(0.f, o[1] * invWt); o[2] = std::max(0.f, o[2] * invWt);
        }
    }
}
// Portable float map, little endian, bottom row first (core/imageio.cpp WriteImagePFM)
int skh_write_pfm(const char *path, const float *rgb, int w, int h) {
    FILE *fp = fopen(path, "wb");
    if (!fp) return 1;
    fprintf(fp, "PF\n%d %d\n-1.000000\n", w, h);
    for (int y = h - 1; y >= 0; --y)
        if (fwrite(rgb + (size_t)3 * w * y, sizeof(float), (size_t)3 * w, fp) != (size_t)3 * w) { fclose(fp); return 1; }
    fclose(fp);
    return 0;
}

}  // extern "C"

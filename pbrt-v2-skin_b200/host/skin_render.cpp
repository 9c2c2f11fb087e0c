// skin_render.cpp -- SamplerRenderer::Render (renderers/samplerrenderer.cpp:191-225) for a parsed scene, every stage on
// the device through include/skin_cuda.h:
//   LayeredSkin construction   materials/layeredskin.cpp:39-123   -> skc_profile_batch, skc_rho_table, skc_upload_profile
//   RenderOptions::MakeScene   core/api.cpp:1190-1225             -> BuildScene (host BVH build) + skc_upload_bvh / skc_upload_lights
//   integrator Preprocess      multipolesubsurface.cpp:170-238    -> skc_tessellate_resident, skc_irradiance_resident, skc_build_octree
//   SamplerRendererTask loop   samplerrenderer.cpp:59-167         -> skc_frame_setup / _samples / _render
//   ImageFilm::WriteImage      film/image.cpp:178-213             -> skc_frame_film_get + skh_film_to_rgb + the image writers below
#include "skin_scene.h"
#include "skin_material.h"
#include "skin_cie.h"
#include "skin_tables.h"
#include <chrono>
#include <cstdio>
#include <cstring>
#include <thread>

extern "C" void skh_film_to_rgb(const float *pixels4, int w, int h, float *rgb);
extern "C" int skh_write_pfm(const char *path, const float *rgb, int w, int h);

namespace skin {

void MakeFrameDesc(const SceneDesc &sd, int nCores, skc_frame_desc *d) {
    memset(d, 0, sizeof *d);
    d->xres = sd.xres; d->yres = sd.yres;
    d->spp = (int)RoundUpPow2((uint32_t)std::max(1, sd.spp));                // lowdiscrepancy.cpp:41-50
    int nPixels = sd.xres * sd.yres;
    d->nTasks = (int)RoundUpPow2((uint32_t)std::max(32 * nCores, nPixels / (16 * 16)));   // samplerrenderer.cpp:206-208
    d->extra1D = 2;                                                           // EmissionIntegrator::RequestSamples
    PerspectiveCameraDesc cam = sd.Camera();
    memcpy(d->cameraToWorld, cam.CameraToWorld.m.m, 64);
    memcpy(d->rasterToCamera, cam.RasterToCamera.m.m, 64);
    d->filterXWidth = sd.filter.xwidth; d->filterYWidth = sd.filter.ywidth;
    sd.filter.Table(d->filterTable);                                          // film/image.cpp:55-68
    for (int k = 0; k < 3; ++k) memcpy(d->cieXYZ[k], kCieXYZ30[k], sizeof(float) * 30);
    d->xyzScale = float(700 - 400) / float(kCieYIntegral * 30);               // core/spectrum.h:381-382
    d->roughness = sd.skin.roughness; d->ior = sd.skin.layerIor[0];
    RGBToSpectrum(sd.skin.Kr, d->Kr);                                         // "color Kr" / "color Kt": Spectrum::FromRGB
    RGBToSpectrum(sd.skin.Kt, d->Kt);
    d->matteKd = 0.5f;                                                        // materials/matte.cpp: Kd default
}

// LayeredSkin::LayeredSkin (materials/layeredskin.cpp:39-123): layer spectra -> 30 diffusion profiles -> rho_hd, all
// generated on the device and installed as the context's material tables.
int BuildSkinMaterial(skc_ctx *ctx, const SkinParams &skin, skc_comm *comm) {
    SkinLayerSpectra sp = ComputeSkinLayerSpectra(skin);
    ProfileTable t;
    int rc = ComputeMultipoleProfile(ctx, sp, skin.desiredlength, skin.lerponthinslab, t, comm);
    if (rc) return rc;
    std::vector<float> rho(1025 * 30);
    if ((rc = skc_rho_table(ctx, skin.roughness, 1.f, skin.layerIor[0], 0, rho.data()))) return rc;
    return skc_upload_profile(ctx, t.rd.data(), t.rcpDsqSpacing, t.len, rho.data());
}

int InstallAlbedo(skc_ctx *ctx, const SkinParams &skin, float mix) {
    if (!skin.hasAlbedo || skin.albedoTex.present) return skc_set_albedo_factors(ctx, nullptr, nullptr);
    float al[30], a[30], b[30];
    RGBToSpectrum(skin.albedo, al);                                           // "color albedo": Spectrum::FromRGB, reflectance
    for (int i = 0; i < 30; ++i) { a[i] = powf(al[i], mix); b[i] = powf(al[i], 1.f - mix); }      // Pow(Spectrum, float): powf per sample
    return skc_set_albedo_factors(ctx, a, b);
}

int UploadScene(skc_ctx *ctx, const BuiltScene &b) {
    const TriangleMesh &m = *b.mesh;
    float o2w[32];
    memcpy(o2w, m.ObjectToWorld.m.m, 64); memcpy(o2w + 16, m.ObjectToWorld.mInv.m, 64);
    int rc = skc_upload_bvh(ctx, b.nodes.data(), (uint32_t)b.nodes.size(), b.prims.data(), (uint32_t)b.prims.size(),
                            &m.p[0].x, m.n.empty() ? nullptr : &m.n[0].x, m.uvs.empty() ? nullptr : m.uvs.data(), (uint32_t)m.nverts,
                            m.vertexIndex.data(), (uint32_t)m.ntris, o2w, m.ReverseOrientation, m.TransformSwapsHandedness,
                            b.lights.data(), (uint32_t)b.lights.size());
    if (rc) return rc;
    return skc_upload_lights(ctx, b.lights.data(), (uint32_t)b.lights.size());
}

int UploadAlbedoTexture(skc_ctx *ctx, const AlbedoTexture &a) {
    if (!a.present) return skc_upload_albedo_texture(ctx, 0, nullptr, nullptr, nullptr, nullptr, 1.f, 1.f, 0.f, 0.f);
    return skc_upload_albedo_texture(ctx, (uint32_t)a.w.size(), a.w.data(), a.h.data(), a.rgb.data(), a.weightLut, a.su, a.sv, a.du, a.dv);
}

int UploadBumpTexture(skc_ctx *ctx, const AlbedoTexture &a) {
    if (!a.present) return skc_upload_bump_texture(ctx, 0, nullptr, nullptr, nullptr, nullptr, 1.f, 1.f, 0.f, 0.f);
    return skc_upload_bump_texture(ctx, (uint32_t)a.w.size(), a.w.data(), a.h.data(), a.rgb.data(), a.weightLut, a.su, a.sv, a.du, a.dv);
}

int UploadAlpha(skc_ctx *ctx, const AlphaTexture &a) {
    if (!a.present) return skc_upload_alpha(ctx, 0, 0, nullptr, 1.f, 1.f, 0.f, 0.f);
    return skc_upload_alpha(ctx, (uint32_t)a.w, (uint32_t)a.h, a.texels.data(), a.su, a.sv, a.du, a.dv);
}

int UploadEnvMap(skc_ctx *ctx, const EnvMap &e) {
    if (!e.present) return skc_upload_envmap(ctx, 0, 0, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0.f);
    return skc_upload_envmap(ctx, (uint32_t)e.w, (uint32_t)e.h, e.rgb.data(), e.condFunc.data(), e.condCdf.data(), e.condInt.data(), e.margFunc.data(), e.margCdf.data(), e.margInt);
}

namespace {
struct PointsFileHeader {            // 32 bytes, little-endian fields as written by the machine that made the file
    char magic[8];                   // "SKINPTS\1"
    uint32_t version;                // 1
    uint32_t recordBytes;            // sizeof(skc_surface_point) = 44
    uint64_t count;
    uint32_t byteOrder;              // 0x01020304 as the writer's uint32
    uint32_t reserved;
};
const char kPointsMagic[8] = {'S', 'K', 'I', 'N', 'P', 'T', 'S', 1};
}  // namespace

bool ReadPointsFile(const std::string &path, std::vector<skc_surface_point> *pts, std::string *err) {
    pts->clear();
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) { *err = "Unable to open file \"" + path + "\""; return false; }
    fseek(f, 0, SEEK_END); const long size = ftell(f); fseek(f, 0, SEEK_SET);
    PointsFileHeader h;
    bool tagged = size >= (long)sizeof h && fread(&h, sizeof h, 1, f) == 1 && !memcmp(h.magic, kPointsMagic, 8);
    bool ok = true;
    if (tagged) {
        if (h.byteOrder != 0x01020304u) { *err = "points file \"" + path + "\" was written with another byte order"; ok = false; }
        else if (h.version != 1 || h.recordBytes != sizeof(skc_surface_point)) { *err = "points file \"" + path + "\": unknown version or record size"; ok = false; }
        else if ((uint64_t)(size - (long)sizeof h) != h.count * sizeof(skc_surface_point)) { *err = "points file \"" + path + "\" is truncated or has trailing bytes"; ok = false; }
        else { pts->resize(h.count); ok = !h.count || fread(pts->data(), sizeof(skc_surface_point), h.count, f) == h.count; if (!ok) *err = "short read from \"" + path + "\""; }
    } else {
        fseek(f, 0, SEEK_SET);
        pts->resize((size_t)size / sizeof(skc_surface_point));
        ok = pts->empty() || fread(pts->data(), sizeof(skc_surface_point), pts->size(), f) == pts->size();
        if (!ok) *err = "short read from \"" + path + "\"";
    }
    fclose(f);
    if (!ok) pts->clear();
    return ok;
}

bool WritePointsFile(const std::string &path, const skc_surface_point *pts, uint64_t n, bool versioned, std::string *err) {
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) { *err = "Unable to open output file \"" + path + "\""; return false; }
    bool ok = true;
    if (versioned) {
        PointsFileHeader h; memset(&h, 0, sizeof h);
        memcpy(h.magic, kPointsMagic, 8); h.version = 1; h.recordBytes = (uint32_t)sizeof(skc_surface_point); h.count = n; h.byteOrder = 0x01020304u;
        ok = fwrite(&h, sizeof h, 1, f) == 1;
    }
    ok = ok && (!n || fwrite(pts, sizeof(skc_surface_point), n, f) == n);
    fclose(f);
    if (!ok) *err = "short write to \"" + path + "\"";
    return ok;
}

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// One rank's share of SamplerRenderer::Render.  comm == nullptr: the whole frame on one GPU.  With a communicator:
// the material's wavelengths, the irradiance points and the render tasks are dealt over the ranks; the points are all-gathered
// (positions first, the rows behind the octree topology pass), every rank builds the same octree, and the films are summed into
// rank 0's, which alone gets the image.  rgb: [yres][xres][3] as ImageFilm::WriteImage hands to ::WriteImage (rank 0).
bool RenderSceneRank(skc_ctx *ctx, skc_comm *comm, const SceneDesc &sd, const BuiltScene &built, int nCores, std::vector<float> *rgb, RenderStats *st, std::string *err) {
    auto dev = [&](int rc) { if (rc) *err = skc_last_error(ctx); return rc == 0; };
    const int world = comm ? skc_comm_world(comm) : 1, rank = comm ? skc_comm_rank(comm) : 0;
    double t0 = now();
    if (!dev(UploadScene(ctx, built)) || !dev(UploadEnvMap(ctx, sd.env)) || !dev(UploadAlpha(ctx, built.mesh->alpha)) || !dev(UploadAlbedoTexture(ctx, sd.skin.albedoTex)) ||
        !dev(UploadBumpTexture(ctx, sd.skin.bumpTex))) return false;
    double t1 = now();
    if (!dev(BuildSkinMaterial(ctx, sd.skin, comm)) || !dev(InstallAlbedo(ctx, sd.skin, sd.mix)) || !dev(skc_synchronize(ctx))) return false;
    double t2 = now();
    // the camera pass's Preprocess-independent part (sampler stream, camera rays, closest hits, direct lighting) goes to the
    // second stream first and runs beside Preprocess
    skc_frame_desc fd;
    MakeFrameDesc(sd, nCores, &fd);
    fd.taskRank = rank; fd.taskWorld = world;
    uint64_t nSamples = 0;
    if (sd.lights.empty()) { *err = "no light sources in the scene: the subsurface integrator has nothing to render"; return false; }
    if (!dev(skc_frame_setup(ctx, &fd, &nSamples)) || !dev(skc_frame_prefetch(ctx))) return false;
    // Preprocess (multipolesubsurface.cpp:170-238)
    uint64_t nPts = 0;
    {
        std::vector<skc_surface_point> filePts;
        if (!sd.pointsFile.empty()) {                                        // :176-181, ReadBinaryFile (core/floatfile.h:45-63); a file that cannot be used
            std::string why;                                                 // leaves pts empty and the points are generated, as in the reference
            if (!ReadPointsFile(sd.pointsFile, &filePts, &why)) fprintf(stderr, "Error: %s\n", why.c_str());
        }
        bool fromFile = !filePts.empty();
        if (fromFile) {
            nPts = filePts.size();
            if (!dev(skc_set_surface_points(ctx, filePts.data(), nPts))) return false;
        } else if (sd.usePoissonPointFinder) {                               // :183-185; deterministic (one SurfacePointTask), so every rank finds the same set
            const float origin[3] = {fd.cameraToWorld[3], fd.cameraToWorld[7], fd.cameraToWorld[11]};      // CameraToWorld(shutterOpen, Point(0, 0, 0))
            if (!dev(skc_poisson_resident(ctx, origin, sd.minsampledistance, built.mesh->materialId, sd.poissonMaxFails, &nPts, nullptr))) return false;
            fromFile = true;                                                  // the whole set is resident from here on
        } else if (world == 1) {                                             // pts.size() == 0 -> tessellate (:182-190)
            if (!dev(skc_tessellate_resident(ctx, sd.minsampledistance, sd.incenter, built.mesh->materialId, &nPts))) return false;
        } else if (!dev(skc_tessellate_range(ctx, sd.minsampledistance, sd.incenter, built.mesh->materialId, 0, 0, &nPts))) return false;   // the count
        st->irradianceTasks = RoundUpPow2((uint32_t)std::max(32 * nCores, (int)nPts / 4096));   // :194-199
        if (world == 1) {
            if (!dev(skc_irradiance_resident(ctx, 0, nPts, st->irradianceTasks, sd.mix, nullptr))) return false;
            if (!dev(skc_build_octree(ctx, &st->octreeNodes, nullptr, nullptr))) return false;
        } else {
            // contiguous equal shards (the scramble stream of a point depends on its task and its place in it, both functions
            // of the point index alone, so any range can be computed anywhere); the last shard may be short
            const uint64_t shard = (nPts + world - 1) / world, first = std::min<uint64_t>((uint64_t)rank * shard, nPts), count = std::min<uint64_t>(shard, nPts - first);
            void *pa = nullptr, *rows = nullptr;
            if (!dev(skc_irradiance_dev(ctx, shard * world, &pa, &rows))) return false;             // arrays sized for the padded exchange
            if (!fromFile && count && !dev(skc_tessellate_range(ctx, sd.minsampledistance, sd.incenter, built.mesh->materialId, first, count, &nPts))) return false;
            if (!dev(skc_irradiance_resident(ctx, first, count, st->irradianceTasks, sd.mix, nullptr))) return false;
            if (!dev(skc_comm_allgather_positions(comm, shard)) || !dev(skc_comm_allgather_rows_begin(comm, shard))) return false;
            if (!dev(skc_set_point_count(ctx, nPts)) || !dev(skc_build_octree_topology(ctx))) return false;
            if (!dev(skc_comm_allgather_rows_wait(comm)) || !dev(skc_build_octree_finalize(ctx, &st->octreeNodes, nullptr, nullptr))) return false;
        }
    }
    if (!dev(skc_synchronize(ctx))) return false;
    double t3 = now();
    // Mo + the subsurface term, the spectral sum and the film
    if (!dev(skc_frame_film_clear(ctx)) || !dev(skc_frame_render(ctx, 0, nSamples, sd.maxerror, sd.mix, nullptr))) return false;
    if (comm && !dev(skc_comm_reduce_film(comm, 0))) return false;
    if (rank == 0) {
        rgb->resize((size_t)sd.xres * sd.yres * 3);
        if (!dev(skc_frame_film_rgb(ctx, rgb->data()))) return false;       // ImageFilm::WriteImage's XYZ -> RGB / weight on the device
    } else if (!dev(skc_synchronize(ctx))) return false;
    double t4 = now();
    st->buildSec = t1 - t0; st->materialSec = t2 - t1; st->preprocessSec = t3 - t2; st->renderSec = t4 - t3;
    st->nPoints = nPts; st->nSamples = nSamples; st->renderTasks = (uint32_t)fd.nTasks;
    return true;
}

bool RenderScene(skc_ctx *ctx, const SceneDesc &sd, int nCores, std::vector<float> *rgb, RenderStats *st, std::string *err) {
    double t0 = now();
    BuiltScene built;
    if (!BuildScene(sd, nullptr, &built, err)) return false;
    double tBuild = now() - t0;
    if (!RenderSceneRank(ctx, nullptr, sd, built, nCores, rgb, st, err)) return false;
    st->buildSec += tBuild;
    return true;
}

// `pbrt --gpus N scene.pbrt`: one process, N contexts on devices firstDevice .. firstDevice + N - 1, one host thread per context,
// the exchanges through NCCL (skc_comm_create_local).  The image and the statistics are rank 0's.
bool RenderSceneMultiGPU(int firstDevice, int nGpus, const SceneDesc &sd, int nCores, std::vector<float> *rgb, RenderStats *st, std::string *err) {
    double t0 = now();
    BuiltScene built;
    if (!BuildScene(sd, nullptr, &built, err)) return false;
    double tBuild = now() - t0;
    std::vector<skc_ctx *> ctxs(nGpus, nullptr);
    std::vector<skc_comm *> comms(nGpus, nullptr);
    bool ok = true;
    for (int i = 0; i < nGpus && ok; ++i) {
        ctxs[i] = skc_create(firstDevice + i);
        if (!ctxs[i]) { *err = std::string("skc_create: ") + skc_last_error(nullptr); ok = false; }
    }
    if (ok && nGpus > 1 && skc_comm_create_local(ctxs.data(), nGpus, comms.data())) { *err = skc_last_error(ctxs[0]); ok = false; }
    if (ok) {
        std::vector<std::string> errs(nGpus);
        std::vector<RenderStats> stats(nGpus);
        std::vector<char> good(nGpus, 0);
        std::vector<std::vector<float>> images(nGpus);
        std::vector<std::thread> threads;
        for (int i = 0; i < nGpus; ++i)
            threads.emplace_back([&, i]() { good[i] = RenderSceneRank(ctxs[i], comms[i], sd, built, nCores, &images[i], &stats[i], &errs[i]) ? 1 : 0; });
        for (std::thread &t : threads) t.join();
        for (int i = 0; i < nGpus; ++i) if (!good[i]) { if (ok) *err = "GPU " + std::to_string(firstDevice + i) + ": " + errs[i]; ok = false; }
        if (ok) { *rgb = std::move(images[0]); *st = stats[0]; st->buildSec += tBuild; }
    }
    for (skc_comm *c : comms) if (c) skc_comm_destroy(c);
    for (skc_ctx *c : ctxs) if (c) skc_destroy(c);
    return ok;
}

bool TessellateSurfacePointsToFile(skc_ctx *ctx, const SceneDesc &sd, uint64_t *nPoints, std::string *err) {
    BuiltScene built;
    if (!BuildScene(sd, nullptr, &built, err)) return false;
    uint64_t n = 0;
    if (UploadScene(ctx, built) || UploadEnvMap(ctx, sd.env) || UploadAlpha(ctx, built.mesh->alpha) || UploadBumpTexture(ctx, sd.skin.bumpTex)) { *err = skc_last_error(ctx); return false; }
    const bool poisson = sd.rendererName == "surfacepoints";                 // SurfacePointsRenderer (the Poisson finder) instead of the tessellator
    if (poisson) {
        const PerspectiveCameraDesc cam = sd.Camera();
        const float *m = cam.CameraToWorld.m.m[0];
        const float origin[3] = {m[3], m[7], m[11]};                         // pCamera = CameraToWorld(shutterOpen, Point(0, 0, 0)), core/api.cpp:1256
        if (skc_poisson_resident(ctx, origin, sd.rendererMinDist, built.mesh->materialId, sd.poissonMaxFails, &n, nullptr)) { *err = skc_last_error(ctx); return false; }
    } else if (skc_tessellate_resident(ctx, sd.rendererMinDist, 0, built.mesh->materialId, &n)) { *err = skc_last_error(ctx); return false; }
    *nPoints = n;
    if (sd.rendererFilename.empty()) return true;                            // "(Not writing to file)"
    std::vector<skc_surface_point> pts(n);
    if (n && skc_get_surface_points(ctx, 0, n, pts.data())) { *err = skc_last_error(ctx); return false; }
    if (poisson) {                                                           // the text file of SurfacePointsRenderer::Render (surfacepoints.cpp:150-166)
        FILE *f = fopen(sd.rendererFilename.c_str(), "w");
        if (!f) { *err = "Unable to open output file \"" + sd.rendererFilename + "\""; return false; }
        fprintf(f, "# points generated by SurfacePointsRenderer\n");
        fprintf(f, "# position (x,y,z), normal (x,y,z), area, rayEpsilon\n");
        for (const skc_surface_point &sp : pts)
            fprintf(f, "%g %g %g %g %g %g %g %g\n", sp.p[0], sp.p[1], sp.p[2], sp.n[0], sp.n[1], sp.n[2], sp.area, sp.rayEpsilon);
        fclose(f);
        return true;
    }
    return WritePointsFile(sd.rendererFilename, pts.data(), pts.size(), sd.rendererVersionedFile, err);
}

// WriteImageTGA (core/imageio.cpp:185-215): 24-bit BGR, gamma 2.2, top row first
static bool writeTGA(const std::string &name, const float *rgb, int w, int h) {
    FILE *f = fopen(name.c_str(), "wb");
    if (!f) return false;
    unsigned char hdr[18] = {0, 0, 2, 0, 0, 0, 0, 0, 0, 0, 0, 0, (unsigned char)(w & 255), (unsigned char)(w >> 8), (unsigned char)(h & 255),
                             (unsigned char)(h >> 8), 24, 0x20};
    bool ok = fwrite(hdr, 1, 18, f) == 18;
    std::vector<unsigned char> row(3 * (size_t)w);
    auto toByte = [](float v) { float x = 255.f * powf(v, 1.f / 2.2f); return (unsigned char)(x < 0.f ? 0.f : x > 255.f ? 255.f : x); };
    for (int y = 0; y < h && ok; ++y) {
        for (int x = 0; x < w; ++x) {
            const float *p = rgb + 3 * ((size_t)y * w + x);
            row[3 * x] = toByte(p[2]); row[3 * x + 1] = toByte(p[1]); row[3 * x + 2] = toByte(p[0]);
        }
        ok = fwrite(row.data(), 1, row.size(), f) == row.size();
    }
    fclose(f);
    return ok;
}

// ::WriteImage (core/imageio.cpp:78-104) without OpenEXR
bool WriteImage(const std::string &name, const float *rgb, int w, int h, std::string *err) {
    auto ends = [&](const char *a, const char *b) { return name.size() >= 5 && (!name.compare(name.size() - 4, 4, a) || !name.compare(name.size() - 4, 4, b)); };
    if (ends(".pfm", ".PFM")) { if (skh_write_pfm(name.c_str(), rgb, w, h)) { *err = "Unable to open output PFM file \"" + name + "\""; return false; } return true; }
    if (ends(".tga", ".TGA")) { if (!writeTGA(name, rgb, w, h)) { *err = "Unable to write image file \"" + name + "\""; return false; } return true; }
    if (ends(".exr", ".EXR")) { *err = "OpenEXR is not available in this build (as in a reference build without PBRT_HAS_OPENEXR): use a .pfm or .tga filename"; return false; }
    *err = "Can't determine image file type from suffix of filename \"" + name + "\"";
    return false;
}

}  // namespace skin

// ---- C interface (tests, bench, the pbrt executable) ------------------------------------------------------------------
using namespace skin;

extern "C" {

// `pbrt [--ncores n] [--outfile f] [--quick] scene` in one call: parse, build, render on `device`, write the image (when
// the film has a filename or outfile is given; "" -> pbrt.tga) and/or return the RGB pixels.  rgbOut may be NULL; when
// not, it must hold xres*yres*3 floats -- call with rgbOut == NULL and dims != NULL first to learn the size (no render).
// stats (optional) = {build s, material s, preprocess s, render s, points, samples, octree nodes}.
int skh_pbrt_render(const char *scenePath, int device, int nCores, int quick, const char *outfile, int writeImage, int *dims2, float *rgbOut,
                    double *stats7, char *errbuf, int errlen) {
    auto fail = [&](const std::string &e) { if (errbuf) snprintf(errbuf, errlen, "%s", e.c_str()); return 1; };
    SceneDesc sd;
    LoadOptions opt;
    opt.quickRender = quick != 0;
    if (outfile) opt.imageFile = outfile;
    if (!sd.Load(scenePath, &opt)) return fail(sd.error);
    if (dims2) { dims2[0] = sd.xres; dims2[1] = sd.yres; }
    if (!rgbOut && !writeImage) return 0;
    skc_ctx *ctx = skc_create(device);
    if (!ctx) return fail(std::string("skc_create: ") + skc_last_error(nullptr));
    if (sd.rendererName == "tessellatesurfacepoints" || sd.rendererName == "surfacepoints") {
        uint64_t n = 0;
        std::string e;
        bool ok = TessellateSurfacePointsToFile(ctx, sd, &n, &e);
        skc_destroy(ctx);
        if (stats7) stats7[4] = (double)n;
        return ok ? 0 : fail(e);
    }
    std::vector<float> rgb;
    RenderStats st;
    std::string err;
    bool ok = RenderScene(ctx, sd, nCores, &rgb, &st, &err);
    skc_destroy(ctx);
    if (!ok) return fail(err);
    if (rgbOut) memcpy(rgbOut, rgb.data(), rgb.size() * sizeof(float));
    if (stats7) { stats7[0] = st.buildSec; stats7[1] = st.materialSec; stats7[2] = st.preprocessSec; stats7[3] = st.renderSec;
                  stats7[4] = (double)st.nPoints; stats7[5] = (double)st.nSamples; stats7[6] = (double)st.octreeNodes; }
    if (writeImage) {
        std::string name = sd.filmFilename.empty() ? "pbrt.tga" : sd.filmFilename;
        if (!WriteImage(name, rgb.data(), sd.xres, sd.yres, &err)) return fail(err);
    }
    return 0;
}

}  // extern "C"

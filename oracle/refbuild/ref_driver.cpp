// ref_driver.cpp -- TEST INFRASTRUCTURE ONLY (oracle/README.md).
//
// Links against objects compiled from the UNMODIFIED-ARITHMETIC reference
// sources (see build_ref.sh / patch_ref.py) and drives the reference's own
// pbrt*() API directly, because the reference's scene parser needs flex, which
// is not in this image.  It runs the reference's own
// MultipoleSubsurfaceIntegrator / BVHAccel / LayeredSkin / MPC_* code and dumps
// the intermediate streams that pin parity for the B200 path:
//   surface points, irradiance points, flat BVH, profile + rho tables,
//   camera rays, closest hits, per-hit Mo, per-sample radiance, PFM image.
// Nothing here is product code; nothing in the product path calls it.
//
// The translation unit textually includes the reference's core/api.cpp so that
// the file-static RenderOptions (MakeScene / MakeRenderer) is reachable.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <functional>
#include <iostream>
#include <limits>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <sstream>
#include <string>
#include <typeinfo>
#include <unordered_map>
#include <utility>
#include <vector>

// Open up the reference classes: the dumps need private members.
#define private public
#define protected public
#include "core/api.cpp"
#include "integrators/diffusionutil.h"
#include "materials/layeredskin.h"
#include "materials/skincoeffs.h"
#include "accelerators/bvh.h"
#include "renderers/samplerrenderer.h"
#include "renderers/surfacepoints.h"
#include "shapes/trianglemesh.h"
#include "shapes/sphere.h"
#include "lights/diffuse.h"
#include "lights/point.h"
#include "lights/spot.h"
#include "lights/distant.h"
#include "lights/infinite.h"
#include "textures/imagemap.h"
#include "core/multipole.h"
#include "core/reflection.h"
#include "core/film.h"
#include "film/image.h"
#include "core/sampler.h"
#include "core/camera.h"
#include "core/parallel.h"
#include "core/rng.h"
#include "multipole/MultipoleProfileCalculator/MultipoleProfileCalculator.h"
#undef private
#undef protected

// parser.cpp is not linked (needs flex/bison output); api.cpp only needs these.
string current_file;
int line_num = 0;
bool ParseFile(const string &) { Severe("ref_driver: parser not available"); return false; }

// Mirrors of the file-local structs in core/multipole.cpp:40-56,95-101 (layout only).
struct DrvProfileEntry { float reflectance; };
struct DrvProfile {
    std::vector<DrvProfileEntry> data;
    float dsqSpacing, rcpDsqSpacing, totalReflectance, totalTransmittance;
};
struct DrvProfileData { DrvProfile spectralProfile[nSpectralSamples]; bool isRGBProfile; };

// Mirror of the file-local LinearBVHNode in accelerators/bvh.cpp:113-123.
struct DrvLinearBVHNode {
    BBox bounds;
    union { uint32_t primitivesOffset; uint32_t secondChildOffset; };
    uint8_t nPrimitives, axis, pad[2];
};

static double now_s() {
    using namespace std::chrono;
    return duration_cast<duration<double>>(steady_clock::now().time_since_epoch()).count();
}

template <class T>
static void dump(const std::string &path, const T *data, size_t n) {
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) { fprintf(stderr, "cannot write %s\n", path.c_str()); exit(2); }
    if (n) fwrite(data, sizeof(T), n, f);
    fclose(f);
}

// ---------------------------------------------------------------------------
// Scene description shared with the B200 host code (tests/golden/*.scene):
//   film W H | lookat 9f | fov f | spp n
//   integrator key value ...   (maxerror minsampledistance mix maxdepth)
//   light tx ty tz radius Lr Lg Lb nsamples
//   skin key value ...  (roughness nmperunit f_mel f_eu f_blood f_ohg desiredlength
//                        layers t0 ior0 t1 ior1)
//   xform translate|rotate|scale ...   (applied, in order, to the next mesh)
//   mesh file.bin       (int32 nv,nt,hasN,hasUV; P[3nv]; N[3nv]; uv[2nv]; idx[3nt])
// ---------------------------------------------------------------------------
struct SceneDesc {
    int xres = 320, yres = 180, spp = 4;
    float lookat[9] = {0, -5, 0, 0, 0, 0, 0, 0, 1};
    float fov = 30.f;
    std::map<std::string, float> integ;
    struct Light { float t[3], radius, L[3]; int ns; int kind = 0; float to[3] = {0, 0, 1}, cone = 30.f, delta = 5.f; std::string map; float rotz = 0.f; };   // kind 0 sphere area, 1 point, 2 spot, 3 distant, 4 infinite
    std::vector<Light> lights;
    std::map<std::string, float> skin;
    float layers[4] = {0.25e6f, 1.4f, 20e6f, 1.4f};
    float Kr[3] = {1, 1, 1}, Kt[3] = {0, 0, 0};     // "color Kr" / "color Kt" as the skin scenes give them
    float albedo[3] = {1, 1, 1}; bool hasAlbedo = false;      // "color albedo" (materials/layeredskin.cpp:247)
    std::string albedoTex;                                     // "texture albedo": PFM of a "color" image map
    std::string bumpTex;                                       // "texture bumpmap": PFM of a "float" image map
    struct Xf { std::string kind; float v[4]; };
    struct Mesh { std::vector<Xf> xf; std::string file; std::string alpha; };      // alpha: PFM of the mesh's "texture alpha" (float imagemap)
    std::vector<Mesh> meshes;
};

static SceneDesc parseScene(const std::string &path) {
    SceneDesc sd;
    std::ifstream in(path);
    if (!in) { fprintf(stderr, "cannot open scene %s\n", path.c_str()); exit(2); }
    std::string dir = path.substr(0, path.find_last_of('/') + 1);
    std::string line;
    std::vector<SceneDesc::Xf> pending;
    std::string pendingAlpha;
    while (std::getline(in, line)) {
        std::istringstream ls(line);
        std::string key;
        if (!(ls >> key) || key[0] == '#') continue;
        if (key == "film") ls >> sd.xres >> sd.yres;
        else if (key == "lookat") { for (int i = 0; i < 9; ++i) ls >> sd.lookat[i]; }
        else if (key == "fov") ls >> sd.fov;
        else if (key == "spp") ls >> sd.spp;
        else if (key == "integrator") { std::string k; float v; while (ls >> k >> v) sd.integ[k] = v; }
        else if (key == "light") {
            SceneDesc::Light l;
            ls >> l.t[0] >> l.t[1] >> l.t[2] >> l.radius >> l.L[0] >> l.L[1] >> l.L[2] >> l.ns;
            sd.lights.push_back(l);
        }
        else if (key == "pointlight" || key == "spotlight" || key == "distantlight") {
            // pointlight fx fy fz r g b | spotlight fx fy fz tx ty tz r g b coneangle conedeltaangle | distantlight fx fy fz tx ty tz r g b
            SceneDesc::Light l; l.radius = 0.f; l.ns = 1;
            l.kind = key == "pointlight" ? 1 : key == "spotlight" ? 2 : 3;
            ls >> l.t[0] >> l.t[1] >> l.t[2];
            if (l.kind != 1) ls >> l.to[0] >> l.to[1] >> l.to[2];
            ls >> l.L[0] >> l.L[1] >> l.L[2];
            if (l.kind == 2) ls >> l.cone >> l.delta;
            sd.lights.push_back(l);
        }
        else if (key == "envlight") {
            // envlight <map.pfm | none> r g b nsamples rotz : LightSource "infinite" "color L" [r g b] "integer nsamples" "string mapname",
            // under Rotate rotz 0 0 1 (lights/infinite.cpp:185-193)
            SceneDesc::Light l; l.kind = 4; l.radius = 0.f; l.t[0] = l.t[1] = l.t[2] = 0.f;
            ls >> l.map >> l.L[0] >> l.L[1] >> l.L[2] >> l.ns >> l.rotz;
            if (l.map == "none") l.map.clear(); else if (l.map[0] != '/') l.map = dir + l.map;
            sd.lights.push_back(l);
        }
        else if (key == "skin") {
            std::string k;
            while (ls >> k) {
                if (k == "layers") { for (int i = 0; i < 4; ++i) ls >> sd.layers[i]; }
                else if (k == "Kr") { for (int i = 0; i < 3; ++i) ls >> sd.Kr[i]; }
                else if (k == "Kt") { for (int i = 0; i < 3; ++i) ls >> sd.Kt[i]; }
                else if (k == "albedo") { for (int i = 0; i < 3; ++i) ls >> sd.albedo[i]; sd.hasAlbedo = true; }
                else if (k == "albedotex") { ls >> sd.albedoTex; if (sd.albedoTex[0] != '/') sd.albedoTex = dir + sd.albedoTex; }
                else if (k == "bumptex") { ls >> sd.bumpTex; if (sd.bumpTex[0] != '/') sd.bumpTex = dir + sd.bumpTex; }
                else { float v; ls >> v; sd.skin[k] = v; }
            }
        }
        else if (key == "xform") {
            SceneDesc::Xf x; ls >> x.kind;
            int n = x.kind == "rotate" ? 4 : 3;
            for (int i = 0; i < n; ++i) ls >> x.v[i];
            pending.push_back(x);
        }
        else if (key == "alpha") { ls >> pendingAlpha; if (pendingAlpha[0] != '/') pendingAlpha = dir + pendingAlpha; }
        else if (key == "mesh") {
            SceneDesc::Mesh m; ls >> m.file;
            if (m.file[0] != '/') m.file = dir + m.file;
            m.alpha = pendingAlpha; pendingAlpha.clear();
            m.xf = pending; pending.clear();
            sd.meshes.push_back(m);
        }
        else { fprintf(stderr, "scene: unknown key %s\n", key.c_str()); exit(2); }
    }
    return sd;
}

static void addF(ParamSet &ps, const char *n, float v) { ps.AddFloat(n, &v, 1); }
static void addI(ParamSet &ps, const char *n, int v) { ps.AddInt(n, &v, 1); }

// Issues the same pbrt*() calls the reference's parser would for this scene.
static void buildScene(const SceneDesc &sd, const std::string &pfmOut) {
    {
        ParamSet ps; addI(ps, "xresolution", sd.xres); addI(ps, "yresolution", sd.yres);
        std::string fn = pfmOut; ps.AddString("filename", &fn, 1);
        pbrtFilm("image", ps);
    }
    pbrtLookAt(sd.lookat[0], sd.lookat[1], sd.lookat[2], sd.lookat[3], sd.lookat[4], sd.lookat[5],
               sd.lookat[6], sd.lookat[7], sd.lookat[8]);
    { ParamSet ps; addF(ps, "fov", sd.fov); pbrtCamera("perspective", ps); }
    { ParamSet ps; addI(ps, "pixelsamples", sd.spp); pbrtSampler("lowdiscrepancy", ps); }
    {
        ParamSet ps;
        for (auto &kv : sd.integ) {
            if (kv.first == "maxdepth") addI(ps, "maxdepth", (int)kv.second);
            else if (kv.first == "usepoissonpointfinder" || kv.first == "incenter") { bool b = kv.second != 0.f; ps.AddBool(kv.first.c_str(), &b, 1); }
            else addF(ps, kv.first.c_str(), kv.second);
        }
        pbrtSurfaceIntegrator("multipolesubsurface", ps);
    }
    pbrtWorldBegin();
    for (auto &l : sd.lights) {
        pbrtAttributeBegin();
        if (l.kind == 4) {
            ParamSet ps;
            ps.AddRGBSpectrum("L", l.L, 3); addI(ps, "nsamples", l.ns);
            if (!l.map.empty()) ps.AddString("mapname", &l.map, 1);
            pbrtRotate(l.rotz, 0.f, 0.f, 1.f);
            pbrtLightSource("infinite", ps);
            pbrtAttributeEnd();
            continue;
        }
        if (l.kind != 0) {
            ParamSet ps;
            Point from(l.t[0], l.t[1], l.t[2]), to(l.to[0], l.to[1], l.to[2]);
            ps.AddRGBSpectrum(l.kind == 3 ? "L" : "I", l.L, 3);
            ps.AddPoint("from", &from, 1);
            if (l.kind != 1) ps.AddPoint("to", &to, 1);
            if (l.kind == 2) { addF(ps, "coneangle", l.cone); addF(ps, "conedeltaangle", l.delta); }
            pbrtLightSource(l.kind == 1 ? "point" : l.kind == 2 ? "spot" : "distant", ps);
            pbrtAttributeEnd();
            continue;
        }
        { ParamSet ps; ps.AddRGBSpectrum("L", l.L, 3); addI(ps, "nsamples", l.ns); pbrtAreaLightSource("area", ps); }
        pbrtTranslate(l.t[0], l.t[1], l.t[2]);
        { ParamSet ps; addF(ps, "radius", l.radius); pbrtShape("sphere", ps); }
        pbrtAttributeEnd();
    }
    for (auto &m : sd.meshes) {
        pbrtAttributeBegin();
        for (auto &x : m.xf) {
            if (x.kind == "translate") pbrtTranslate(x.v[0], x.v[1], x.v[2]);
            else if (x.kind == "rotate") pbrtRotate(x.v[0], x.v[1], x.v[2], x.v[3]);
            else if (x.kind == "scale") pbrtScale(x.v[0], x.v[1], x.v[2]);
        }
        {
            ParamSet ps;
            for (auto &kv : sd.skin) {
                if (kv.first == "desiredlength") addI(ps, "desiredlength", (int)kv.second);
                else addF(ps, kv.first.c_str(), kv.second);
            }
            ps.AddRGBSpectrum("Kr", sd.Kr, 3); ps.AddRGBSpectrum("Kt", sd.Kt, 3);
            if (sd.hasAlbedo) ps.AddRGBSpectrum("albedo", sd.albedo, 3);
            if (!sd.albedoTex.empty()) {      // Texture "albedomap" "color" "imagemap" "string filename" + "texture albedo" (materials/layeredskin.cpp:247)
                ParamSet tp; tp.AddString("filename", &sd.albedoTex, 1);
                pbrtTexture("albedomap", "color", "imagemap", tp);
                std::string nm = "albedomap"; ps.AddTexture("albedo", nm);
            }
            if (!sd.bumpTex.empty()) {        // Texture "bumpmap" "float" "imagemap" + "texture bumpmap" (materials/layeredskin.cpp:246)
                ParamSet tp; tp.AddString("filename", &sd.bumpTex, 1);
                pbrtTexture("skinbump", "float", "imagemap", tp);
                std::string nm = "skinbump"; ps.AddTexture("bumpmap", nm);
            }
            SkinLayer sl[2]; sl[0].thickness = sd.layers[0]; sl[0].ior = sd.layers[1];
            sl[1].thickness = sd.layers[2]; sl[1].ior = sd.layers[3];
            ps.AddSkinLayer("layers", sl, 2);
            pbrtMaterial("layeredskin", ps);
        }
        {
            FILE *f = fopen(m.file.c_str(), "rb");
            if (!f) { fprintf(stderr, "cannot open mesh %s\n", m.file.c_str()); exit(2); }
            int32_t hdr[4];
            if (fread(hdr, 4, 4, f) != 4) exit(2);
            int nv = hdr[0], nt = hdr[1];
            std::vector<Point> P(nv); std::vector<Normal> N(nv); std::vector<float> uv(2 * nv); std::vector<int> idx(3 * nt);
            if (fread(P.data(), 12, nv, f) != (size_t)nv) exit(2);
            if (hdr[2] && fread(N.data(), 12, nv, f) != (size_t)nv) exit(2);
            if (hdr[3] && fread(uv.data(), 8, nv, f) != (size_t)nv) exit(2);
            if (fread(idx.data(), 12, nt, f) != (size_t)nt) exit(2);
            fclose(f);
            ParamSet ps;
            if (!m.alpha.empty()) {      // Texture "alphamap" "float" "imagemap" "string filename" + "texture alpha" (shapes/trianglemesh.cpp:411-421)
                ParamSet tp; tp.AddString("filename", &m.alpha, 1);
                pbrtTexture("alphamap", "float", "imagemap", tp);
                std::string nm = "alphamap"; ps.AddTexture("alpha", nm);
            }
            ps.AddInt("indices", idx.data(), 3 * nt);
            ps.AddPoint("P", P.data(), nv);
            if (hdr[2]) ps.AddNormal("N", N.data(), nv);
            if (hdr[3]) ps.AddFloat("uv", uv.data(), 2 * nv);
            pbrtShape("trianglemesh", ps);
        }
        pbrtAttributeEnd();
    }
}

// ---------------------------------------------------------------------------
// Dumps.  All files are raw little-endian arrays; meta.txt lists counts.
// ---------------------------------------------------------------------------
struct Meta {
    std::ostringstream s;
    template <class T> void kv(const char *k, T v) { s << k << " " << v << "\n"; }
};

static void dumpTransform(std::vector<float> &out, const Transform &t) {
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) out.push_back(t.GetMatrix().m[i][j]);
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) out.push_back(t.GetInverseMatrix().m[i][j]);
}

static void dumpSpectrumConsts(const std::string &dir) {
    // SampledSpectrum::X/Y/Z (core/spectrum.cpp Init) + the CIE_Y integral used by ToXYZ.
    std::vector<float> v;
    for (int i = 0; i < nSpectralSamples; ++i) v.push_back(SampledSpectrum::X.c[i]);
    for (int i = 0; i < nSpectralSamples; ++i) v.push_back(SampledSpectrum::Y.c[i]);
    for (int i = 0; i < nSpectralSamples; ++i) v.push_back(SampledSpectrum::Z.c[i]);
    dump(dir + "/cie_xyz30.bin", v.data(), v.size());
    // Spectrum::FromRGB((1,1,1), SPECTRUM_REFLECTANCE): what "color Kr" [1 1 1] becomes (core/spectrum.cpp)
    float one[3] = {1.f, 1.f, 1.f};
    Spectrum w = Spectrum::FromRGB(one);
    dump(dir + "/rgb_unit30.bin", w.c, nSpectralSamples);
    // the binned basis spectra FromRGB(.., SPECTRUM_REFLECTANCE) mixes (SampledSpectrum::Init, core/spectrum.cpp): data for
    // pbrt-v2-skin_b200/host/skin_tables.h kRGBRefl2Spect30 (white, cyan, magenta, yellow, red, green, blue)
    std::vector<float> bases;
    const SampledSpectrum *bs[7] = {&SampledSpectrum::rgbRefl2SpectWhite, &SampledSpectrum::rgbRefl2SpectCyan, &SampledSpectrum::rgbRefl2SpectMagenta,
                                    &SampledSpectrum::rgbRefl2SpectYellow, &SampledSpectrum::rgbRefl2SpectRed, &SampledSpectrum::rgbRefl2SpectGreen,
                                    &SampledSpectrum::rgbRefl2SpectBlue};
    for (int k = 0; k < 7; ++k) for (int i = 0; i < nSpectralSamples; ++i) bases.push_back(bs[k]->c[i]);
    dump(dir + "/rgb_refl_bases30.bin", bases.data(), bases.size());
    // ... and the illuminant set (FromRGB(.., SPECTRUM_ILLUMINANT), what InfiniteAreaLight turns its map's texels into): kRGBIllum2Spect30
    std::vector<float> ill;
    const SampledSpectrum *is[7] = {&SampledSpectrum::rgbIllum2SpectWhite, &SampledSpectrum::rgbIllum2SpectCyan, &SampledSpectrum::rgbIllum2SpectMagenta,
                                    &SampledSpectrum::rgbIllum2SpectYellow, &SampledSpectrum::rgbIllum2SpectRed, &SampledSpectrum::rgbIllum2SpectGreen,
                                    &SampledSpectrum::rgbIllum2SpectBlue};
    for (int k = 0; k < 7; ++k) for (int i = 0; i < nSpectralSamples; ++i) ill.push_back(is[k]->c[i]);
    dump(dir + "/rgb_illum_bases30.bin", ill.data(), ill.size());
}

static void dumpMaterial(const std::string &dir, const LayeredSkin *skin, Meta &meta) {
    if (const ImageTexture<float, float> *it = dynamic_cast<const ImageTexture<float, float> *>(skin->bumpMap.GetPtr())) {
        // "texture bumpmap": the MIPMap<float> pyramid -- int32 nLevels, then per level int32 w, h and [h][w] floats
        std::vector<float> a;
        auto pushInt = [&](int32_t v) { float f; memcpy(&f, &v, 4); a.push_back(f); };
        pushInt((int32_t)it->mipmap->nLevels);
        for (uint32_t l = 0; l < it->mipmap->nLevels; ++l) {
            const int w = (int)it->mipmap->pyramid[l]->uSize(), h = (int)it->mipmap->pyramid[l]->vSize();
            pushInt(w); pushInt(h);
            for (int t = 0; t < h; ++t) for (int s2 = 0; s2 < w; ++s2) a.push_back(it->mipmap->Texel(l, s2, t));
        }
        dump(dir + "/bump_tex.bin", a.data(), a.size());
    }
    if (const ImageTexture<RGBSpectrum, Spectrum> *it = dynamic_cast<const ImageTexture<RGBSpectrum, Spectrum> *>(skin->albedo.GetPtr())) {
        // "texture albedo": the MIPMap's whole pyramid -- int32 nLevels, then per level int32 w, h and rgb[h][w][3] -- and the EWA weight table
        std::vector<float> a;
        auto pushInt = [&](int32_t v) { float f; memcpy(&f, &v, 4); a.push_back(f); };
        pushInt((int32_t)it->mipmap->nLevels);
        for (uint32_t l = 0; l < it->mipmap->nLevels; ++l) {
            const int w = (int)it->mipmap->pyramid[l]->uSize(), h = (int)it->mipmap->pyramid[l]->vSize();
            pushInt(w); pushInt(h);
            for (int t = 0; t < h; ++t) for (int s2 = 0; s2 < w; ++s2) { float c3[3]; it->mipmap->Texel(l, s2, t).ToRGB(c3); a.push_back(c3[0]); a.push_back(c3[1]); a.push_back(c3[2]); }
        }
        for (int i = 0; i < WEIGHT_LUT_SIZE; ++i) a.push_back(MIPMap<RGBSpectrum>::weightLut[i]);
        dump(dir + "/albedo_tex.bin", a.data(), a.size());
    }
    const MultipoleBSSRDFData *bd = skin->preparedBSSRDFData;
    std::vector<float> coeff;   // [layer][mua30 | musp30], then eta[2], thickness[2]
    for (int l = 0; l < 2; ++l) {
        for (int i = 0; i < nSpectralSamples; ++i) coeff.push_back(bd->sig_a[l].c[i]);
        for (int i = 0; i < nSpectralSamples; ++i) coeff.push_back(bd->sigp_s[l].c[i]);
    }
    for (int l = 0; l < 2; ++l) coeff.push_back(bd->e[l]);
    for (int l = 0; l < 2; ++l) coeff.push_back(bd->d[l]);
    dump(dir + "/skin_coeffs.bin", coeff.data(), coeff.size());
    const DrvProfileData *pd = reinterpret_cast<const DrvProfileData *>(skin->profileData);
    std::vector<float> hdr, tab;
    uint32_t len = (uint32_t)pd->spectralProfile[0].data.size();
    for (int i = 0; i < nSpectralSamples; ++i) {
        const DrvProfile &p = pd->spectralProfile[i];
        if (p.data.size() != len) { fprintf(stderr, "profile lengths differ\n"); exit(2); }
        hdr.push_back(p.dsqSpacing); hdr.push_back(p.rcpDsqSpacing);
        hdr.push_back(p.totalReflectance); hdr.push_back(p.totalTransmittance);
        for (uint32_t k = 0; k < len; ++k) tab.push_back(p.data[k].reflectance);
    }
    meta.kv("profile_len", len);
    dump(dir + "/profile_hdr.bin", hdr.data(), hdr.size());     // [30][4]
    dump(dir + "/profile_rd.bin", tab.data(), tab.size());      // [30][len]
    std::vector<float> rho;
    for (size_t k = 0; k < bd->rhoData.hd.size(); ++k)
        for (int i = 0; i < nSpectralSamples; ++i) rho.push_back(bd->rhoData.hd[k].c[i]);
    for (int i = 0; i < nSpectralSamples; ++i) rho.push_back(bd->rhoData.hh.c[i]);
    meta.kv("rho_len", bd->rhoData.hd.size());
    dump(dir + "/rho_hd.bin", rho.data(), rho.size());          // [1025][30] + hh[30]
}

static uint32_t countNodes(const DrvLinearBVHNode *nodes, uint32_t i) {
    if (nodes[i].nPrimitives > 0) return 1;
    return 1 + countNodes(nodes, i + 1) + countNodes(nodes, nodes[i].secondChildOffset);
}

struct DrvPrim { int32_t type, mesh, index, primitiveId; };   // type 0 = triangle, 1 = sphere (index = light no.)

static void dumpBVH(const std::string &dir, const Scene *scene, Meta &meta,
                    std::vector<const TriangleMesh *> &meshes) {
    const BVHAccel *bvh = dynamic_cast<const BVHAccel *>(scene->aggregate);
    if (!bvh) { fprintf(stderr, "aggregate is not a BVHAccel\n"); exit(2); }
    const DrvLinearBVHNode *nodes = reinterpret_cast<const DrvLinearBVHNode *>(bvh->nodes);
    uint32_t nNodes = countNodes(nodes, 0);
    dump(dir + "/bvh_nodes.bin", nodes, nNodes);
    std::vector<DrvPrim> prims;
    // sphere index = light number (the emitter shapes are the only quadrics in these scenes)
    std::vector<const Sphere *> spheres;
    for (const Light *l : scene->lights) {
        const DiffuseAreaLight *dl = dynamic_cast<const DiffuseAreaLight *>(l);
        if (dl && dl->shapeSet->shapes.size() == 1)
            spheres.push_back(dynamic_cast<const Sphere *>(dl->shapeSet->shapes[0].GetPtr()));
        else spheres.push_back(NULL);
    }
    for (auto &pr : bvh->primitives) {
        const GeometricPrimitive *gp = dynamic_cast<const GeometricPrimitive *>(pr.GetPtr());
        if (!gp) { fprintf(stderr, "non-geometric primitive in BVH\n"); exit(2); }
        const Shape *sh = gp->shape.GetPtr();
        DrvPrim dp; dp.primitiveId = (int32_t)pr->primitiveId;
        if (const Triangle *tri = dynamic_cast<const Triangle *>(sh)) {
            const TriangleMesh *m = tri->mesh.GetPtr();
            size_t mi = std::find(meshes.begin(), meshes.end(), m) - meshes.begin();
            if (mi == meshes.size()) meshes.push_back(m);
            dp.type = 0; dp.mesh = (int32_t)mi; dp.index = (int32_t)((tri->v - m->vertexIndex) / 3);
        } else if (const Sphere *sp = dynamic_cast<const Sphere *>(sh)) {
            size_t si = std::find(spheres.begin(), spheres.end(), sp) - spheres.begin();
            if (si == spheres.size()) { fprintf(stderr, "sphere primitive is not a light shape\n"); exit(2); }
            dp.type = 1; dp.mesh = -1; dp.index = (int32_t)si;
        } else { fprintf(stderr, "unsupported shape in BVH\n"); exit(2); }
        prims.push_back(dp);
    }
    dump(dir + "/bvh_prims.bin", prims.data(), prims.size());
    meta.kv("bvh_nodes", nNodes);
    meta.kv("bvh_prims", prims.size());
    // World-space meshes exactly as the reference holds them (shapes/trianglemesh.cpp:72-73).
    meta.kv("meshes", meshes.size());
    for (size_t mi = 0; mi < meshes.size(); ++mi) {
        const TriangleMesh *m = meshes[mi];
        char nm[64];
        snprintf(nm, sizeof nm, "/mesh%zu_", mi);
        std::string b = dir + nm;
        dump(b + "P.bin", m->p, m->nverts);
        if (m->n) dump(b + "N.bin", m->n, m->nverts);
        if (m->uvs) dump(b + "uv.bin", m->uvs, 2 * m->nverts);
        dump(b + "idx.bin", m->vertexIndex, 3 * m->ntris);
        std::vector<float> xf; dumpTransform(xf, *m->ObjectToWorld);
        dump(b + "o2w.bin", xf.data(), xf.size());
        if (m->alphaTexture) {      // its MIPMap's level-0 texels: int32 w, h, then [h][w] floats
            const ImageTexture<float, float> *it = dynamic_cast<const ImageTexture<float, float> *>(m->alphaTexture.GetPtr());
            if (!it) { fprintf(stderr, "alpha texture is not a float image map\n"); exit(2); }
            const int w = (int)it->mipmap->width, h = (int)it->mipmap->height;
            std::vector<float> a; int32_t wh[2] = {w, h}; float whf[2]; memcpy(whf, wh, 8); a.push_back(whf[0]); a.push_back(whf[1]);
            for (int t = 0; t < h; ++t) for (int s2 = 0; s2 < w; ++s2) a.push_back(it->mipmap->Texel(0, s2, t));
            dump(b + "alpha.bin", a.data(), a.size());
        }
        meta.kv((std::string("mesh") + std::to_string(mi) + "_nverts").c_str(), m->nverts);
        meta.kv((std::string("mesh") + std::to_string(mi) + "_ntris").c_str(), m->ntris);
        meta.kv((std::string("mesh") + std::to_string(mi) + "_hasN").c_str(), m->n ? 1 : 0);
        meta.kv((std::string("mesh") + std::to_string(mi) + "_hasUV").c_str(), m->uvs ? 1 : 0);
        meta.kv((std::string("mesh") + std::to_string(mi) + "_reverse").c_str(), (int)m->ReverseOrientation);
        meta.kv((std::string("mesh") + std::to_string(mi) + "_swaps").c_str(), (int)m->TransformSwapsHandedness);
    }
}

static void dumpLights(const std::string &dir, const Scene *scene, Meta &meta) {
    // Per light: o2w(16) w2o(16) radius zmin zmax phiMax nSamples reverse Lemit[30]
    std::vector<float> out;
    for (const Light *l : scene->lights) {
        // delta lights in the record convention of include/skin_cuda.h: radius -1 / -2 / -3, LightToWorld, cone cosines or direction
        if (const PointLight *pl = dynamic_cast<const PointLight *>(l)) {
            dumpTransform(out, pl->LightToWorld);
            float r[8] = {-1.f, 0, 0, 0, 1.f, 0, 0, 0}; out.insert(out.end(), r, r + 8);
            for (int i = 0; i < nSpectralSamples; ++i) out.push_back(pl->Intensity.c[i]);
            continue;
        }
        if (const SpotLight *sl = dynamic_cast<const SpotLight *>(l)) {
            dumpTransform(out, sl->LightToWorld);
            float r[8] = {-2.f, sl->cosTotalWidth, sl->cosFalloffStart, 0, 1.f, 0, 0, 0}; out.insert(out.end(), r, r + 8);
            for (int i = 0; i < nSpectralSamples; ++i) out.push_back(sl->Intensity.c[i]);
            continue;
        }
        if (const DistantLight *dd = dynamic_cast<const DistantLight *>(l)) {
            dumpTransform(out, dd->LightToWorld);
            float r[8] = {-3.f, dd->lightDir.x, dd->lightDir.y, dd->lightDir.z, 1.f, 0, 0, 0}; out.insert(out.end(), r, r + 8);
            for (int i = 0; i < nSpectralSamples; ++i) out.push_back(dd->L.c[i]);
            continue;
        }
        if (const InfiniteAreaLight *il = dynamic_cast<const InfiniteAreaLight *>(l)) {
            // radius -4; zmin / zmax = the map's width / height; the map and its sampling distribution go to envmap.bin:
            // int32 w, h; level-0 texels rgb[h][w][3]; conditional func[h][w], cdf[h][w+1], funcInt[h]; marginal func[h], cdf[h+1], funcInt
            dumpTransform(out, il->LightToWorld);
            const int w = (int)il->radianceMap->width, h = (int)il->radianceMap->height;
            float r[8] = {-4.f, (float)w, (float)h, 0, (float)il->nSamples, 0, 0, 0}; out.insert(out.end(), r, r + 8);
            for (int i = 0; i < nSpectralSamples; ++i) out.push_back(0.f);
            std::vector<float> env;
            int32_t wh[2] = {w, h}; float whf[2]; memcpy(whf, wh, 8); env.push_back(whf[0]); env.push_back(whf[1]);
            for (int t = 0; t < h; ++t) for (int s2 = 0; s2 < w; ++s2) { float c3[3]; il->radianceMap->Texel(0, s2, t).ToRGB(c3); env.push_back(c3[0]); env.push_back(c3[1]); env.push_back(c3[2]); }
            const Distribution2D *d2 = il->distribution;
            for (int v = 0; v < h; ++v) env.insert(env.end(), d2->pConditionalV[v]->func, d2->pConditionalV[v]->func + w);
            for (int v = 0; v < h; ++v) env.insert(env.end(), d2->pConditionalV[v]->cdf, d2->pConditionalV[v]->cdf + w + 1);
            for (int v = 0; v < h; ++v) env.push_back(d2->pConditionalV[v]->funcInt);
            env.insert(env.end(), d2->pMarginal->func, d2->pMarginal->func + h);
            env.insert(env.end(), d2->pMarginal->cdf, d2->pMarginal->cdf + h + 1);
            env.push_back(d2->pMarginal->funcInt);
            dump(dir + "/envmap.bin", env.data(), env.size());
            continue;
        }
        const DiffuseAreaLight *dl = dynamic_cast<const DiffuseAreaLight *>(l);
        if (!dl || dl->shapeSet->shapes.size() != 1) { fprintf(stderr, "unsupported light\n"); exit(2); }
        const Sphere *sp = dynamic_cast<const Sphere *>(dl->shapeSet->shapes[0].GetPtr());
        if (!sp) { fprintf(stderr, "light shape is not a sphere\n"); exit(2); }
        dumpTransform(out, *sp->ObjectToWorld);
        out.push_back(sp->radius); out.push_back(sp->zmin); out.push_back(sp->zmax); out.push_back(sp->phiMax);
        out.push_back((float)dl->nSamples); out.push_back(sp->ReverseOrientation ? 1.f : 0.f);
        out.push_back(sp->thetaMin); out.push_back(sp->thetaMax);
        for (int i = 0; i < nSpectralSamples; ++i) out.push_back(dl->Lemit.c[i]);
    }
    meta.kv("lights", scene->lights.size());
    dump(dir + "/lights.bin", out.data(), out.size());      // [nLights][32+8+30]
}

// Same functor as integrators/multipolesubsurface.cpp:241-251 (file-local there).
class MultipoleReflectanceDrv {
public:
    MultipoleReflectanceDrv(const MultipoleBSSRDF *b) : bssrdf(b) {}
    Spectrum operator()(float d2) const { return bssrdf->reflectance(d2); }
private:
    const MultipoleBSSRDF *bssrdf;
};

// The subsurface part of the camera pass alone, on the reference's own thread pool: the sample loop of
// SamplerRendererTask::Run (renderers/samplerrenderer.cpp:60-167) with Li reduced to what
// MultipoleSubsurfaceIntegrator::Li does for the subsurface term (multipolesubsurface.cpp:262-287): Scene::Intersect,
// GetBSDF, GetMultipoleBSSRDF, octree->Mo, rho, clamp.  Every function called is the reference's; direct lighting and
// film accumulation are left out so that the timed work equals the device path's camera pass.
class SssOnlyTask : public Task {
public:
    SssOnlyTask(const Scene *sc, SamplerRenderer *r, Camera *c, Sampler *ms, Sample *sam, const MultipoleSubsurfaceIntegrator *in, int tn, int tc, double *sink)
        : scene(sc), sr(r), camera(c), mainSampler(ms), origSample(sam), integ(in), taskNum(tn), taskCount(tc), sink(sink) {}
    void Run() override {
        Sampler *sampler = mainSampler->GetSubSampler(taskNum, taskCount);
        if (!sampler) return;
        MemoryArena arena; RNG rng(taskNum);
        int maxSamples = sampler->MaximumSampleCount();
        Sample *samples = origSample->Duplicate(maxSamples);
        RayDifferential *rds = new RayDifferential[maxSamples];
        int sampleCount; double acc = 0.;
        while ((sampleCount = sampler->GetMoreSamples(samples, rng)) > 0) {
            for (int i = 0; i < sampleCount; ++i) {
                camera->GenerateRayDifferential(samples[i], &rds[i]);
                rds[i].ScaleDifferentials(1.f / sqrtf(sampler->samplesPerPixel));
                Intersection is;
                if (scene->Intersect(rds[i], &is)) {
                    Vector wo = -rds[i].d;
                    BSDF *bsdf = is.GetBSDF(rds[i], arena);
                    const Point &p = bsdf->dgShading.p; const Normal &n = bsdf->dgShading.nn;
                    const MultipoleBSSRDF *bssrdf = is.GetMultipoleBSSRDF(rds[i], arena);
                    if (bssrdf && integ->octree) {
                        MultipoleReflectanceDrv mr(bssrdf);
                        Spectrum Mo = integ->octree->Mo(integ->octreeBounds, p, n, mr, integ->maxError);
                        float costheta = min(AbsDot(wo, n), 1.f);
                        Spectrum Ft = Spectrum(1.f) - bssrdf->rho(costheta);
                        Spectrum sss = ((INV_PI * Ft) * Mo * Pow(bssrdf->albedo(), 1.f - integ->mix)).Clamp(0.f);
                        acc += sss.y();
                    }
                }
                arena.FreeAll();
            }
        }
        *sink = acc;      // keeps the work observable
        delete sampler; delete[] samples; delete[] rds;
    }
private:
    const Scene *scene; SamplerRenderer *sr; Camera *camera; Sampler *mainSampler; Sample *origSample;
    const MultipoleSubsurfaceIntegrator *integ; int taskNum, taskCount; double *sink;
};

struct DrvRay { float o[3], d[3], mint, maxt; };                       // 32 B
struct DrvHit { int32_t hit, prim; float t, rayEpsilon; float p[3], n[3], wo[3]; int32_t hasBssrdf; };  // 56 B

static int runScene(int argc, char **argv) {
    if (argc < 4) { fprintf(stderr, "usage: ref_driver scene <scene.txt> <outdir> [--ncores N] [--noexport] [--nocamera]\n"); return 2; }
    std::string scenePath = argv[2], dir = argv[3];
    int ncores = 1; bool doExport = true, doCamera = true;
    for (int i = 4; i < argc; ++i) {
        if (!strcmp(argv[i], "--ncores")) ncores = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--noexport")) doExport = false;
        else if (!strcmp(argv[i], "--nocamera")) doCamera = false;
    }
    Options options; options.nCores = ncores; options.quiet = true;
    pbrtInit(options);
    SceneDesc sd = parseScene(scenePath);
    Meta meta; meta.kv("ncores", ncores);
    double t0 = now_s();
    buildScene(sd, dir + "/image.pfm");     // LayeredSkin ctor builds profile + rho tables here
    double tMaterial = now_s() - t0;
    t0 = now_s();
    Renderer *renderer = renderOptions->MakeRenderer();
    Scene *scene = renderOptions->MakeScene();
    double tAccel = now_s() - t0;
    SamplerRenderer *sr = dynamic_cast<SamplerRenderer *>(renderer);
    MultipoleSubsurfaceIntegrator *integ = dynamic_cast<MultipoleSubsurfaceIntegrator *>(sr->surfaceIntegrator);
    Camera *camera = sr->camera;
    std::vector<const TriangleMesh *> meshes;
    if (doExport) {
        dumpSpectrumConsts(dir);
        for (auto &kv : scene->materials) {
            const LayeredSkin *skin = dynamic_cast<const LayeredSkin *>(kv.second.GetPtr());
            if (skin) { meta.kv("skin_material_id", kv.first); dumpMaterial(dir, skin, meta); break; }
        }
        dumpBVH(dir, scene, meta, meshes);
        dumpLights(dir, scene, meta);
        // Surface points exactly as Preprocess will regenerate them (deterministic tessellation; the Poisson finder is
        // deterministic with one task, i.e. --ncores 1).
        vector<SurfacePoint> pts;
        if (integ->usePoissonPointFinder) {
            Point pCamera = camera->CameraToWorld(camera->shutterOpen, Point(0, 0, 0));
            FindPoissonPointDistribution(pCamera, camera->shutterOpen, integ->minSampleDist, scene, &pts);
        } else
        GetSurfacePointsThroughTessellation(camera->shutterOpen, integ->minSampleDist, scene,
                                            &integ->originalPrimitives, &pts, integ->incenter);
        dump(dir + "/surface_points.bin", pts.data(), pts.size());
        meta.kv("surface_points", pts.size());
    }
    // ---- Preprocess: irradiance pass + octree (integrators/multipolesubsurface.cpp:170-238)
    t0 = now_s();
    sr->surfaceIntegrator->Preprocess(scene, camera, sr);
    double tPre = now_s() - t0;
    meta.kv("irradiance_points", integ->irradiancePoints.size());
    {
        int nPoints = (int)integ->irradiancePoints.size();
        int nTasks = RoundUpPow2((uint32_t)max(32 * NumSystemCores(), nPoints / 4096));
        meta.kv("irradiance_tasks", nTasks);
    }
    if (doExport) {
        dump(dir + "/irradiance_points.bin", integ->irradiancePoints.data(), integ->irradiancePoints.size());
        float ob[6] = {integ->octreeBounds.pMin.x, integ->octreeBounds.pMin.y, integ->octreeBounds.pMin.z,
                       integ->octreeBounds.pMax.x, integ->octreeBounds.pMax.y, integ->octreeBounds.pMax.z};
        dump(dir + "/octree_bounds.bin", ob, 6);
        if (integ->octree) {
            std::vector<float> root;
            root.push_back(integ->octree->p.x); root.push_back(integ->octree->p.y); root.push_back(integ->octree->p.z);
            root.push_back(integ->octree->sumArea);
            for (int i = 0; i < nSpectralSamples; ++i) root.push_back(integ->octree->Et.c[i]);
            dump(dir + "/octree_root.bin", root.data(), root.size());
        }
    }
    meta.kv("maxError", integ->maxError);
    meta.kv("mix", integ->mix);
    meta.kv("minSampleDist", integ->minSampleDist);
    // ---- Camera pass (renderers/samplerrenderer.cpp:59-167,191-225), single thread, task order
    double tRender = 0;
    if (doCamera && !doExport) {
        t0 = now_s();
        // timing path: the reference's own render loop (Preprocess already done -> skip it)
        Sample *sample = new Sample(sr->sampler, sr->surfaceIntegrator, sr->volumeIntegrator, scene);
        int nPixels = camera->film->xResolution * camera->film->yResolution;
        int nTasks = RoundUpPow2((uint32_t)max(32 * NumSystemCores(), nPixels / (16 * 16)));
        ProgressReporter reporter(nTasks, "Rendering");
        vector<Task *> tasks;
        for (int i = 0; i < nTasks; ++i)
            tasks.push_back(new SamplerRendererTask(scene, sr, camera, reporter, sr->sampler, sample, false,
                                                    nTasks - 1 - i, nTasks));
        EnqueueTasks(tasks); WaitForAllTasks();
        for (Task *t : tasks) delete t;
        reporter.Done();
        delete sample;
        camera->film->WriteImage();
        tRender = now_s() - t0;
    }
    if (doCamera && doExport) {
        t0 = now_s();
        const BVHAccel *bvh = dynamic_cast<const BVHAccel *>(scene->aggregate);
        std::unordered_map<const Primitive *, int> primIndex;
        for (size_t i = 0; i < bvh->primitives.size(); ++i) primIndex[bvh->primitives[i].GetPtr()] = (int)i;
        Sample *origSample = new Sample(sr->sampler, sr->surfaceIntegrator, sr->volumeIntegrator, scene);
        int nPixels = camera->film->xResolution * camera->film->yResolution;
        int nTasks = RoundUpPow2((uint32_t)max(32 * NumSystemCores(), nPixels / (16 * 16)));
        std::vector<DrvRay> rays; std::vector<DrvHit> hits; std::vector<float> imgxy, mo, Lout, Lsss;
        std::vector<float> smp;   // per camera sample: every oneD array then every twoD array (Sample::oneD/twoD, core/sampler.h)
        for (int ti = 0; ti < nTasks; ++ti) {
            int taskNum = nTasks - 1 - ti;
            Sampler *sampler = sr->sampler->GetSubSampler(taskNum, nTasks);
            if (!sampler) continue;
            MemoryArena arena; RNG rng(taskNum);
            int maxSamples = sampler->MaximumSampleCount();
            Sample *samples = origSample->Duplicate(maxSamples);
            RayDifferential *rds = new RayDifferential[maxSamples];
            Spectrum *Ls = new Spectrum[maxSamples];
            Intersection *isects = new Intersection[maxSamples];
            int sampleCount;
            while ((sampleCount = sampler->GetMoreSamples(samples, rng)) > 0) {
                for (int i = 0; i < sampleCount; ++i) {
                    float rayWeight = camera->GenerateRayDifferential(samples[i], &rds[i]);
                    rds[i].ScaleDifferentials(1.f / sqrtf(sampler->samplesPerPixel));
                    DrvRay dr = {{rds[i].o.x, rds[i].o.y, rds[i].o.z}, {rds[i].d.x, rds[i].d.y, rds[i].d.z}, rds[i].mint, rds[i].maxt};
                    rays.push_back(dr);
                    imgxy.push_back(samples[i].imageX); imgxy.push_back(samples[i].imageY);
                    for (size_t j = 0; j < samples[i].n1D.size(); ++j) for (uint32_t k = 0; k < samples[i].n1D[j]; ++k) smp.push_back(samples[i].oneD[j][k]);
                    for (size_t j = 0; j < samples[i].n2D.size(); ++j) for (uint32_t k = 0; k < 2 * samples[i].n2D[j]; ++k) smp.push_back(samples[i].twoD[j][k]);
                    DrvHit h; memset(&h, 0, sizeof h); h.prim = -1;
                    Spectrum Mo(0.f), sss(0.f);
                    {   // closest hit on a copy of the ray (Intersect shrinks maxt)
                        RayDifferential r2 = rds[i]; Intersection is;
                        if (scene->Intersect(r2, &is)) {
                            h.hit = 1; h.t = r2.maxt; h.rayEpsilon = is.rayEpsilon;
                            auto it = primIndex.find(is.primitive); h.prim = it == primIndex.end() ? -2 : it->second;
                            Vector wo = -r2.d;
                            BSDF *bsdf = is.GetBSDF(r2, arena);
                            const Point &p = bsdf->dgShading.p; const Normal &n = bsdf->dgShading.nn;
                            h.p[0] = p.x; h.p[1] = p.y; h.p[2] = p.z; h.n[0] = n.x; h.n[1] = n.y; h.n[2] = n.z;
                            h.wo[0] = wo.x; h.wo[1] = wo.y; h.wo[2] = wo.z;
                            const MultipoleBSSRDF *bssrdf = is.GetMultipoleBSSRDF(r2, arena);
                            if (bssrdf && integ->octree) {
                                h.hasBssrdf = 1;
                                MultipoleReflectanceDrv mr(bssrdf);
                                Mo = integ->octree->Mo(integ->octreeBounds, p, n, mr, integ->maxError);
                                float costheta = min(AbsDot(wo, n), 1.f);
                                Spectrum Ft = Spectrum(1.f) - bssrdf->rho(costheta);
                                sss = ((INV_PI * Ft) * Mo * Pow(bssrdf->albedo(), 1.f - integ->mix)).Clamp(0.f);
                            }
                        }
                    }
                    hits.push_back(h);
                    for (int c = 0; c < nSpectralSamples; ++c) { mo.push_back(Mo.c[c]); Lsss.push_back(sss.c[c]); }
                    if (rayWeight > 0.f) Ls[i] = rayWeight * sr->Li(scene, rds[i], &samples[i], rng, arena, &isects[i], NULL);
                    else Ls[i] = 0.f;
                    if (Ls[i].HasNaNs() || Ls[i].y() < -1e-5 || isinf(Ls[i].y())) Ls[i] = Spectrum(0.f);
                    for (int c = 0; c < nSpectralSamples; ++c) Lout.push_back(Ls[i].c[c]);
                    arena.FreeAll();
                }
                if (sampler->ReportResults(samples, rds, Ls, isects, sampleCount))
                    for (int i = 0; i < sampleCount; ++i) camera->film->AddSample(samples[i], Ls[i]);
                arena.FreeAll();
            }
            delete sampler; delete[] samples; delete[] rds; delete[] Ls; delete[] isects;
        }
        camera->film->WriteImage();
        tRender = now_s() - t0;
        dump(dir + "/camera_rays.bin", rays.data(), rays.size());
        dump(dir + "/camera_xy.bin", imgxy.data(), imgxy.size());
        dump(dir + "/camera_samples.bin", smp.data(), smp.size());
        meta.kv("sample_arrays_1d", origSample->n1D.size()); meta.kv("sample_arrays_2d", origSample->n2D.size());
        meta.kv("sample_floats", rays.empty() ? 0 : smp.size() / rays.size());
        dump(dir + "/hits.bin", hits.data(), hits.size());
        dump(dir + "/mo.bin", mo.data(), mo.size());
        dump(dir + "/L_sss.bin", Lsss.data(), Lsss.size());
        dump(dir + "/L.bin", Lout.data(), Lout.size());
        meta.kv("camera_samples", rays.size());
        meta.kv("camera_tasks", nTasks);
        std::vector<float> cam; dumpTransform(cam, camera->CameraToWorld.startTransform ? *camera->CameraToWorld.startTransform : Transform());
        dump(dir + "/camera_c2w.bin", cam.data(), cam.size());
    }
    meta.kv("t_material_s", tMaterial); meta.kv("t_accel_s", tAccel);
    meta.kv("t_preprocess_s", tPre); meta.kv("t_render_s", tRender);
    std::string ms = meta.s.str();
    dump(dir + "/meta.txt", ms.data(), ms.size());
    fputs(ms.c_str(), stdout);
    TasksCleanup();
    return 0;
}

// ---------------------------------------------------------------------------
// mpc mode: the reference's MPC_* DLL API on a list of layer stacks.
//   in.txt line:  nLayers desiredLength lerpOnThinSlab resample stepSize  {ior thickness mua musp} x nLayers
//   stepSize <= 0  ->  12*mfp/desiredLength with mfp as in core/multipole.cpp:247-264
//   out.bin per line: uint32 len; float totalR,totalT; float d2[len],R[len],T[len]
// ---------------------------------------------------------------------------
static int runMPC(int argc, char **argv) {
    if (argc < 4) { fprintf(stderr, "usage: ref_driver mpc <in.txt> <out.bin>\n"); return 2; }
    std::ifstream in(argv[2]);
    FILE *out = fopen(argv[3], "wb");
    if (!in || !out) { fprintf(stderr, "mpc: cannot open files\n"); return 2; }
    std::string line; int n = 0; double total = 0;
    while (std::getline(in, line)) {
        std::istringstream ls(line);
        int nLayers, desiredLength, lerp, resample; float stepSize;
        if (!(ls >> nLayers >> desiredLength >> lerp >> resample >> stepSize)) continue;
        std::vector<MPC_LayerSpec> specs(nLayers);
        float mfpTotal = 0.f;
        for (int l = 0; l < nLayers; ++l) {
            ls >> specs[l].ior >> specs[l].thickness >> specs[l].mua >> specs[l].musp;
            specs[l].g_HG = 0.f;
            mfpTotal += 1.f / (specs[l].mua + specs[l].musp);
        }
        MPC_Options opt; opt.desiredLength = desiredLength; opt.lerpOnThinSlab = lerp != 0;
        float mfp = mfpTotal / (float)nLayers;
        opt.desiredStepSize = stepSize > 0.f ? stepSize : 12.f * mfp / (float)opt.desiredLength;
        double t0 = now_s();
        MPC_Output *o = NULL;
        MPC_ComputeDiffusionProfile(nLayers, specs.data(), &opt, &o);
        // resample: 0 raw, 1 MPC_ResampleForUniformDistanceSquaredDistribution (default target), 3 the same with target length 1000,
        // 2 MPC_ResampleDistribution at 777 radii r_i = i * 1.05 * sqrt(extent) / 776 (the last ones lie beyond the table: zeros)
        if (resample == 1) MPC_ResampleForUniformDistanceSquaredDistribution(o);
        else if (resample == 3) MPC_ResampleForUniformDistanceSquaredDistribution(o, 1000);
        else if (resample == 2) {
            std::vector<float> r(777);
            float rmax = 1.05f * sqrtf(o->pDistanceSquared[o->length - 1]);
            for (int i = 0; i < 777; ++i) r[i] = (float)i * rmax / 776.f;
            MPC_ResampleDistribution(o, 777, r.data());
        }
        total += now_s() - t0;
        uint32_t len = o->length;
        fwrite(&len, 4, 1, out);
        fwrite(&o->totalReflectance, 4, 1, out); fwrite(&o->totalTransmittance, 4, 1, out);
        fwrite(o->pDistanceSquared, 4, len, out); fwrite(o->pReflectance, 4, len, out); fwrite(o->pTransmittance, 4, len, out);
        MPC_FreeOutput(o);
        ++n;
    }
    MPC_ClearCache();
    fclose(out);
    printf("mpc_profiles %d\nmpc_seconds %.6f\n", n, total);
    return 0;
}

// ---------------------------------------------------------------------------
// bench mode: the reference's own hot path, timed, K times in one process (material tables
// are built once, outside the timed steps -- they are not part of the per-frame hot path).
//   ref_driver bench <scene.txt> --ncores N --steps K --warmup W
// One step = MultipoleSubsurfaceIntegrator::Preprocess (tessellation + irradiance + octree)
// followed by the reference's own SamplerRendererTask loop (Scene::Intersect + Li per sample).
// Prints one line per step: "step i points P samples S preprocess_s a render_s b".
// ---------------------------------------------------------------------------
static int runBench(int argc, char **argv) {
    if (argc < 3) { fprintf(stderr, "usage: ref_driver bench <scene.txt> [--ncores N] [--steps K] [--warmup W]\n"); return 2; }
    std::string scenePath = argv[2];
    int ncores = 0, steps = 1, warmup = 0; bool sssOnly = false;
    for (int i = 3; i < argc; ++i) {
        if (!strcmp(argv[i], "--ncores")) ncores = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--steps")) steps = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--warmup")) warmup = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--sss-only")) sssOnly = true;
    }
    Options options; options.nCores = ncores; options.quiet = true;
    pbrtInit(options);
    SceneDesc sd = parseScene(scenePath);
    double t0 = now_s();
    buildScene(sd, "/dev/null");
    Renderer *renderer = renderOptions->MakeRenderer();
    Scene *scene = renderOptions->MakeScene();
    printf("cores %d\nsetup_s %.4f\n", NumSystemCores(), now_s() - t0);
    SamplerRenderer *sr = dynamic_cast<SamplerRenderer *>(renderer);
    MultipoleSubsurfaceIntegrator *integ = dynamic_cast<MultipoleSubsurfaceIntegrator *>(sr->surfaceIntegrator);
    Camera *camera = sr->camera;
    for (int it = 0; it < warmup + steps; ++it) {
        // reset what Preprocess accumulates (multipolesubsurface.cpp:146-150,218-237)
        integ->irradiancePoints.clear();
        integ->octreeArena.FreeAll();
        integ->octree = NULL;
        integ->octreeBounds = BBox();
        t0 = now_s();
        sr->surfaceIntegrator->Preprocess(scene, camera, sr);
        double tPre = now_s() - t0;
        t0 = now_s();
        Sample *sample = new Sample(sr->sampler, sr->surfaceIntegrator, sr->volumeIntegrator, scene);
        int nPixels = camera->film->xResolution * camera->film->yResolution;
        int nTasks = RoundUpPow2((uint32_t)max(32 * NumSystemCores(), nPixels / (16 * 16)));
        ProgressReporter reporter(nTasks, "Rendering");
        vector<Task *> tasks;
        std::vector<double> sinks(nTasks, 0.);
        for (int i = 0; i < nTasks; ++i) {
            if (sssOnly) tasks.push_back(new SssOnlyTask(scene, sr, camera, sr->sampler, sample, integ, nTasks - 1 - i, nTasks, &sinks[i]));
            else tasks.push_back(new SamplerRendererTask(scene, sr, camera, reporter, sr->sampler, sample, false, nTasks - 1 - i, nTasks));
        }
        EnqueueTasks(tasks); WaitForAllTasks();
        for (Task *t : tasks) delete t;
        reporter.Done();
        delete sample;
        double tRender = now_s() - t0;
        // every sample of the sampler's extent is generated, traced and shaded (Film::GetSampleExtent adds a row and a column)
        long nSamples = (long)(sr->sampler->xPixelEnd - sr->sampler->xPixelStart) * (sr->sampler->yPixelEnd - sr->sampler->yPixelStart) * sr->sampler->samplesPerPixel;
        double chk = 0.; for (double v : sinks) chk += v;
        printf("step %d points %zu samples %ld preprocess_s %.4f render_s %.4f %s checksum %.6g\n", it - warmup, integ->irradiancePoints.size(), nSamples, tPre, tRender,
               sssOnly ? "sss_only" : "full_li", chk);
        fflush(stdout);
    }
    TasksCleanup();
    return 0;
}

// ---------------------------------------------------------------------------
// trace mode: the reference's BVHAccel::Intersect / IntersectP (accelerators/bvh.cpp:383-502) on a file of rays, timed.
//   ref_driver trace <scene.txt> <rays.bin> [--any] [--ncores N] [--reps R]
// rays.bin: DrvRay records (o[3], d[3], mint, maxt).  The rays are cut into 4096-ray tasks for the reference's own task
// system.  Prints "rays N hits H seconds s" per repetition.
// ---------------------------------------------------------------------------
class TraceTask : public Task {
public:
    TraceTask(const Scene *sc, const DrvRay *r, size_t n, bool any, long *hits) : scene(sc), rays(r), n(n), any(any), hits(hits) {}
    void Run() override {
        long h = 0;
        for (size_t i = 0; i < n; ++i) {
            const DrvRay &d = rays[i];
            Ray ray(Point(d.o[0], d.o[1], d.o[2]), Vector(d.d[0], d.d[1], d.d[2]), d.mint, d.maxt);
            if (any) h += scene->IntersectP(ray) ? 1 : 0;
            else { Intersection is; h += scene->Intersect(ray, &is) ? 1 : 0; }
        }
        *hits = h;
    }
private:
    const Scene *scene; const DrvRay *rays; size_t n; bool any; long *hits;
};

static int runTrace(int argc, char **argv) {
    if (argc < 4) { fprintf(stderr, "usage: ref_driver trace <scene.txt> <rays.bin> [--any] [--ncores N] [--reps R]\n"); return 2; }
    int ncores = 0, reps = 1; bool any = false;
    for (int i = 4; i < argc; ++i) {
        if (!strcmp(argv[i], "--ncores")) ncores = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--reps")) reps = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--any")) any = true;
    }
    Options options; options.nCores = ncores; options.quiet = true;
    pbrtInit(options);
    SceneDesc sd = parseScene(argv[2]);
    buildScene(sd, "/dev/null");
    renderOptions->MakeRenderer();
    Scene *scene = renderOptions->MakeScene();
    std::ifstream f(argv[3], std::ios::binary | std::ios::ate);
    if (!f) { fprintf(stderr, "cannot read %s\n", argv[3]); return 1; }
    const size_t n = (size_t)f.tellg() / sizeof(DrvRay);
    std::vector<DrvRay> rays(n);
    f.seekg(0); f.read((char *)rays.data(), n * sizeof(DrvRay));
    printf("cores %d\n", NumSystemCores());
    const size_t chunk = 4096, nTasks = (n + chunk - 1) / chunk;
    for (int r = 0; r < reps; ++r) {
        std::vector<long> hits(nTasks, 0);
        vector<Task *> tasks;
        for (size_t t = 0; t < nTasks; ++t) tasks.push_back(new TraceTask(scene, rays.data() + t * chunk, std::min(chunk, n - t * chunk), any, &hits[t]));
        double t0 = now_s();
        EnqueueTasks(tasks); WaitForAllTasks();
        double dt = now_s() - t0;
        for (Task *t : tasks) delete t;
        long h = 0; for (long v : hits) h += v;
        printf("rays %zu hits %ld seconds %.4f %s\n", n, h, dt, any ? "any" : "closest");
        fflush(stdout);
    }
    TasksCleanup();
    return 0;
}

int main(int argc, char **argv) {
    if (argc < 2) { fprintf(stderr, "usage: ref_driver scene|mpc|bench|trace ...\n"); return 2; }
    if (!strcmp(argv[1], "bench")) return runBench(argc, argv);
    if (!strcmp(argv[1], "scene")) return runScene(argc, argv);
    if (!strcmp(argv[1], "mpc")) return runMPC(argc, argv);
    if (!strcmp(argv[1], "trace")) return runTrace(argc, argv);
    fprintf(stderr, "unknown mode %s\n", argv[1]);
    return 2;
}

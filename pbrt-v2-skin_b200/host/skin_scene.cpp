// skin_scene.cpp -- transforms, scene description, camera (host mirror; see skin_scene.h).
#include "skin_scene.h"
#include <cstdio>
#include <fstream>
#include <sstream>
#include "skin_tables.h"

namespace skin {

// Matrix inverse.  The camera and mesh matrices must equal the reference's bit for bit (camera rays and world-space vertices
// are compared exactly), and in fp32 that pins the elimination: in-place Gauss-Jordan with FULL pivoting, the pivot of each
// round being the LAST largest |entry| among rows and columns not used yet, the pivot row scaled by the reciprocal, every other
// row reduced by (row entry) x (scaled pivot row) -- core/transform.cpp:75-133 performs the same rounds.  Written over a flat
// array with a used-column mask and the list of pivot moves, which are undone in reverse at the end.
Mat4 Inverse(const Mat4 &m) {
    float a[16];
    memcpy(a, m.m, sizeof a);
    auto at = [&a](int r, int c) -> float & { return a[4 * r + c]; };
    struct Move { int from, to; } moves[4];
    unsigned pivoted = 0;                                    // bit c: column (= row, after the swap) c holds a finished pivot
    for (int round = 0; round < 4; ++round) {
        Move mv = {-1, -1};
        float largest = 0.f;
        for (int r = 0; r < 4; ++r) {
            if (pivoted >> r & 1u) continue;
            for (int c = 0; c < 4; ++c)
                if (!(pivoted >> c & 1u) && fabsf(at(r, c)) >= largest) { largest = fabsf(at(r, c)); mv.from = r; mv.to = c; }
        }
        pivoted |= 1u << mv.to;
        if (mv.from != mv.to) for (int c = 0; c < 4; ++c) std::swap(at(mv.from, c), at(mv.to, c));
        moves[round] = mv;
        const int p = mv.to;
        const float scale = 1.f / at(p, p);
        at(p, p) = 1.f;
        for (int c = 0; c < 4; ++c) at(p, c) *= scale;
        for (int r = 0; r < 4; ++r) {
            if (r == p) continue;
            const float factor = at(r, p);
            at(r, p) = 0.f;
            for (int c = 0; c < 4; ++c) at(r, c) -= at(p, c) * factor;
        }
    }
    for (int round = 3; round >= 0; --round)
        if (moves[round].from != moves[round].to) for (int r = 0; r < 4; ++r) std::swap(at(r, moves[round].from), at(r, moves[round].to));
    Mat4 out;
    memcpy(out.m, a, sizeof a);
    return out;
}

// scale on the diagonal, offset in the last column
static Mat4 scaleOffset(float sx, float sy, float sz, V3 t) {
    Mat4 r;
    r.m[0][0] = sx; r.m[1][1] = sy; r.m[2][2] = sz;
    r.m[0][3] = t.x; r.m[1][3] = t.y; r.m[2][3] = t.z;
    return r;
}
Transform Translate(V3 d) { return Transform(scaleOffset(1, 1, 1, d), scaleOffset(1, 1, 1, V3(-d.x, -d.y, -d.z))); }
Transform Scale(float x, float y, float z) { return Transform(scaleOffset(x, y, z, V3(0, 0, 0)), scaleOffset(1.f / x, 1.f / y, 1.f / z, V3(0, 0, 0))); }
// Rotation by `angle` degrees about `axis` (core/transform.cpp:213-241): u u^T (1 - cos) + cos on the diagonal, -/+ the third
// axis component times sin off it; the inverse is the transpose.
Transform Rotate(float angle, V3 axis) {
    const V3 unit = Normalize(axis);
    const float u[3] = {unit.x, unit.y, unit.z};
    const float sine = sinf(Radians(angle)), cosine = cosf(Radians(angle));
    Mat4 r;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            if (i == j) { r.m[i][i] = u[i] * u[i] + (1.f - u[i] * u[i]) * cosine; continue; }
            const int lo = i < j ? i : j, hi = i < j ? j : i, third = 3 - i - j;
            const float sym = u[lo] * u[hi] * (1.f - cosine), skew = u[third] * sine;
            r.m[i][j] = (j - i + 3) % 3 == 1 ? sym - skew : sym + skew;
        }
    return Transform(r, Transpose(r));
}
// Camera-to-world frame from an eye point, a target and an up hint (core/transform.cpp:244-265): columns = left, up', view
// direction, eye; the Transform returned is world-to-camera, i.e. (inverse, frame).
Transform LookAt(V3 pos, V3 look, V3 up) {
    const V3 view = Normalize(look - pos);
    const V3 left = Normalize(Cross(Normalize(up), view));
    const V3 upOrtho = Cross(view, left);
    const V3 col[4] = {left, upOrtho, view, pos};
    Mat4 frame;
    for (int c = 0; c < 4; ++c) {
        frame.m[0][c] = col[c].x; frame.m[1][c] = col[c].y; frame.m[2][c] = col[c].z;
        frame.m[3][c] = c == 3 ? 1.f : 0.f;
    }
    return Transform(Inverse(frame), frame);
}
// Perspective projection onto z in [0, 1] between the planes n and f, x/y scaled by cot(fov / 2) (core/transform.cpp:309-319)
Transform Perspective(float fov, float n, float f) {
    Mat4 proj;                                               // identity, then the two rows that matter
    proj.m[2][2] = f / (f - n); proj.m[2][3] = -f * n / (f - n);
    proj.m[3][2] = 1.f;         proj.m[3][3] = 0.f;
    const float cot = 1.f / tanf(Radians(fov) / 2.f);
    return Scale(cot, cot, 1) * Transform(proj);
}

// core/camera.cpp:83-102, cameras/perspective.cpp:124-155
PerspectiveCamera::PerspectiveCamera(const Transform &cam2world, float fov, int xres, int yres, float frameAspect, const float *screenWindow)
    : CameraToWorld(cam2world) {
    float frame = frameAspect > 0.f ? frameAspect : float(xres) / float(yres);
    float screen[4];
    if (frame > 1.f) { screen[0] = -frame; screen[1] = frame; screen[2] = -1.f; screen[3] = 1.f; }
    else { screen[0] = -1.f; screen[1] = 1.f; screen[2] = -1.f / frame; screen[3] = 1.f / frame; }
    if (screenWindow) memcpy(screen, screenWindow, sizeof screen);
    Transform CameraToScreen = Perspective(fov, 1e-2f, 1000.f);
    Transform ScreenToRaster = Scale(float(xres), float(yres), 1.f) *
                               Scale(1.f / (screen[1] - screen[0]), 1.f / (screen[2] - screen[3]), 1.f) *
                               Translate(V3(-screen[0], -screen[3], 0.f));
    Transform RasterToScreen = InverseT(ScreenToRaster);
    RasterToCamera = InverseT(CameraToScreen) * RasterToScreen;
}

void PerspectiveCamera::GenerateRay(float imageX, float imageY, skc_ray *ray) const {   // perspective.cpp:79-113
    V3 Pcamera = RasterToCamera.Point(V3(imageX, imageY, 0));
    V3 dir = Normalize(Pcamera);
    V3 o = CameraToWorld.Point(V3(0, 0, 0));
    V3 d = CameraToWorld.Vector(dir);
    ray->o[0] = o.x; ray->o[1] = o.y; ray->o[2] = o.z;
    ray->d[0] = d.x; ray->d[1] = d.y; ray->d[2] = d.z;
    ray->mint = 0.f; ray->maxt = INFINITY;
}

// ParamSet::AddRGBSpectrum -> Spectrum::FromRGB(rgb, SPECTRUM_REFLECTANCE) (core/paramset.cpp, core/spectrum.cpp:103-145,185):
// the smallest component's share of the white basis plus the two remaining differences on the cyan/magenta/yellow and
// red/green/blue bases, times .94, clamped at 0 -- same operation order (0 + a, then += b, += c; the .94 last).
void RGBToSpectrum(const float rgb[3], float out30[30]) {
    enum { W, C, M, Y, R, G, B };
    float a, b, c; int ib, ic;
    if (rgb[0] <= rgb[1] && rgb[0] <= rgb[2]) {
        a = rgb[0];
        if (rgb[1] <= rgb[2]) { b = rgb[1] - rgb[0]; ib = C; c = rgb[2] - rgb[1]; ic = B; }
        else { b = rgb[2] - rgb[0]; ib = C; c = rgb[1] - rgb[2]; ic = G; }
    } else if (rgb[1] <= rgb[0] && rgb[1] <= rgb[2]) {
        a = rgb[1];
        if (rgb[0] <= rgb[2]) { b = rgb[0] - rgb[1]; ib = M; c = rgb[2] - rgb[0]; ic = B; }
        else { b = rgb[2] - rgb[1]; ib = M; c = rgb[0] - rgb[2]; ic = R; }
    } else {
        a = rgb[2];
        if (rgb[0] <= rgb[1]) { b = rgb[0] - rgb[2]; ib = Y; c = rgb[1] - rgb[0]; ic = G; }
        else { b = rgb[1] - rgb[2]; ib = Y; c = rgb[0] - rgb[1]; ic = R; }
    }
    for (int i = 0; i < 30; ++i) {
        float r = 0.f;
        r += kRGBRefl2Spect30[W][i] * a;
        r += kRGBRefl2Spect30[ib][i] * b;
        r += kRGBRefl2Spect30[ic][i] * c;
        r *= .94f;
        out30[i] = r < 0.f ? 0.f : r;       // Clamp(0, INFINITY)
    }
}
void RGBIlluminantToSpectrum(const float rgb[3], float out30[30]) { RGBToSpectrum(rgb, out30); }

static void CoordinateSystemV(V3 v1, V3 *v2, V3 *v3) {      // core/geometry.h:564-574
    if (fabsf(v1.x) > fabsf(v1.y)) { float invLen = 1.f / sqrtf(v1.x * v1.x + v1.z * v1.z); *v2 = V3(-v1.z * invLen, 0.f, v1.x * invLen); }
    else { float invLen = 1.f / sqrtf(v1.y * v1.y + v1.z * v1.z); *v2 = V3(0.f, v1.z * invLen, -v1.y * invLen); }
    *v3 = Cross(v1, *v2);
}

SphereLight MakePointLight(const Transform &l2w, V3 from, const float rgb[3]) {
    SphereLight l; l.kind = 1; l.nSamples = 1; l.radius = 0.f;
    l.ObjectToWorld = Translate(from) * l2w;                  // lights/point.cpp:68 (this order)
    memcpy(l.Lrgb, rgb, 12);
    return l;
}
SphereLight MakeSpotLight(const Transform &l2w, V3 from, V3 to, const float rgb[3], float coneangle, float conedelta) {
    SphereLight l; l.kind = 2; l.nSamples = 1; l.radius = 0.f;
    V3 dir = Normalize(to - from), du, dv;
    CoordinateSystemV(dir, &du, &dv);
    Transform dirToZ(Mat4(du.x, du.y, du.z, 0.f, dv.x, dv.y, dv.z, 0.f, dir.x, dir.y, dir.z, 0.f, 0.f, 0.f, 0.f, 1.f));
    l.ObjectToWorld = l2w * Translate(from) * InverseT(dirToZ);          // lights/spot.cpp:97-98
    l.cosTotalWidth = cosf(Radians(coneangle));                           // SpotLight ctor, width = coneangle, fall = coneangle - conedelta
    l.cosFalloffStart = cosf(Radians(coneangle - conedelta));
    memcpy(l.Lrgb, rgb, 12);
    return l;
}
SphereLight MakeDistantLight(const Transform &l2w, V3 from, V3 to, const float rgb[3]) {
    SphereLight l; l.kind = 3; l.nSamples = 1; l.radius = 0.f;
    l.ObjectToWorld = l2w;
    l.lightDir = Normalize(l2w.Vector(from - to));
    memcpy(l.Lrgb, rgb, 12);
    return l;
}

skc_light SphereLight::Pack(const float *Lemit30) const {
    skc_light L;
    memcpy(L.o2w, ObjectToWorld.m.m, 64); memcpy(L.w2o, ObjectToWorld.mInv.m, 64);
    if (kind != 0) {      // delta lights in the sphere record (include/skin_cuda.h)
        L.radius = -(float)kind; L.zmin = L.zmax = L.phiMax = 0.f;
        if (kind == 2) { L.zmin = cosTotalWidth; L.zmax = cosFalloffStart; }
        if (kind == 3) { L.zmin = lightDir.x; L.zmax = lightDir.y; L.phiMax = lightDir.z; }
        if (kind == 4) {      // infinite: the radiance is the map's; Lemit is not used
            L.zmin = (float)envW; L.zmax = (float)envH; L.nSamples = (float)nSamples; L.reverseOrientation = 0.f; L.thetaMin = L.thetaMax = 0.f;
            for (int i = 0; i < 30; ++i) L.Lemit[i] = 0.f;
            return L;
        }
        L.nSamples = 1.f; L.reverseOrientation = 0.f; L.thetaMin = L.thetaMax = 0.f;
        for (int i = 0; i < 30; ++i) L.Lemit[i] = Lemit30[i];
        return L;
    }
    // Sphere ctor (shapes/sphere.cpp:40-50)
    float zmn = std::min(std::max(std::min(zmin, zmax), -radius), radius);
    float zmx = std::min(std::max(std::max(zmin, zmax), -radius), radius);
    L.radius = radius; L.zmin = zmn; L.zmax = zmx;
    L.thetaMin = acosf(std::min(std::max(zmn / radius, -1.f), 1.f));
    L.thetaMax = acosf(std::min(std::max(zmx / radius, -1.f), 1.f));
    L.phiMax = Radians(std::min(std::max(phiMaxDeg, 0.0f), 360.0f));
    L.nSamples = (float)nSamples; L.reverseOrientation = ReverseOrientation ? 1.f : 0.f;
    for (int i = 0; i < 30; ++i) L.Lemit[i] = Lemit30[i];
    return L;
}

BBox SphereLight::WorldBound() const {
    float zmn = std::min(std::max(std::min(zmin, zmax), -radius), radius);
    float zmx = std::min(std::max(std::max(zmin, zmax), -radius), radius);
    return ObjectToWorld.Box(BBox(V3(-radius, -radius, zmn), V3(radius, radius, zmx)));
}

void InitMesh(TriangleMesh *m, const Transform &o2w, bool reverseOrientation, int nverts, const float *P, const float *N, const float *uv,
              int ntris, const int *indices) {
    m->nverts = nverts; m->ntris = ntris;
    m->p.resize(nverts);
    for (int i = 0; i < nverts; ++i) m->p[i] = o2w.Point(V3(P[3 * i], P[3 * i + 1], P[3 * i + 2]));   // trianglemesh.cpp:72-73
    m->n.clear(); m->uvs.clear();
    if (N) { m->n.resize(nverts); memcpy((void *)m->n.data(), N, 12 * (size_t)nverts); }
    if (uv) m->uvs.assign(uv, uv + 2 * (size_t)nverts);
    m->vertexIndex.assign(indices, indices + 3 * (size_t)ntris);
    m->ObjectToWorld = o2w;
    m->ReverseOrientation = reverseOrientation;
    m->TransformSwapsHandedness = o2w.SwapsHandedness();
}

static bool loadMesh(const std::string &file, const Transform &o2w, TriangleMesh *m, std::string *err) {
    FILE *f = fopen(file.c_str(), "rb");
    if (!f) { *err = "cannot open mesh " + file; return false; }
    int32_t hdr[4];
    bool ok = fread(hdr, 4, 4, f) == 4;
    if (ok) {
        int nv = hdr[0], nt = hdr[1];
        std::vector<float> P(3 * (size_t)nv), N, uv;
        std::vector<int> idx;
        ok = fread(P.data(), 12, nv, f) == (size_t)nv;
        if (ok && hdr[2]) { N.resize(3 * (size_t)nv); ok = fread(N.data(), 12, nv, f) == (size_t)nv; }
        if (ok && hdr[3]) { uv.resize(2 * (size_t)nv); ok = fread(uv.data(), 8, nv, f) == (size_t)nv; }
        if (ok) { idx.resize(3 * (size_t)nt); ok = fread(idx.data(), 12, nt, f) == (size_t)nt; }
        if (ok) InitMesh(m, o2w, false, nv, P.data(), N.empty() ? nullptr : N.data(), uv.empty() ? nullptr : uv.data(), nt, idx.data());
    }
    fclose(f);
    if (!ok) { *err = "short read in mesh " + file; return false; }
    return true;
}

// ---- pixel filters (filters/*.cpp) ----------------------------------------------------------------------------------
float PixelFilter::Evaluate(float x, float y) const {
    const float invX = 1.f / xwidth, invY = 1.f / ywidth;     // core/filter.h:50-52
    if (name == "box") return 1.f;
    if (name == "triangle") return std::max(0.f, xwidth - fabsf(x)) * std::max(0.f, ywidth - fabsf(y));
    if (name == "gaussian") {                                  // filters/gaussian.{h,cpp}
        float expX = expf(-alpha * xwidth * xwidth), expY = expf(-alpha * ywidth * ywidth);
        float gx = std::max(0.f, float(expf(-alpha * x * x) - expX)), gy = std::max(0.f, float(expf(-alpha * y * y) - expY));
        return gx * gy;
    }
    if (name == "mitchell") {                                  // filters/mitchell.h:49-58
        auto m1 = [&](float v) {
            v = fabsf(2.f * v);
            if (v > 1.f) return ((-B - 6 * C) * v * v * v + (6 * B + 30 * C) * v * v + (-12 * B - 48 * C) * v + (8 * B + 24 * C)) * (1.f / 6.f);
            return ((12 - 9 * B - 6 * C) * v * v * v + (-18 + 12 * B + 6 * C) * v * v + (6 - 2 * B)) * (1.f / 6.f);
        };
        return m1(x * invX) * m1(y * invY);
    }
    // "sinc": filters/sinc.h:49-57 (M_PI is the float literal of core/pbrt.h:193-196)
    auto s1 = [&](float v) {
        v = fabsf(v);
        if (v < 1e-5) return 1.f;
        if (v > 1.) return 0.f;
        v *= 3.14159265358979323846f;
        float sinc = sinf(v) / v;
        float lanczos = sinf(v * tau) / (v * tau);
        return sinc * lanczos;
    };
    return s1(x * invX) * s1(y * invY);
}
void PixelFilter::Table(float out[256]) const {
    float *ftp = out;
    for (int y = 0; y < 16; ++y) {
        float fy = ((float)y + .5f) * ywidth / 16;
        for (int x = 0; x < 16; ++x) {
            float fx = ((float)x + .5f) * xwidth / 16;
            *ftp++ = Evaluate(fx, fy);
        }
    }
}

PerspectiveCameraDesc SceneDesc::Camera() const {
    PerspectiveCamera cam(InverseT(WorldToCamera), fov, xres, yres, frameAspect, haveScreenWindow ? screenWindow : nullptr);   // api.cpp:862
    PerspectiveCameraDesc d; d.CameraToWorld = cam.CameraToWorld; d.RasterToCamera = cam.RasterToCamera;
    return d;
}

bool SceneDesc::Load(const std::string &path, const LoadOptions *opt) {
    bool pbrt = path.size() >= 5 && path.compare(path.size() - 5, 5, ".pbrt") == 0;
    return pbrt ? LoadPbrt(path, opt) : LoadLines(path);
}

// Line format shared with oracle/refbuild/ref_driver.cpp ("Scene description")
bool SceneDesc::LoadLines(const std::string &path) {
    std::ifstream in(path);
    if (!in) { error = "cannot open scene " + path; return false; }
    std::string dir = path.substr(0, path.find_last_of('/') + 1), line;
    AlphaTexture pendingAlpha;
    Transform cur;   // CTM of the next mesh (pbrtTranslate etc. post-multiply: core/api.cpp:720-756)
    uint32_t nextMaterialId = 0;
    while (std::getline(in, line)) {
        std::istringstream ls(line);
        std::string key;
        if (!(ls >> key) || key[0] == '#') continue;
        if (key == "film") ls >> xres >> yres;
        else if (key == "lookat") { for (int i = 0; i < 9; ++i) ls >> lookat[i]; }
        else if (key == "fov") ls >> fov;
        else if (key == "spp") ls >> spp;
        else if (key == "integrator") {
            std::string k; float v;
            while (ls >> k >> v) {
                if (k == "maxdepth") maxdepth = (int)v; else if (k == "maxerror") maxerror = v;
                else if (k == "minsampledistance") minsampledistance = v; else if (k == "mix") mix = v;
                else if (k == "incenter") incenter = v != 0.f;
                else if (k == "usepoissonpointfinder") usePoissonPointFinder = v != 0.f;
                else { error = "unknown integrator parameter " + k; return false; }
            }
        } else if (key == "light") {
            SphereLight l; float t[3];
            ls >> t[0] >> t[1] >> t[2] >> l.radius >> l.Lrgb[0] >> l.Lrgb[1] >> l.Lrgb[2] >> l.nSamples;
            l.zmin = -l.radius; l.zmax = l.radius;
            l.ObjectToWorld = Transform() * Translate(V3(t[0], t[1], t[2]));
            lights.push_back(l);
            nextMaterialId++;   // each emitter shape gets the default matte material (api.cpp MakeMaterial)
        } else if (key == "pointlight" || key == "spotlight" || key == "distantlight") {
            // pointlight fx fy fz r g b | spotlight fx fy fz tx ty tz r g b coneangle conedeltaangle | distantlight fx fy fz tx ty tz r g b
            float f[3], t[3] = {0, 0, 1}, c[3], cone = 30.f, delta = 5.f;
            ls >> f[0] >> f[1] >> f[2];
            if (key != "pointlight") ls >> t[0] >> t[1] >> t[2];
            ls >> c[0] >> c[1] >> c[2];
            if (key == "spotlight") ls >> cone >> delta;
            V3 from(f[0], f[1], f[2]), to(t[0], t[1], t[2]);
            lights.push_back(key == "pointlight" ? MakePointLight(Transform(), from, c) : key == "spotlight" ? MakeSpotLight(Transform(), from, to, c, cone, delta)
                                                 : MakeDistantLight(Transform(), from, to, c));
        } else if (key == "envlight") {
            // envlight <map.pfm | none> r g b nsamples rotz: LightSource "infinite" under Rotate rotz 0 0 1
            std::string map; float c[3], rotz = 0.f; int ns = 1;
            ls >> map >> c[0] >> c[1] >> c[2] >> ns >> rotz;
            if (map == "none") map.clear(); else if (map[0] != '/') map = dir + map;
            if (env.present) { error = "more than one infinite light"; return false; }
            SphereLight l;
            if (!MakeInfiniteLight(Transform() * Rotate(rotz, V3(0, 0, 1)), c, nullptr, ns, map, false, &l, &env, &error)) return false;
            lights.push_back(l);
        } else if (key == "skin") {
            std::string k;
            while (ls >> k) {
                if (k == "layers") ls >> skin.layerThickness[0] >> skin.layerIor[0] >> skin.layerThickness[1] >> skin.layerIor[1];
                else if (k == "Kr") ls >> skin.Kr[0] >> skin.Kr[1] >> skin.Kr[2];
                else if (k == "Kt") ls >> skin.Kt[0] >> skin.Kt[1] >> skin.Kt[2];
                else if (k == "albedo") { ls >> skin.albedo[0] >> skin.albedo[1] >> skin.albedo[2]; skin.hasAlbedo = true; }
                else if (k == "albedotex") {      // Texture ".." "color" "imagemap" + "texture albedo"
                    std::string file; ls >> file;
                    if (file[0] != '/') file = dir + file;
                    if (!LoadAlbedoTexture(file, 0.f, 1.f, 1.f, &skin.albedoTex, &error)) return false;
                }
                else if (k == "bumptex") {        // Texture ".." "float" "imagemap" + "texture bumpmap"
                    std::string file; ls >> file;
                    if (file[0] != '/') file = dir + file;
                    if (!LoadBumpTexture(file, 0.f, 1.f, 1.f, &skin.bumpTex, &error)) return false;
                }
                else {
                    float v; ls >> v;
                    if (k == "roughness") skin.roughness = v; else if (k == "nmperunit") skin.nmperunit = v;
                    else if (k == "f_mel") skin.f_mel = v; else if (k == "f_eu") skin.f_eu = v;
                    else if (k == "f_blood") skin.f_blood = v; else if (k == "f_ohg") skin.f_ohg = v;
                    else if (k == "desiredlength") skin.desiredlength = (int)v; else if (k == "lerponthinslab") skin.lerponthinslab = v != 0.f;
                    else { error = "unknown skin parameter " + k; return false; }
                }
            }
        } else if (key == "xform") {
            std::string kind; float v[4] = {0, 0, 0, 0};
            ls >> kind;
            int n = kind == "rotate" ? 4 : 3;
            for (int i = 0; i < n; ++i) ls >> v[i];
            if (kind == "translate") cur = cur * Translate(V3(v[0], v[1], v[2]));
            else if (kind == "rotate") cur = cur * Rotate(v[0], V3(v[1], v[2], v[3]));
            else if (kind == "scale") cur = cur * Scale(v[0], v[1], v[2]);
            else { error = "unknown xform " + kind; return false; }
        } else if (key == "alpha") {
            // alpha <map.pfm>: Texture "a" "float" "imagemap" "string filename" + "texture alpha" "a" on the next mesh
            std::string file; ls >> file;
            if (file[0] != '/') file = dir + file;
            if (!LoadAlphaTexture(file, 0.f, 1.f, 1.f, &pendingAlpha, &error)) return false;
        } else if (key == "mesh") {
            std::string file; ls >> file;
            if (file[0] != '/') file = dir + file;
            TriangleMesh m;
            if (!loadMesh(file, cur, &m, &error)) return false;
            m.alpha = pendingAlpha; pendingAlpha = AlphaTexture();
            m.materialId = nextMaterialId++;
            meshes.push_back(std::move(m));
            cur = Transform();
        } else { error = "scene: unknown key " + key; return false; }
    }
    WorldToCamera = Transform() * LookAt(V3(lookat[0], lookat[1], lookat[2]), V3(lookat[3], lookat[4], lookat[5]), V3(lookat[6], lookat[7], lookat[8]));
    return true;
}

bool BuildScene(const SceneDesc &sd, const float *Lemit30PerLight, BuiltScene *out, std::string *err) {
    if (sd.meshes.empty()) { *err = "no layeredskin trianglemesh in the scene"; return false; }
    // The device holds ONE mesh.  Several Shape "trianglemesh" directives under the same layeredskin material are joined into
    // one (vertices are already in world space, indices are offset); this is exact as long as the per-mesh state the device
    // keeps once -- orientation flags, and ObjectToWorld when there are vertex normals to transform -- agrees.
    std::vector<int> triOffset(sd.meshes.size(), 0);
    if (sd.meshes.size() > 1) {
        TriangleMesh &m = out->merged;
        m = TriangleMesh();
        const TriangleMesh &m0 = sd.meshes[0];
        m.ObjectToWorld = m0.ObjectToWorld; m.ReverseOrientation = m0.ReverseOrientation; m.TransformSwapsHandedness = m0.TransformSwapsHandedness;
        m.materialId = m0.materialId;
        m.alpha = m0.alpha;
        for (size_t k = 0; k < sd.meshes.size(); ++k) {
            const TriangleMesh &mk = sd.meshes[k];
            if (!(mk.alpha == m0.alpha)) { *err = "several skin meshes with different alpha textures cannot be joined into the one mesh the device path holds"; return false; }
            if (mk.ReverseOrientation != m0.ReverseOrientation || mk.TransformSwapsHandedness != m0.TransformSwapsHandedness || mk.n.empty() != m0.n.empty() ||
                mk.uvs.empty() != m0.uvs.empty() ||
                (!m0.n.empty() && memcmp(&mk.ObjectToWorld, &m0.ObjectToWorld, sizeof(Transform)) != 0)) {
                *err = "several skin meshes that differ in orientation, in having normals / uvs, or (with normals) in their transform cannot be joined into the one mesh the device path holds";
                return false;
            }
            triOffset[k] = m.ntris;
            for (int v : mk.vertexIndex) m.vertexIndex.push_back(v + m.nverts);
            m.p.insert(m.p.end(), mk.p.begin(), mk.p.end());
            m.n.insert(m.n.end(), mk.n.begin(), mk.n.end());
            m.uvs.insert(m.uvs.end(), mk.uvs.begin(), mk.uvs.end());
            m.nverts += mk.nverts; m.ntris += mk.ntris;
        }
    }
    const TriangleMesh &mesh = sd.meshes.size() > 1 ? out->merged : sd.meshes[0];
    // renderOptions->primitives order: emitter spheres, then the mesh triangles (api.cpp:1066-1070)
    // (a .pbrt file fixes the order itself: primOrder)
    std::vector<BBox> bounds;
    std::vector<skc_prim> prims;
    std::vector<PrimRef> order = sd.primOrder;
    if (order.empty()) {
        for (size_t i = 0; i < sd.lights.size(); ++i) if (sd.lights[i].kind == 0) order.push_back(PrimRef{1, (int)i});
        for (size_t k = 0; k < sd.meshes.size(); ++k) order.push_back(PrimRef{0, (int)k});
    }
    for (const PrimRef &r : order) {
        if (r.kind == 1) {
            bounds.push_back(sd.lights[r.index].WorldBound());
            skc_prim p = {1, -1, (int32_t)r.index, (int32_t)prims.size()};
            prims.push_back(p);
        } else for (int t = triOffset[r.index], tEnd = triOffset[r.index] + sd.meshes[r.index].ntris; t < tEnd; ++t) {
            const V3 &p1 = mesh.p[mesh.vertexIndex[3 * t]], &p2 = mesh.p[mesh.vertexIndex[3 * t + 1]], &p3 = mesh.p[mesh.vertexIndex[3 * t + 2]];
            bounds.push_back(Union(BBox(p1, p2), p3));   // trianglemesh.inl:45-51
            skc_prim p = {0, 0, t, (int32_t)prims.size()};
            prims.push_back(p);
        }
    }
    BVHAccel bvh(bounds, (uint32_t)sd.maxNodePrims, sd.splitMethod);
    out->nodes = bvh.nodes;
    out->prims.clear();
    for (uint32_t pn : bvh.orderedPrims) out->prims.push_back(prims[pn]);
    out->lights.clear();
    for (size_t i = 0; i < sd.lights.size(); ++i) {
        float Le[30];
        if (Lemit30PerLight) memcpy(Le, Lemit30PerLight + 30 * i, sizeof Le);
        else {
            RGBToSpectrum(sd.lights[i].Lrgb, Le);
        }
        out->lights.push_back(sd.lights[i].Pack(Le));
    }
    out->mesh = &mesh;
    return true;
}

}  // namespace skin

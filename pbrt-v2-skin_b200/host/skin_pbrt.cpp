// skin_pbrt.cpp -- the pbrt scene-description front end for the subsurface path: `pbrt scene.pbrt` with the
// reference's own scene files.
//
//   tokens          core/pbrtlex.ll:112-200   (comments, quoted strings with escapes, NUMBER -> (float)atof, [ ], Include)
//   statements      core/pbrtparse.yy:352-640 (directive + "type name" value lists; a value is one item or a [ ] array)
//   parameter types core/pbrtparse.yy:805-840 (integer float bool point vector normal string texture color/rgb xyz
//                                              blackbody spectrum skinlayer layer)
//   state machine   core/api.cpp:700-1330     (CTM, attribute/transform stacks, named coordinate systems, graphics state,
//                                              render options, Shape -> primitive / area light, WorldEnd -> MakeScene)
//   plugin factories: CreateImageFilm film/image.cpp:219-252, CreatePerspectiveCamera cameras/perspective.cpp:117-156,
//     CreateLowDiscrepancySampler samplers/lowdiscrepancy.cpp:82-91, Create*Filter filters/*.cpp,
//     CreateMultipoleSubsurfaceIntegrator integrators/multipolesubsurface.cpp:307-320, CreateBVHAccelerator
//     accelerators/bvh.cpp:547-552, CreateSphereShape shapes/sphere.cpp:217-225, CreateTriangleMeshShape
//     shapes/trianglemesh.cpp:354-420, CreateDiffuseAreaLight lights/diffuse.cpp:59-66, CreateLayeredSkinMaterial
//     materials/layeredskin.cpp:222-262 -- same parameter names and defaults.
//
// Everything the device path cannot render is refused with an error that names the directive (the reference's
// Severe()), never silently replaced: other integrators / cameras / samplers / lights / materials on visible geometry,
// textures bound to the skin, object instancing, motion blur, participating media.
#include "skin_scene.h"
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace skin {
namespace {

struct Param {
    std::string type, name;
    std::vector<float> nums;
    std::vector<std::string> strs;
    mutable bool lookedUp = false;
};

struct ParamSet {
    std::vector<Param> ps;
    const Param *find(const char *name, const char *t1, const char *t2 = nullptr) const {
        for (const Param &p : ps)
            if (p.name == name && (p.type == t1 || (t2 && p.type == t2))) { p.lookedUp = true; return &p; }
        return nullptr;
    }
    float FindOneFloat(const char *n, float d) const { const Param *p = find(n, "float"); return p && p->nums.size() == 1 ? p->nums[0] : d; }
    int FindOneInt(const char *n, int d) const { const Param *p = find(n, "integer"); return p && p->nums.size() == 1 ? (int)p->nums[0] : d; }
    bool FindOneBool(const char *n, bool d) const { const Param *p = find(n, "bool"); return p && p->strs.size() == 1 ? p->strs[0] == "true" : d; }
    std::string FindOneString(const char *n, const std::string &d) const { const Param *p = find(n, "string"); return p && p->strs.size() == 1 ? p->strs[0] : d; }
    std::string FindTexture(const char *n) const { const Param *p = find(n, "texture"); return p && p->strs.size() == 1 ? p->strs[0] : ""; }
    const std::vector<float> *FindFloats(const char *n, const char *t1, const char *t2 = nullptr) const { const Param *p = find(n, t1, t2); return p ? &p->nums : nullptr; }
    bool Has(const char *n) const { for (const Param &p : ps) if (p.name == n) return true; return false; }
};

struct Source {
    std::string text, file;
    size_t pos = 0;
    int line = 1;
};

enum TokKind { T_EOF, T_ID, T_NUM, T_STR, T_LB, T_RB };
struct Token { TokKind kind = T_EOF; std::string s; float num = 0.f; };

struct GraphicsState {       // core/api.cpp:166-196
    std::string material = "matte";
    ParamSet materialParams;
    std::string areaLight;
    ParamSet areaLightParams;
    std::string currentNamedMaterial;
    bool reverseOrientation = false;
};

class Front {
public:
    Front(SceneDesc *sd, const LoadOptions *opt) : sd(sd) { if (opt) options = *opt; }
    bool Run(const std::string &path);

private:
    SceneDesc *sd;
    LoadOptions options;
    std::vector<Source> stack;
    std::string searchDirectory;
    bool failed = false;
    // api state
    enum { STATE_OPTIONS, STATE_WORLD, STATE_DONE } state = STATE_OPTIONS;
    Transform ctm;
    std::map<std::string, Transform> namedCoordinateSystems;
    GraphicsState gs;
    std::vector<GraphicsState> pushedGraphicsStates;
    std::vector<Transform> pushedTransforms;
    std::map<std::string, std::string> textures;              // name -> class, declared but not evaluable here
    std::map<std::string, AlbedoTexture> floatPyramids;        // the same float image maps with their MIPMap (EWA, maxanisotropy 8): usable as the skin's "texture bumpmap"
    std::map<std::string, AlbedoTexture> colorImageMaps;       // Texture ".." "color" | "spectrum" "imagemap": usable as the skin's "texture albedo"
    std::map<std::string, std::string> floatImageMapErrors;
    std::map<std::string, AlphaTexture> floatImageMaps;        // Texture ".." "float" "imagemap": usable as a mesh's "texture alpha"
    std::map<std::string, GraphicsState> namedMaterials;      // MakeNamedMaterial: type in .material, params in .materialParams
    uint32_t nextMaterialId = 0;                              // core/material.h:63 (one Material object per Shape / MakeNamedMaterial)
    std::map<std::string, uint32_t> namedMaterialIds;
    bool haveSkin = false;
    SkinParams skinOfScene;
    bool haveToken = false;
    Token pushed;
    bool haveFilm = false, haveCamera = false, haveSampler = false, haveIntegrator = false;
    ParamSet cameraParams;
    std::vector<char> attrKinds;                              // 'a' AttributeBegin, 't' TransformBegin (api.cpp:877-925)

    void Error(const char *fmt, ...);       // fatal for this front end (the reference's Severe, or an Error that would change the image)
    void Warn(const char *fmt, ...);
    bool next(Token *t);
    bool lex(Token *t);
    void unget(const Token &t) { pushed = t; haveToken = true; }
    bool number(float *v, const char *what);
    bool string(std::string *s, const char *what);
    bool paramlist(ParamSet *ps);
    bool numArray(std::vector<float> *v, const char *what);
    void reportUnused(const ParamSet &ps, const char *what);
    bool include(const std::string &name);
    std::string resolve(const std::string &name) const;       // ParamSet::FindOneFilename -> AbsolutePath(ResolveFilename()), core/fileutil.cpp:51-73
    bool statement(const std::string &id);
    void shape(const std::string &name, const ParamSet &ps);
    bool skinFromParams(const ParamSet &mp, const ParamSet &shapeParams, SkinParams *out);
    bool greyColor(const ParamSet &ps, const char *name, float *rgb, const char *what);
};

void Front::Error(const char *fmt, ...) {
    char buf[1024];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    if (!failed) {
        char where[512] = "";
        if (!stack.empty()) snprintf(where, sizeof where, "%s(%d): ", stack.back().file.c_str(), stack.back().line);
        sd->error = std::string(where) + buf;
    }
    failed = true;
}
void Front::Warn(const char *fmt, ...) {
    char buf[1024];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    char where[512] = "";
    if (!stack.empty()) snprintf(where, sizeof where, "%s(%d): ", stack.back().file.c_str(), stack.back().line);
    sd->warnings.push_back(std::string(where) + buf);
}

static bool isAbsolute(const std::string &f) { return !f.empty() && (f[0] == '/' || (f.size() > 2 && f[1] == ':')); }

static bool readFile(const std::string &path, std::string *out) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END); long n = ftell(f); fseek(f, 0, SEEK_SET);
    out->resize((size_t)std::max(0L, n));
    size_t got = n > 0 ? fread(&(*out)[0], 1, (size_t)n, f) : 0;
    fclose(f);
    return got == (size_t)std::max(0L, n);
}

// One token of the innermost open file; at its end the Include stack is popped (pbrtlex.ll include_pop / yywrap).
bool Front::lex(Token *t) {
    while (!stack.empty()) {
        Source &s = stack.back();
        const std::string &x = s.text;
        // whitespace and comments
        while (s.pos < x.size()) {
            char c = x[s.pos];
            if (c == '\n') { ++s.line; ++s.pos; }
            else if (c == ' ' || c == '\t' || c == '\r') ++s.pos;
            else if (c == '#') { while (s.pos < x.size() && x[s.pos] != '\n') ++s.pos; }
            else break;
        }
        if (s.pos >= x.size()) {
            if (stack.size() == 1) { t->kind = T_EOF; return true; }
            stack.pop_back();
            continue;
        }
        char c = x[s.pos];
        if (c == '[') { ++s.pos; t->kind = T_LB; return true; }
        if (c == ']') { ++s.pos; t->kind = T_RB; return true; }
        if (c == '"') {                                       // <STR> rules, pbrtlex.ll:177-198
            ++s.pos; t->kind = T_STR; t->s.clear();
            while (true) {
                if (s.pos >= x.size() || x[s.pos] == '\n') { Error("Unterminated string!"); return false; }
                char d = x[s.pos++];
                if (d == '"') return true;
                if (d != '\\') { t->s.push_back(d); continue; }
                if (s.pos >= x.size()) { Error("Unterminated string!"); return false; }
                char e = x[s.pos++];
                if (e == 'n') t->s.push_back('\n'); else if (e == 't') t->s.push_back('\t'); else if (e == 'r') t->s.push_back('\r');
                else if (e == 'b') t->s.push_back('\b'); else if (e == 'f') t->s.push_back('\f');
                else if (e == '\n') ++s.line;
                else if (e >= '0' && e <= '9' && s.pos + 1 < x.size() && isdigit((unsigned char)x[s.pos]) && isdigit((unsigned char)x[s.pos + 1])) {
                    int val = (e - '0') * 100 + (x[s.pos] - '0') * 10 + (x[s.pos + 1] - '0');
                    s.pos += 2;
                    while (val > 256) val -= 256;
                    t->s.push_back((char)val);
                } else t->s.push_back(e);
            }
        }
        if (isalpha((unsigned char)c) || c == '_') {
            size_t b = s.pos;
            while (s.pos < x.size() && (isalnum((unsigned char)x[s.pos]) || x[s.pos] == '_')) ++s.pos;
            t->kind = T_ID; t->s = x.substr(b, s.pos - b);
            return true;
        }
        // NUMBER [-+]?([0-9]+|(([0-9]+\.[0-9]*)|(\.[0-9]+)))([eE][-+]?[0-9]+)?
        size_t b = s.pos, p = s.pos;
        if (p < x.size() && (x[p] == '-' || x[p] == '+')) ++p;
        size_t d0 = p;
        while (p < x.size() && isdigit((unsigned char)x[p])) ++p;
        size_t nInt = p - d0, nFrac = 0;
        if (p < x.size() && x[p] == '.') {
            size_t q = p + 1;
            while (q < x.size() && isdigit((unsigned char)x[q])) ++q;
            nFrac = q - p - 1;
            if (nInt > 0 || nFrac > 0) p = q;
        }
        if (nInt == 0 && nFrac == 0) { Error("Illegal character: %c (0x%x)", c, (int)(unsigned char)c); return false; }
        if (p < x.size() && (x[p] == 'e' || x[p] == 'E')) {
            size_t q = p + 1;
            if (q < x.size() && (x[q] == '-' || x[q] == '+')) ++q;
            if (q < x.size() && isdigit((unsigned char)x[q])) { while (q < x.size() && isdigit((unsigned char)x[q])) ++q; p = q; }
        }
        s.pos = p;
        t->kind = T_NUM; t->s = x.substr(b, p - b);
        t->num = (float)atof(t->s.c_str());                   // pbrtlex.ll:165-168
        return true;
    }
    t->kind = T_EOF;
    return true;
}

bool Front::next(Token *t) {
    if (haveToken) { *t = pushed; haveToken = false; return true; }
    return lex(t);
}

bool Front::number(float *v, const char *what) {
    Token t;
    if (!next(&t)) return false;
    if (t.kind != T_NUM) { Error("%s: number expected", what); return false; }
    *v = t.num;
    return true;
}
bool Front::string(std::string *s, const char *what) {
    Token t;
    if (!next(&t)) return false;
    if (t.kind != T_STR) { Error("%s: quoted string expected", what); return false; }
    *s = t.s;
    return true;
}
bool Front::numArray(std::vector<float> *v, const char *what) {
    Token t;
    if (!next(&t)) return false;
    if (t.kind != T_LB) { Error("%s: '[' expected", what); return false; }
    while (true) {
        if (!next(&t)) return false;
        if (t.kind == T_RB) return true;
        if (t.kind != T_NUM) { Error("%s: number expected in array", what); return false; }
        v->push_back(t.num);
    }
}

// paramlist: ("type name" (item | '[' items ']'))*   -- pbrtparse.yy:303-339 and InitParamSet :661-803
bool Front::paramlist(ParamSet *ps) {
    static const char *kTypes[] = {"float", "integer", "bool", "point", "vector", "normal", "string", "texture", "color", "rgb", "xyz",
                                   "blackbody", "spectrum", "skinlayer", "layer"};
    while (true) {
        Token t;
        if (!next(&t)) return false;
        if (t.kind != T_STR) { unget(t); return true; }
        Param p;
        // "type name": lookupType, pbrtparse.yy:805-858
        size_t b = t.s.find_first_not_of(" \t");
        size_t e = b == std::string::npos ? b : t.s.find_first_of(" \t", b);
        if (b == std::string::npos || e == std::string::npos) { Error("Parameter \"%s\" doesn't have a type declaration?!", t.s.c_str()); return false; }
        p.type = t.s.substr(b, e - b);
        size_t nb = t.s.find_first_not_of(" \t", e);
        if (nb == std::string::npos) { Error("Unable to find parameter name from \"%s\"", t.s.c_str()); return false; }
        size_t ne = t.s.find_first_of(" \t", nb);
        p.name = t.s.substr(nb, ne == std::string::npos ? std::string::npos : ne - nb);
        bool known = false;
        for (const char *k : kTypes) known |= p.type == k;
        if (!known) { Error("Unable to decode type for name \"%s\"", t.s.c_str()); return false; }
        if (p.type == "rgb") p.type = "color";
        bool wantStrings = p.type == "string" || p.type == "texture" || p.type == "bool" || p.type == "layer";
        Token v;
        if (!next(&v)) return false;
        auto take = [&](const Token &it) -> bool {
            if (it.kind == T_STR) { p.strs.push_back(it.s); return true; }
            if (it.kind == T_NUM) { p.nums.push_back(it.num); return true; }
            Error("Parameter \"%s\": value expected", p.name.c_str());
            return false;
        };
        if (v.kind == T_LB) {
            while (true) {
                Token it;
                if (!next(&it)) return false;
                if (it.kind == T_RB) break;
                if (!take(it)) return false;
            }
        } else if (!take(v)) return false;
        if (!p.strs.empty() && !p.nums.empty()) { Error("Parameter \"%s\" mixes strings and numbers", p.name.c_str()); return false; }
        if (p.type == "spectrum" && !p.strs.empty()) { Error("Parameter \"%s\": spectrum files (SPD) are not supported by this front end", p.name.c_str()); return false; }
        if (p.type != "spectrum" && wantStrings != !p.strs.empty() && !(p.strs.empty() && p.nums.empty())) {
            Error("Parameter \"%s\" has wrong type of values for \"%s\"", p.name.c_str(), p.type.c_str());
            return false;
        }
        if (p.type == "bool")
            for (const std::string &s : p.strs)
                if (s != "true" && s != "false") Warn("Value \"%s\" unknown for boolean parameter \"%s\". Using \"false\".", s.c_str(), p.name.c_str());
        if (p.type == "integer")
            for (float &f : p.nums) f = (float)(int)f;        // InitParamSet truncates (pbrtparse.yy:686-694)
        ps->ps.push_back(std::move(p));
    }
}

void Front::reportUnused(const ParamSet &ps, const char *what) {   // ParamSet::ReportUnused, core/paramset.cpp
    for (const Param &p : ps.ps)
        if (!p.lookedUp) Warn("Parameter \"%s\" not used (%s)", p.name.c_str(), what);
}

std::string Front::resolve(const std::string &name) const {
    // ResolveFilename (core/fileutil.cpp:61-73): relative names live beside the scene file given on the command line
    std::string f = name;
    for (char &c : f) if (c == '\\') c = '/';
    if (!isAbsolute(f)) f = searchDirectory + f;
    return f;
}

bool Front::include(const std::string &name) {
    if (stack.size() > 32) { Error("Only 32 levels of nested Include allowed in scene files."); return false; }
    std::string f = resolve(name);
    Source s;
    s.file = f;
    if (!readFile(f, &s.text)) { Error("Unable to open included scene file \"%s\"", f.c_str()); return false; }
    stack.push_back(std::move(s));
    return true;
}

bool Front::greyColor(const ParamSet &ps, const char *name, float *rgb, const char *what) {
    const std::vector<float> *c = ps.FindFloats(name, "color", "spectrum");
    if (!c) return true;
    if (c->size() != 3) { Error("%s \"%s\": 3 RGB values expected (sampled spectra are not supported by this front end)", what, name); return false; }
    rgb[0] = (*c)[0]; rgb[1] = (*c)[1]; rgb[2] = (*c)[2];
    return true;
}

// CreateLayeredSkinMaterial (materials/layeredskin.cpp:222-262).  mp = Material directive parameters overlaid by the
// Shape's own (TextureParams looks in the shape's list first: core/paramset.cpp TextureParams::Find*).
bool Front::skinFromParams(const ParamSet &mp, const ParamSet &shapeParams, SkinParams *out) {
    auto F = [&](const char *n, float d) { float v = mp.FindOneFloat(n, d); return shapeParams.FindOneFloat(n, v); };
    const std::vector<float> *layers = shapeParams.FindFloats("layers", "skinlayer");
    if (!layers) layers = mp.FindFloats("layers", "skinlayer");
    if (!layers || layers->empty()) { Error("No layers param set for LayeredSkin material!"); return false; }   // :226-233 abort()
    if (layers->size() != 4) { Error("LayeredSkin: exactly two skin layers (epidermis, dermis) are supported on the device path; %d given", (int)layers->size() / 2); return false; }
    SkinParams k;
    k.layerThickness[0] = (*layers)[0]; k.layerIor[0] = (*layers)[1]; k.layerThickness[1] = (*layers)[2]; k.layerIor[1] = (*layers)[3];
    k.roughness = F("roughness", 0.4f);                      // materials/layeredskin.cpp:234
    k.nmperunit = F("nmperunit", 100e6f);
    k.f_mel = F("f_mel", 0.15f); k.f_eu = F("f_eu", 1.f); k.f_blood = F("f_blood", 0.002f); k.f_ohg = F("f_ohg", 0.3f);
    k.desiredlength = shapeParams.FindOneInt("desiredlength", mp.FindOneInt("desiredlength", 512));
    k.lerponthinslab = shapeParams.FindOneBool("lerponthinslab", mp.FindOneBool("lerponthinslab", true));
    k.Kr[0] = k.Kr[1] = k.Kr[2] = 1.f;                        // GetSpectrumTexture("Kr", Spectrum(1.f)) :235-236
    k.Kt[0] = k.Kt[1] = k.Kt[2] = 1.f;
    if (!greyColor(mp, "Kr", k.Kr, "layeredskin") || !greyColor(shapeParams, "Kr", k.Kr, "layeredskin")) return false;
    if (!greyColor(mp, "Kt", k.Kt, "layeredskin") || !greyColor(shapeParams, "Kt", k.Kt, "layeredskin")) return false;
    for (const char *tex : {"albedo", "bumpmap", "Kr", "Kt"}) {
        std::string t = shapeParams.FindTexture(tex);
        if (t.empty()) t = mp.FindTexture(tex);
        if (!t.empty() && !strcmp(tex, "albedo")) {      // GetSpectrumTexture("albedo", ..) :247 -- an image map is supported
            if (colorImageMaps.count(t)) { k.albedoTex = colorImageMaps[t]; continue; }
            Error("layeredskin \"texture albedo\" \"%s\": %s", t.c_str(), floatImageMapErrors.count(t) ? floatImageMapErrors[t].c_str()
                  : "only \"color\" / \"spectrum\" \"imagemap\" textures are supported on the device path");
            return false;
        }
        if (!t.empty() && !strcmp(tex, "bumpmap")) {     // GetFloatTextureOrNull("bumpmap"), materials/layeredskin.cpp -- an image map is supported
            if (floatPyramids.count(t)) { k.bumpTex = floatPyramids[t]; continue; }
            Error("layeredskin \"texture bumpmap\" \"%s\": %s", t.c_str(), floatImageMapErrors.count(t) ? floatImageMapErrors[t].c_str()
                  : "only \"float\" \"imagemap\" textures (EWA filter, maxanisotropy 8) are supported on the device path");
            return false;
        }
        if (!t.empty()) { Error("layeredskin \"texture %s\" \"%s\": this texture slot of the skin is not supported on the device path (\"texture albedo\" and \"texture bumpmap\" with image maps are; constant Kr / Kt)", tex, t.c_str()); return false; }
    }
    // Parameters the device path cannot honour are refused by name, not dropped: "doublerefsslf" swaps in FixedFresnelDielectric
    // (layeredskin.cpp:103-109), "genprofile false" / "usemontecarlo true" select other profile sources (:97-101).
    auto B = [&](const char *n, bool d) { return shapeParams.FindOneBool(n, mp.FindOneBool(n, d)); };
    if (mp.FindFloats("albedo", "color", "spectrum") || shapeParams.FindFloats("albedo", "color", "spectrum")) {      // GetSpectrumTexture("albedo", Spectrum(1.f)) :247
        if (!greyColor(mp, "albedo", k.albedo, "layeredskin") || !greyColor(shapeParams, "albedo", k.albedo, "layeredskin")) return false;
        k.hasAlbedo = true;
    }
    if (B("doublerefsslf", false)) { Error("layeredskin \"bool doublerefsslf\" \"true\" is not supported on the device path"); return false; }
    if (!B("genprofile", true)) { Error("layeredskin \"bool genprofile\" \"false\" is not supported on the device path (the multipole profile is always generated)"); return false; }
    if (B("usemontecarlo", false)) { Error("layeredskin \"bool usemontecarlo\" \"true\" is not supported on the device path (multipole profiles only)"); return false; }
    // read by the reference's factory but without effect on this path: Henyey-Greenstein g / b of the LayeredIntegrator's random
    // walk (layeredskin.cpp:53-60), photon count of the Monte-Carlo profile, the irradiance-point display of the material
    for (const char *n : {"ga_epi", "ga_derm", "b_derm", "irradiancepointsize"}) { mp.FindOneFloat(n, 0.f); shapeParams.FindOneFloat(n, 0.f); }
    mp.FindOneString("photons", ""); shapeParams.FindOneString("photons", "");
    if (B("showirradiancepoints", false)) Warn("layeredskin \"bool showirradiancepoints\" is ignored on the device path");
    if (mp.Has("rgbprofile") || shapeParams.Has("rgbprofile"))
        if (shapeParams.FindOneBool("rgbprofile", mp.FindOneBool("rgbprofile", false))) { Error("layeredskin \"bool rgbprofile\" \"true\" is not supported on the device path (30-bin spectral profiles only)"); return false; }
    *out = k;
    return true;
}

static bool sameSkin(const SkinParams &a, const SkinParams &b) {
    return a.roughness == b.roughness && a.nmperunit == b.nmperunit && a.f_mel == b.f_mel && a.f_eu == b.f_eu && a.f_blood == b.f_blood && a.f_ohg == b.f_ohg &&
           a.desiredlength == b.desiredlength && !memcmp(a.Kr, b.Kr, 12) && !memcmp(a.Kt, b.Kt, 12) && !memcmp(a.albedo, b.albedo, 12) && a.hasAlbedo == b.hasAlbedo &&
           a.albedoTex == b.albedoTex && a.bumpTex == b.bumpTex && a.lerponthinslab == b.lerponthinslab && !memcmp(a.layerThickness, b.layerThickness, 8) && !memcmp(a.layerIor, b.layerIor, 8);
}

// pbrtShape (core/api.cpp:982-1070)
void Front::shape(const std::string &name, const ParamSet &ps) {
    // one Material object per Shape (GraphicsState::CreateMaterial, api.cpp:1073-1088)
    uint32_t materialId;
    std::string matName = gs.material;
    const ParamSet *matParams = &gs.materialParams;
    if (!gs.currentNamedMaterial.empty() && namedMaterials.count(gs.currentNamedMaterial)) {
        materialId = namedMaterialIds[gs.currentNamedMaterial];
        matName = namedMaterials[gs.currentNamedMaterial].material;
        matParams = &namedMaterials[gs.currentNamedMaterial].materialParams;
    } else materialId = nextMaterialId++;

    if (name == "sphere") {
        if (gs.areaLight.empty()) { Error("Shape \"sphere\" without an AreaLightSource: only emitter spheres and layeredskin triangle meshes are supported on the device path"); return; }
        if (gs.areaLight != "area" && gs.areaLight != "diffuse") { Error("Area light \"%s\" unknown.", gs.areaLight.c_str()); return; }
        SphereLight l;
        l.ObjectToWorld = ctm;
        l.ReverseOrientation = gs.reverseOrientation;
        l.radius = ps.FindOneFloat("radius", 1.f);
        l.zmin = ps.FindOneFloat("zmin", -l.radius);
        l.zmax = ps.FindOneFloat("zmax", l.radius);
        l.phiMaxDeg = ps.FindOneFloat("phimax", 360.f);
        // CreateDiffuseAreaLight (lights/diffuse.cpp:59-66)
        float L[3] = {1, 1, 1}, sc[3] = {1, 1, 1};
        if (!greyColor(gs.areaLightParams, "L", L, "AreaLightSource") || !greyColor(gs.areaLightParams, "scale", sc, "AreaLightSource")) return;
        if (sc[0] != 1.f || sc[1] != 1.f || sc[2] != 1.f) { Error("AreaLightSource \"scale\" is not supported on the device path; fold it into \"L\""); return; }
        memcpy(l.Lrgb, L, sizeof L);
        l.nSamples = gs.areaLightParams.FindOneInt("nsamples", 1);
        if (options.quickRender) l.nSamples = std::max(1, l.nSamples / 4);
        if (matName != "matte") { Error("emitter sphere with Material \"%s\": emitters carry the default matte on the device path", matName.c_str()); return; }
        if (matParams->Has("Kd") || matParams->Has("sigma") || matParams->Has("bumpmap")) { Error("emitter sphere: only the default matte (Kd 0.5) is supported on the device path"); return; }
        reportUnused(ps, "Shape \"sphere\"");
        sd->primOrder.push_back(PrimRef{1, (int)sd->lights.size()});
        sd->lights.push_back(l);
        return;
    }
    if (name != "trianglemesh") { Error("Shape \"%s\" is not supported on the device path (trianglemesh with layeredskin, emitter spheres)", name.c_str()); return; }
    if (!gs.areaLight.empty()) { Error("AreaLightSource on a trianglemesh is not supported on the device path (sphere emitters only)"); return; }
    if (matName != "layeredskin") { Error("Shape \"trianglemesh\" with Material \"%s\": only layeredskin meshes are supported on the device path", matName.c_str()); return; }
    // CreateTriangleMeshShape (shapes/trianglemesh.cpp:354-420)
    const std::vector<float> *vi = ps.FindFloats("indices", "integer");
    const std::vector<float> *P = ps.FindFloats("P", "point");
    const std::vector<float> *uvs = ps.FindFloats("uv", "float");
    if (!uvs) uvs = ps.FindFloats("st", "float");
    if (ps.FindOneBool("discarddegenerateUVs", false)) { Error("trianglemesh \"discarddegenerateUVs\" is not supported by this front end"); return; }
    if (!vi || !P) { Error("trianglemesh needs \"integer indices\" and \"point P\""); return; }
    int npi = (int)P->size() / 3, nvi = (int)vi->size();
    if (uvs) {
        if ((int)uvs->size() < 2 * npi) { Warn("Not enough of \"uv\"s for triangle mesh.  Expencted %d, found %d.  Discarding.", 2 * npi, (int)uvs->size()); uvs = nullptr; }
        else if ((int)uvs->size() > 2 * npi) Warn("More \"uv\"s provided than will be used for triangle mesh.  (%d expcted, %d found)", 2 * npi, (int)uvs->size());
    }
    const std::vector<float> *S = ps.FindFloats("S", "vector");
    if (S && (int)S->size() / 3 != npi) { Warn("Number of \"S\"s for triangle mesh must match \"P\"s"); S = nullptr; }
    if (S) Warn("trianglemesh \"vector S\": per-vertex tangents are ignored on the device path (the LayeredSkin lobes are isotropic, so they only rotate the local shading frame about the normal)");
    const std::vector<float> *N = ps.FindFloats("N", "normal");
    if (N && (int)N->size() / 3 != npi) { Warn("Number of \"N\"s for triangle mesh must match \"P\"s"); N = nullptr; }
    std::vector<int> idx((size_t)nvi);
    for (int i = 0; i < nvi; ++i) {
        idx[i] = (int)(*vi)[i];
        if (idx[i] >= npi) { Error("trianglemesh has out of-bounds vertex index %d (%d \"P\" values were given", idx[i], npi); return; }
    }
    AlphaTexture alpha;                                       // shapes/trianglemesh.cpp:411-421
    const std::string alphaName = ps.FindTexture("alpha");
    if (!alphaName.empty()) {
        if (!floatImageMaps.count(alphaName)) {
            Error("trianglemesh \"texture alpha\" \"%s\": %s", alphaName.c_str(), floatImageMapErrors.count(alphaName) ? floatImageMapErrors[alphaName].c_str()
                  : textures.count(alphaName) ? "only float \"imagemap\" textures are supported as alpha on the device path" : "Couldn't find float texture");
            return;
        }
        alpha = floatImageMaps[alphaName];
    } else if (ps.FindOneFloat("alpha", 1.f) == 0.f) { Error("trianglemesh \"float alpha\" 0 (a fully transparent mesh) is not supported on the device path"); return; }
    SkinParams k;
    if (!skinFromParams(*matParams, ps, &k)) return;
    reportUnused(*matParams, "Material \"layeredskin\"");
    if (haveSkin && !sameSkin(k, skinOfScene)) { Error("more than one distinct layeredskin material in the scene: the device path holds one profile table"); return; }
    haveSkin = true; skinOfScene = k; sd->skin = k;
    TriangleMesh m;
    InitMesh(&m, ctm, gs.reverseOrientation, npi, P->data(), N ? N->data() : nullptr, uvs ? uvs->data() : nullptr, nvi / 3, idx.data());
    m.materialId = materialId;
    m.alpha = alpha;
    reportUnused(ps, "Shape \"trianglemesh\"");
    sd->primOrder.push_back(PrimRef{0, (int)sd->meshes.size()});
    sd->meshes.push_back(std::move(m));
}

bool Front::statement(const std::string &id) {
    auto needWorld = [&](const char *w) { if (state != STATE_WORLD) { Error("Scene description must be inside world block; \"%s\" not allowed.", w); return false; } return true; };
    auto needOptions = [&](const char *w) { if (state != STATE_OPTIONS) { Error("Options cannot be set inside world block; \"%s\" not allowed.", w); return false; } return true; };
    float v[16];
    std::string name;
    ParamSet ps;
    auto named = [&](const char *w) { return string(&name, w) && paramlist(&ps); };

    if (id == "Identity") { ctm = Transform(); return true; }
    if (id == "Translate") { for (int i = 0; i < 3; ++i) if (!number(&v[i], "Translate")) return false; ctm = ctm * Translate(V3(v[0], v[1], v[2])); return true; }
    if (id == "Scale") { for (int i = 0; i < 3; ++i) if (!number(&v[i], "Scale")) return false; ctm = ctm * Scale(v[0], v[1], v[2]); return true; }
    if (id == "Rotate") { for (int i = 0; i < 4; ++i) if (!number(&v[i], "Rotate")) return false; ctm = ctm * Rotate(v[0], V3(v[1], v[2], v[3])); return true; }
    if (id == "LookAt") {
        for (int i = 0; i < 9; ++i) if (!number(&v[i], "LookAt")) return false;
        memcpy(sd->lookat, v, 9 * sizeof(float));
        ctm = ctm * LookAt(V3(v[0], v[1], v[2]), V3(v[3], v[4], v[5]), V3(v[6], v[7], v[8]));     // api.cpp:759-767
        return true;
    }
    if (id == "Transform" || id == "ConcatTransform") {
        std::vector<float> a;
        if (!numArray(&a, id.c_str())) return false;
        if (a.size() != 16) { Error("Array argument to %s must have 16 elements.", id.c_str()); return false; }
        Transform t(Mat4(a[0], a[4], a[8], a[12], a[1], a[5], a[9], a[13], a[2], a[6], a[10], a[14], a[3], a[7], a[11], a[15]));   // api.cpp:727-744
        ctm = id == "Transform" ? t : ctm * t;
        return true;
    }
    if (id == "CoordinateSystem") { if (!string(&name, "CoordinateSystem")) return false; namedCoordinateSystems[name] = ctm; return true; }
    if (id == "CoordSysTransform") {
        if (!string(&name, "CoordSysTransform")) return false;
        if (namedCoordinateSystems.count(name)) ctm = namedCoordinateSystems[name];
        else Warn("Couldn't find named coordinate system \"%s\"", name.c_str());
        return true;
    }
    if (id == "ActiveTransform" || id == "TransformTimes") { Error("%s: animated transforms (motion blur) are not supported on the device path", id.c_str()); return false; }
    if (id == "Include") { return string(&name, "Include") && include(name); }

    if (id == "Film") {
        if (!named("Film") || !needOptions("Film")) return false;
        if (name != "image") { Error("Film \"%s\" unknown.", name.c_str()); return false; }
        // CreateImageFilm (film/image.cpp:219-252)
        std::string filename = ps.FindOneString("filename", "");
        if (!options.imageFile.empty()) {
            if (!filename.empty()) Warn("Output filename supplied on command line, \"%s\", ignored due to filename provided in scene description file, \"%s\".", options.imageFile.c_str(), filename.c_str());
            else filename = options.imageFile;
        }
        sd->filmFilename = filename;
        sd->xres = ps.FindOneInt("xresolution", 640);
        sd->yres = ps.FindOneInt("yresolution", 480);
        if (options.quickRender) { sd->xres = std::max(1, sd->xres / 4); sd->yres = std::max(1, sd->yres / 4); }
        ps.FindOneBool("display", false);
        if (const std::vector<float> *cr = ps.FindFloats("cropwindow", "float"))
            if (cr->size() == 4 && !((*cr)[0] == 0.f && (*cr)[1] == 1.f && (*cr)[2] == 0.f && (*cr)[3] == 1.f)) { Error("Film \"cropwindow\" is not supported on the device path (full frame only)"); return false; }
        haveFilm = true;
        reportUnused(ps, "Film");
        return true;
    }
    if (id == "Camera") {
        if (!named("Camera") || !needOptions("Camera")) return false;
        if (name != "perspective") { Error("Camera \"%s\": only \"perspective\" is supported on the device path", name.c_str()); return false; }
        sd->WorldToCamera = ctm;                              // CameraToWorld = Inverse(curTransform), api.cpp:862
        namedCoordinateSystems["camera"] = InverseT(ctm);
        cameraParams = ps;
        haveCamera = true;
        return true;
    }
    if (id == "Sampler") {
        if (!named("Sampler") || !needOptions("Sampler")) return false;
        if (name != "lowdiscrepancy") { Error("Sampler \"%s\": only \"lowdiscrepancy\" is supported on the device path", name.c_str()); return false; }
        sd->spp = ps.FindOneInt("pixelsamples", 4);
        if (options.quickRender) sd->spp = 1;
        haveSampler = true;
        reportUnused(ps, "Sampler");
        return true;
    }
    if (id == "PixelFilter") {
        if (!named("PixelFilter") || !needOptions("PixelFilter")) return false;
        PixelFilter f;
        f.name = name;
        if (name == "box") { f.xwidth = ps.FindOneFloat("xwidth", .5f); f.ywidth = ps.FindOneFloat("ywidth", .5f); }
        else if (name == "gaussian") { f.xwidth = ps.FindOneFloat("xwidth", 2.f); f.ywidth = ps.FindOneFloat("ywidth", 2.f); f.alpha = ps.FindOneFloat("alpha", 2.f); }
        else if (name == "mitchell") { f.xwidth = ps.FindOneFloat("xwidth", 2.f); f.ywidth = ps.FindOneFloat("ywidth", 2.f); f.B = ps.FindOneFloat("B", 1.f / 3.f); f.C = ps.FindOneFloat("C", 1.f / 3.f); }
        else if (name == "sinc") { f.xwidth = ps.FindOneFloat("xwidth", 4.f); f.ywidth = ps.FindOneFloat("ywidth", 4.f); f.tau = ps.FindOneFloat("tau", 3.f); }
        else if (name == "triangle") { f.xwidth = ps.FindOneFloat("xwidth", 2.f); f.ywidth = ps.FindOneFloat("ywidth", 2.f); }
        else { Error("Filter \"%s\" unknown.", name.c_str()); return false; }
        sd->filter = f;
        reportUnused(ps, "PixelFilter");
        return true;
    }
    if (id == "SurfaceIntegrator") {
        if (!named("SurfaceIntegrator") || !needOptions("SurfaceIntegrator")) return false;
        if (name != "multipolesubsurface") { Error("SurfaceIntegrator \"%s\": only \"multipolesubsurface\" runs on the device path", name.c_str()); return false; }
        // CreateMultipoleSubsurfaceIntegrator (integrators/multipolesubsurface.cpp:307-320)
        sd->maxdepth = ps.FindOneInt("maxdepth", 5);
        sd->maxerror = ps.FindOneFloat("maxerror", .05f);
        sd->minsampledistance = ps.FindOneFloat("minsampledistance", .25f);
        sd->mix = ps.FindOneFloat("mix", .5f);
        sd->incenter = ps.FindOneBool("incenter", false);
        sd->pointsFile = ps.FindOneString("pointsfile", "");
        if (!sd->pointsFile.empty()) sd->pointsFile = resolve(sd->pointsFile);
        if (ps.FindOneBool("showirradiancepoints", false)) { Error("multipolesubsurface \"showirradiancepoints\" is not supported on the device path"); return false; }
        sd->usePoissonPointFinder = ps.FindOneBool("usepoissonpointfinder", false);
        if (options.quickRender) sd->poissonMaxFails = std::max(10, 2000 / 10);
        if (options.quickRender) { /* the reference's integrator has no quick-render branch */ }
        haveIntegrator = true;
        reportUnused(ps, "SurfaceIntegrator");
        return true;
    }
    if (id == "VolumeIntegrator") {
        if (!named("VolumeIntegrator") || !needOptions("VolumeIntegrator")) return false;
        if (name != "emission") { Error("VolumeIntegrator \"%s\": only the default \"emission\" is supported (no participating media on the device path)", name.c_str()); return false; }
        return true;
    }
    if (id == "Renderer") {
        if (!named("Renderer") || !needOptions("Renderer")) return false;
        if (name == "tessellatesurfacepoints" || name == "surfacepoints") {      // CreateTessellateSurfacePointsRenderer / CreateSurfacePointsRenderer, renderers/surfacepoints.cpp:372-388
            sd->rendererName = name;
            sd->rendererMinDist = ps.FindOneFloat("minsampledistance", .25f);
            sd->rendererFilename = ps.FindOneString("filename", "");
            if (!sd->rendererFilename.empty()) sd->rendererFilename = resolve(sd->rendererFilename);
            sd->rendererVersionedFile = ps.FindOneBool("versioned", false);      // not in the reference: the tagged points file (skin_scene.h)
            if (options.quickRender) { sd->rendererMinDist *= 4.f; sd->poissonMaxFails = std::max(10, 2000 / 10); }
            haveIntegrator = true;                            // this renderer never calls the surface integrator
            reportUnused(ps, "Renderer");
            return true;
        }
        if (name != "sampler") { Error("Renderer \"%s\": only \"sampler\", \"tessellatesurfacepoints\" and \"surfacepoints\" are supported on the device path", name.c_str()); return false; }
        if (ps.FindOneBool("visualizeobjectids", false)) { Error("Renderer \"visualizeobjectids\" is not supported on the device path"); return false; }
        sd->rendererName = name;
        return true;
    }
    if (id == "Accelerator") {
        if (!named("Accelerator") || !needOptions("Accelerator")) return false;
        if (name == "bvh") {                                  // CreateBVHAccelerator (accelerators/bvh.cpp:547-552)
            sd->splitMethod = ps.FindOneString("splitmethod", "sah");
            sd->maxNodePrims = ps.FindOneInt("maxnodeprims", 4);
            if (sd->splitMethod != "sah" && sd->splitMethod != "middle" && sd->splitMethod != "equal") {
                Warn("BVH split method \"%s\" unknown.  Using \"sah\".", sd->splitMethod.c_str());
                sd->splitMethod = "sah";
            }
        } else Warn("Accelerator \"%s\": the device path always traverses a BVH (same intersections; ties between coincident hits may resolve differently)", name.c_str());
        return true;
    }
    if (id == "WorldBegin") {
        if (!needOptions("WorldBegin")) return false;
        state = STATE_WORLD;
        ctm = Transform();
        namedCoordinateSystems["world"] = ctm;
        return true;
    }
    if (id == "AttributeBegin") { if (!needWorld("AttributeBegin")) return false; pushedGraphicsStates.push_back(gs); pushedTransforms.push_back(ctm); attrKinds.push_back('a'); return true; }
    if (id == "AttributeEnd") {
        if (!needWorld("AttributeEnd")) return false;
        if (pushedGraphicsStates.empty() || attrKinds.empty() || attrKinds.back() != 'a') { Warn("Unmatched pbrtAttributeEnd() encountered. Ignoring it."); return true; }
        gs = pushedGraphicsStates.back(); pushedGraphicsStates.pop_back();
        ctm = pushedTransforms.back(); pushedTransforms.pop_back(); attrKinds.pop_back();
        return true;
    }
    if (id == "TransformBegin") { if (!needWorld("TransformBegin")) return false; pushedTransforms.push_back(ctm); attrKinds.push_back('t'); return true; }
    if (id == "TransformEnd") {
        if (!needWorld("TransformEnd")) return false;
        if (pushedTransforms.empty() || attrKinds.empty() || attrKinds.back() != 't') { Warn("Unmatched pbrtTransformEnd() encountered. Ignoring it."); return true; }
        ctm = pushedTransforms.back(); pushedTransforms.pop_back(); attrKinds.pop_back();
        return true;
    }
    if (id == "ReverseOrientation") { if (!needWorld("ReverseOrientation")) return false; gs.reverseOrientation = !gs.reverseOrientation; return true; }
    if (id == "AreaLightSource") {
        if (!named("AreaLightSource") || !needWorld("AreaLightSource")) return false;
        gs.areaLight = name; gs.areaLightParams = ps;
        return true;
    }
    if (id == "LightSource") {
        if (!named("LightSource") || !needWorld("LightSource")) return false;
        // MakeLight (core/api.cpp:470-500) -> CreatePointLight / CreateSpotLight / CreateDistantLight / CreateInfiniteLight
        if (name == "infinite") {      // lights/infinite.cpp:185-193
            float L[3] = {1, 1, 1}, scl[3] = {1, 1, 1};
            const bool haveScale = ps.FindFloats("scale", "color", "spectrum") != nullptr;
            if (!greyColor(ps, "L", L, "LightSource") || !greyColor(ps, "scale", scl, "LightSource")) return false;
            std::string map = ps.FindOneString("mapname", "");
            if (!map.empty()) map = resolve(map);
            if (sd->env.present) { Error("LightSource \"infinite\": one infinite light per scene is supported on the device path"); return false; }
            SphereLight l; std::string err;
            if (!MakeInfiniteLight(ctm, L, haveScale ? scl : nullptr, ps.FindOneInt("nsamples", 1), map, options.quickRender, &l, &sd->env, &err)) { Error("%s", err.c_str()); return false; }
            sd->lights.push_back(l);
            reportUnused(ps, "LightSource");
            return true;
        }
        if (name != "point" && name != "spot" && name != "distant") {
            Error("LightSource \"%s\": the device path has sphere AreaLightSource emitters and point / spot / distant / infinite lights", name.c_str());
            return false;
        }
        float rgb[3] = {1, 1, 1}, sc[3] = {1, 1, 1};
        if (!greyColor(ps, name == "distant" ? "L" : "I", rgb, "LightSource") || !greyColor(ps, "scale", sc, "LightSource")) return false;
        if (sc[0] != 1.f || sc[1] != 1.f || sc[2] != 1.f) { Error("LightSource \"scale\" is not supported on the device path; fold it into the intensity"); return false; }
        auto pt = [&](const char *n, V3 d) { const std::vector<float> *v = ps.FindFloats(n, "point"); return v && v->size() == 3 ? V3((*v)[0], (*v)[1], (*v)[2]) : d; };
        V3 from = pt("from", V3(0, 0, 0)), to = pt("to", V3(0, 0, 1));
        if (name == "point") sd->lights.push_back(MakePointLight(ctm, from, rgb));
        else if (name == "spot") sd->lights.push_back(MakeSpotLight(ctm, from, to, rgb, ps.FindOneFloat("coneangle", 30.f), ps.FindOneFloat("conedeltaangle", 5.f)));
        else sd->lights.push_back(MakeDistantLight(ctm, from, to, rgb));
        reportUnused(ps, "LightSource");
        return true;
    }
    if (id == "Material") {
        if (!named("Material") || !needWorld("Material")) return false;
        gs.material = name; gs.materialParams = ps; gs.currentNamedMaterial = "";
        return true;
    }
    if (id == "MakeNamedMaterial") {
        if (!named("MakeNamedMaterial") || !needWorld("MakeNamedMaterial")) return false;
        GraphicsState g;
        g.material = ps.FindOneString("type", "");
        if (g.material.empty()) { Error("No parameter string \"type\" found in named_material"); return false; }
        g.materialParams = ps;
        namedMaterials[name] = g;
        namedMaterialIds[name] = nextMaterialId++;
        return true;
    }
    if (id == "NamedMaterial") { if (!string(&name, "NamedMaterial") || !needWorld("NamedMaterial")) return false; gs.currentNamedMaterial = name; return true; }
    if (id == "Texture") {
        std::string type, texname;
        if (!string(&name, "Texture") || !string(&type, "Texture") || !string(&texname, "Texture") || !paramlist(&ps) || !needWorld("Texture")) return false;
        textures[name] = texname;                             // declared; refused where a skin material binds it
        if (type == "float" && texname == "imagemap") {       // CreateImageFloatTexture (textures/imagemap.cpp:103-140): kept for "texture alpha"
            const std::string mapping = ps.FindOneString("mapping", "uv"), wrap = ps.FindOneString("wrap", "repeat");
            AlphaTexture a; std::string err;
            if (mapping == "uv" && wrap == "repeat" &&
                LoadAlphaTexture(resolve(ps.FindOneString("filename", "")), ps.FindOneFloat("shift", 0.f), ps.FindOneFloat("scale", 1.f), ps.FindOneFloat("gamma", 1.f), &a, &err)) {
                a.su = ps.FindOneFloat("uscale", 1.f); a.sv = ps.FindOneFloat("vscale", 1.f); a.du = ps.FindOneFloat("udelta", 0.f); a.dv = ps.FindOneFloat("vdelta", 0.f);
                ps.FindOneFloat("maxanisotropy", 8.f); ps.FindOneBool("trilinear", false);      // without effect on a lookup with zero differentials
                floatImageMaps[name] = a;
                AlbedoTexture b;
                if (!ps.FindOneBool("trilinear", false) && ps.FindOneFloat("maxanisotropy", 8.f) == 8.f &&
                    LoadBumpTexture(resolve(ps.FindOneString("filename", "")), ps.FindOneFloat("shift", 0.f), ps.FindOneFloat("scale", 1.f), ps.FindOneFloat("gamma", 1.f), &b, &err)) {
                    b.su = a.su; b.sv = a.sv; b.du = a.du; b.dv = a.dv;
                    floatPyramids[name] = b;
                }
            } else floatImageMapErrors[name] = err.empty() ? "only \"uv\" mapping with \"repeat\" wrap is supported" : err;
        }
        if ((type == "color" || type == "spectrum") && texname == "imagemap") {      // CreateImageSpectrumTexture (textures/imagemap.cpp): kept for "texture albedo"
            const std::string mapping = ps.FindOneString("mapping", "uv"), wrap = ps.FindOneString("wrap", "repeat");
            AlbedoTexture a; std::string err;
            if (mapping == "uv" && wrap == "repeat" && !ps.FindOneBool("trilinear", false) && ps.FindOneFloat("maxanisotropy", 8.f) == 8.f &&
                LoadAlbedoTexture(resolve(ps.FindOneString("filename", "")), ps.FindOneFloat("shift", 0.f), ps.FindOneFloat("scale", 1.f), ps.FindOneFloat("gamma", 1.f), &a, &err)) {
                a.su = ps.FindOneFloat("uscale", 1.f); a.sv = ps.FindOneFloat("vscale", 1.f); a.du = ps.FindOneFloat("udelta", 0.f); a.dv = ps.FindOneFloat("vdelta", 0.f);
                colorImageMaps[name] = a;
            } else floatImageMapErrors[name] = err.empty() ? "only \"uv\" mapping, \"repeat\" wrap, the EWA filter and maxanisotropy 8 are supported" : err;
        }
        return true;
    }
    if (id == "Shape") {
        if (!named("Shape") || !needWorld("Shape")) return false;
        shape(name, ps);
        return !failed;
    }
    if (id == "Volume") { Error("Volume: participating media are not supported on the device path"); return false; }
    if (id == "ObjectBegin" || id == "ObjectEnd" || id == "ObjectInstance") { Error("%s: object instancing is not supported on the device path", id.c_str()); return false; }
    if (id == "WorldEnd") {
        if (!needWorld("WorldEnd")) return false;
        while (!pushedGraphicsStates.empty()) { Warn("Missing end to pbrtAttributeBegin()"); pushedGraphicsStates.pop_back(); }
        state = STATE_DONE;
        return true;
    }
    Error("syntax error: unknown directive \"%s\"", id.c_str());
    return false;
}

bool Front::Run(const std::string &path) {
    size_t slash = path.find_last_of('/');
    searchDirectory = slash == std::string::npos ? "" : path.substr(0, slash + 1);
    Source s;
    s.file = path;
    if (!readFile(path, &s.text)) { sd->error = "Unable to open scene file \"" + path + "\""; return false; }
    stack.push_back(std::move(s));
    // api.cpp RenderOptions defaults (:205-230): Film "image" 640x480, Camera "perspective" fov 90, Sampler
    // "lowdiscrepancy" 4, PixelFilter "box"; SurfaceIntegrator defaults to "directlighting", which this path refuses
    sd->xres = 640; sd->yres = 480; sd->spp = 4; sd->fov = 90.f;
    sd->lights.clear(); sd->meshes.clear(); sd->primOrder.clear(); sd->warnings.clear();
    sd->WorldToCamera = Transform();
    while (!failed && state != STATE_DONE) {
        Token t;
        if (!next(&t)) break;
        if (t.kind == T_EOF) break;
        if (t.kind != T_ID) { Error("syntax error: directive expected"); break; }
        if (!statement(t.s)) break;
    }
    if (failed) return false;
    if (state != STATE_DONE) { Error("scene file ended without WorldEnd"); return false; }
    if (!haveIntegrator) { sd->error = "no SurfaceIntegrator \"multipolesubsurface\" in the scene: the default \"directlighting\" does not run on the device path"; return false; }
    // CreatePerspectiveCamera (cameras/perspective.cpp:117-156), after the film is known
    {
        const ParamSet &cp = cameraParams;
        float so = cp.FindOneFloat("shutteropen", 0.f), sc = cp.FindOneFloat("shutterclose", 1.f);
        (void)so; (void)sc;                                   // no moving geometry: the time sample is drawn but never used
        if (cp.FindOneFloat("lensradius", 0.f) > 0.f) { sd->error = "Camera \"lensradius\" > 0 (depth of field) is not supported on the device path"; return false; }
        cp.FindOneFloat("focaldistance", 1e30f);
        sd->frameAspect = cp.FindOneFloat("frameaspectratio", 0.f);
        if (const std::vector<float> *sw = cp.FindFloats("screenwindow", "float"))
            if (sw->size() == 4) { sd->haveScreenWindow = true; memcpy(sd->screenWindow, sw->data(), 16); }
        sd->fov = cp.FindOneFloat("fov", 90.f);
        float halffov = cp.FindOneFloat("halffov", -1.f);
        if (halffov > 0.f) sd->fov = 2.f * halffov;
        reportUnused(cp, "Camera");
    }
    if (sd->meshes.empty()) { sd->error = "no layeredskin trianglemesh in the scene"; return false; }
    return true;
}

}  // namespace

bool SceneDesc::LoadPbrt(const std::string &path, const LoadOptions *opt) {
    Front f(this, opt);
    return f.Run(path);
}

}  // namespace skin

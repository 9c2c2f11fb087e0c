// skin_envmap.cpp -- LightSource "infinite" on the host side: what InfiniteAreaLight's constructor prepares
// (lights/infinite.cpp:67-103), as flat arrays for the device.
//
//   texels      ReadImagePFM (core/imageio.cpp:315-395: rows in file order, no flip) times L.ToRGBSpectrum(); without a map one
//               texel = L.ToRGBSpectrum()
//   pyramid     MIPMap<RGBSpectrum> (core/mipmap.h:111-196): each level the .25f-weighted sum of four texels of the finer one;
//               images that are not a power of two in size are Lanczos-zoomed to the next one first (textures; refused for the
//               environment map, whose distribution keeps the file's size)
//   img         Lookup(u / w, v / h, 1 / max(w, h)).y() * sinf(M_PI * (v + .5f) / h): the trilinear lookup of mipmap.h:239-271
//   sampling    Distribution2D / Distribution1D (core/montecarlo.h:55-77, montecarlo.cpp:358-370): step-function cdfs,
//               fp32 running sums in index order
//
// The results must equal the reference's arrays bit for bit (tests/test_envlight.py compares them with its export), so the
// arithmetic is spelt in its operation order and uses the same libm (logf for the level, sinf for the row weight).
#include "skin_scene.h"
#include "skin_cie.h"
#include <cstdio>

namespace skin {

namespace {

inline int WrapIndex(int a, int b) { int n = a / b; a -= n * b; if (a < 0) a += b; return a; }      // Mod (core/pbrt.h)
inline int FloorToInt(float v) { return (int)floorf(v); }
const float kPiF = 3.14159265358979323846f;          // M_PI is a float literal in the reference (core/pbrt.h:193-196)

struct Rgb { float r, g, b; };
inline Rgb operator+(Rgb a, Rgb b) { return {a.r + b.r, a.g + b.g, a.b + b.b}; }
inline Rgb operator*(float s, Rgb a) { return {a.r * s, a.g * s, a.b * s}; }         // Spectrum * float, per component

struct Pyramid {
    struct Level { int w, h; std::vector<Rgb> px; };
    std::vector<Level> levels;
    const Rgb &At(int level, int s, int t) const {                                   // Texel, TEXTURE_REPEAT
        const Level &l = levels[level];
        return l.px[(size_t)WrapIndex(t, l.h) * l.w + WrapIndex(s, l.w)];
    }
    // Lanczos (core/texture.cpp:266-274) and MIPMap::resampleWeights (core/mipmap.h:67-87): four weights per texel of the finer grid
    static float LanczosWindow(float x, float tau = 2.f) {
        x = fabsf(x);
        if (x < 1e-5) return 1;
        if (x > 1.) return 0;
        x *= kPiF;
        const float s = sinf(x * tau) / (x * tau);
        const float lanczos = sinf(x) / x;
        return s * lanczos;
    }
    struct Tap { int first; float w[4]; };
    static std::vector<Tap> ResampleTaps(unsigned oldres, unsigned newres) {
        std::vector<Tap> taps(newres);
        const float filterwidth = 2.f;
        for (unsigned i = 0; i < newres; ++i) {
            const float center = (i + .5f) * oldres / newres;
            taps[i].first = FloorToInt((center - filterwidth) + 0.5f);
            for (int j = 0; j < 4; ++j) {
                const float pos = taps[i].first + j + .5f;
                taps[i].w[j] = LanczosWindow((pos - center) / filterwidth);
            }
            const float invSum = 1.f / (taps[i].w[0] + taps[i].w[1] + taps[i].w[2] + taps[i].w[3]);
            for (int j = 0; j < 4; ++j) taps[i].w[j] *= invSum;
        }
        return taps;
    }
    // An image whose size is not a power of two is first zoomed to the next one, rows then columns, with TEXTURE_REPEAT
    // (core/mipmap.h:119-170); only the second pass clamps at 0
    static void ZoomToPow2(int *w, int *h, std::vector<Rgb> *img) {
        auto pow2 = [](unsigned v) { unsigned p = 1; while (p < v) p <<= 1; return p; };
        const unsigned sres = (unsigned)*w, tres = (unsigned)*h, sP = pow2(sres), tP = pow2(tres);
        if (sP == sres && tP == tres) return;
        std::vector<Rgb> zoomed((size_t)sP * tP, Rgb{0.f, 0.f, 0.f});
        const std::vector<Tap> sTaps = ResampleTaps(sres, sP);
        for (unsigned t = 0; t < tres; ++t)
            for (unsigned s = 0; s < sP; ++s) {
                Rgb acc{0.f, 0.f, 0.f};
                for (int j = 0; j < 4; ++j) acc = acc + sTaps[s].w[j] * (*img)[(size_t)t * sres + WrapIndex(sTaps[s].first + j, (int)sres)];
                zoomed[(size_t)t * sP + s] = acc;
            }
        const std::vector<Tap> tTaps = ResampleTaps(tres, tP);
        std::vector<Rgb> column(tP);
        for (unsigned s = 0; s < sP; ++s) {
            for (unsigned t = 0; t < tP; ++t) {
                Rgb acc{0.f, 0.f, 0.f};
                for (int j = 0; j < 4; ++j) acc = acc + tTaps[t].w[j] * zoomed[(size_t)WrapIndex(tTaps[t].first + j, (int)tres) * sP + s];
                column[t] = acc;
            }
            for (unsigned t = 0; t < tP; ++t) zoomed[(size_t)t * sP + s] = Rgb{std::max(column[t].r, 0.f), std::max(column[t].g, 0.f), std::max(column[t].b, 0.f)};
        }
        img->swap(zoomed); *w = (int)sP; *h = (int)tP;
    }
    void Build(int w, int h, std::vector<Rgb> img) {
        ZoomToPow2(&w, &h, &img);
        const float invLog2 = 1.f / logf(2.f);
        const int n = 1 + FloorToInt(logf(float(std::max(w, h))) * invLog2);          // 1 + Log2Int(float(max(sres, tres)))
        levels.resize(n);
        levels[0] = {w, h, img};
        for (int i = 1; i < n; ++i) {
            const int sw = std::max(1, levels[i - 1].w / 2), sh = std::max(1, levels[i - 1].h / 2);
            levels[i].w = sw; levels[i].h = sh; levels[i].px.resize((size_t)sw * sh);
            for (int t = 0; t < sh; ++t)
                for (int s = 0; s < sw; ++s)
                    levels[i].px[(size_t)t * sw + s] = .25f * (At(i - 1, 2 * s, 2 * t) + At(i - 1, 2 * s + 1, 2 * t) + At(i - 1, 2 * s, 2 * t + 1) + At(i - 1, 2 * s + 1, 2 * t + 1));
        }
    }
    Rgb Bilinear(int level, float s, float t) const {                               // MIPMap::triangle
        level = std::min(std::max(level, 0), (int)levels.size() - 1);
        s = s * levels[level].w - 0.5f;
        t = t * levels[level].h - 0.5f;
        const int s0 = FloorToInt(s), t0 = FloorToInt(t);
        const float ds = s - s0, dt = t - t0;
        return (1.f - ds) * (1.f - dt) * At(level, s0, t0) + (1.f - ds) * dt * At(level, s0, t0 + 1) +
               ds * (1.f - dt) * At(level, s0 + 1, t0) + ds * dt * At(level, s0 + 1, t0 + 1);
    }
    Rgb Lookup(float s, float t, float width) const {                               // MIPMap::Lookup(s, t, width)
        const unsigned nLevels = (unsigned)levels.size();
        const float invLog2 = 1.f / logf(2.f);
        const float level = nLevels - 1 + logf(std::max(width, 1e-8f)) * invLog2;
        if (level < 0) return Bilinear(0, s, t);
        if (level >= nLevels - 1) return At((int)nLevels - 1, 0, 0);
        const unsigned iLevel = (unsigned)FloorToInt(level);
        const float delta = level - iLevel;
        return (1.f - delta) * Bilinear((int)iLevel, s, t) + delta * Bilinear((int)iLevel + 1, s, t);
    }
};

// Distribution1D's constructor: cdf[n + 1] and the integral of the step function
float StepCdf(const float *f, int n, float *cdf) {
    cdf[0] = 0.f;
    for (int i = 1; i < n + 1; ++i) cdf[i] = cdf[i - 1] + f[i - 1] / n;
    const float funcInt = cdf[n];
    if (funcInt == 0.f) for (int i = 1; i < n + 1; ++i) cdf[i] = float(i) / float(n);
    else for (int i = 1; i < n + 1; ++i) cdf[i] /= funcInt;
    return funcInt;
}

bool ReadPfm(const std::string &path, int *w, int *h, std::vector<Rgb> *out, std::string *err) {
    FILE *fp = fopen(path.c_str(), "rb");
    if (!fp) { *err = "Unable to read image \"" + path + "\""; return false; }
    auto word = [&](char *buf, int cap) {          // one whitespace-terminated word, as the reference's reader takes them
        int n = 0, c = fgetc(fp);
        while (c != EOF && c != ' ' && c != '\n' && c != '\t' && c != '\r' && n < cap - 1) { buf[n++] = (char)c; c = fgetc(fp); }
        buf[n] = 0;
        return n;
    };
    char buf[80];
    bool ok = word(buf, 80) > 0;
    const int nCh = !strcmp(buf, "PF") ? 3 : !strcmp(buf, "Pf") ? 1 : 0;
    float scale = 0.f;
    ok = ok && nCh && word(buf, 80) > 0; *w = atoi(buf);
    ok = ok && word(buf, 80) > 0; *h = atoi(buf);
    ok = ok && word(buf, 80) > 0 && sscanf(buf, "%f", &scale) == 1 && *w > 0 && *h > 0;
    std::vector<float> data;
    if (ok) { data.resize((size_t)nCh * *w * *h); ok = fread(data.data(), 4, data.size(), fp) == data.size(); }
    fclose(fp);
    if (!ok) { *err = "Unable to read image \"" + path + "\" (PFM expected)"; return false; }
    if (!(scale < 0.f))                            // big-endian file on a little-endian host
        for (float &v : data) { unsigned char b[4]; memcpy(b, &v, 4); std::swap(b[0], b[3]); std::swap(b[1], b[2]); memcpy(&v, b, 4); }
    if (fabsf(scale) != 1.f) for (float &v : data) v *= fabsf(scale);
    out->resize((size_t)*w * *h);
    for (size_t i = 0; i < out->size(); ++i) (*out)[i] = nCh == 1 ? Rgb{data[i], data[i], data[i]} : Rgb{data[3 * i], data[3 * i + 1], data[3 * i + 2]};
    return true;
}

// ReadImageTGA (core/imageio.cpp:214-256 over core/targa.c) for true-colour (24 / 32 bit, types 2 and 10) and grey (8 bit, types 3
// and 11) files, raw or run-length encoded.  The reference normalises the picture to left-to-right / top-to-bottom and then copies
// its rows from the LAST to the first: texel row 0 is the bottom of the picture.  Colour-mapped files are refused.
bool ReadTga(const std::string &path, int *w, int *h, std::vector<Rgb> *out, std::string *err) {
    FILE *fp = fopen(path.c_str(), "rb");
    if (!fp) { *err = "Unable to read from TGA file \"" + path + "\""; return false; }
    std::vector<unsigned char> file;
    { unsigned char buf[65536]; size_t n; while ((n = fread(buf, 1, sizeof buf, fp)) > 0) file.insert(file.end(), buf, buf + n); }
    fclose(fp);
    auto fail = [&](const char *why) { *err = "Unable to read from TGA file \"" + path + "\" (" + why + ")"; return false; };
    if (file.size() < 18) return fail("short header");
    const int idLen = file[0], cmapType = file[1], type = file[2], cmapLen = file[5] | (file[6] << 8), cmapDepth = file[7];
    const int width = file[12] | (file[13] << 8), height = file[14] | (file[15] << 8), depth = file[16], desc = file[17];
    const bool rle = type == 10 || type == 11, grey = type == 3 || type == 11;
    if (type == 1 || type == 9) return fail("colour-mapped files are not supported on the device path");
    if (!(type == 2 || type == 10 || grey)) return fail("unknown image type");
    if (width <= 0 || height <= 0) return fail("empty image");
    if (!(grey ? depth == 8 : (depth == 24 || depth == 32))) return fail("unsupported pixel depth");
    const size_t bpp = (size_t)depth / 8, nPix = (size_t)width * height;
    size_t at = 18 + (size_t)idLen + (cmapType ? (size_t)cmapLen * ((cmapDepth + 7) / 8) : 0);
    std::vector<unsigned char> px(nPix * bpp);
    if (!rle) {
        if (file.size() < at + px.size()) return fail("short image data");
        memcpy(px.data(), file.data() + at, px.size());
    } else {
        size_t done = 0;
        while (done < nPix) {
            if (at >= file.size()) return fail("short run-length data");
            const int head = file[at++], count = (head & 0x7f) + 1;
            if (done + count > nPix) return fail("run-length packet past the end of the image");
            if (head & 0x80) {
                if (at + bpp > file.size()) return fail("short run-length data");
                for (int k = 0; k < count; ++k) memcpy(&px[(done + k) * bpp], &file[at], bpp);
                at += bpp;
            } else {
                if (at + bpp * count > file.size()) return fail("short run-length data");
                memcpy(&px[done * bpp], &file[at], bpp * count);
                at += bpp * count;
            }
            done += count;
        }
    }
    const bool topToBottom = (desc & 0x20) != 0, rightToLeft = (desc & 0x10) != 0;
    out->resize(nPix);
    for (int row = 0; row < height; ++row) {                  // texel row `row` = picture row height-1-row counted from the top
        const int fromTop = height - 1 - row, fileRow = topToBottom ? fromTop : height - 1 - fromTop;
        for (int x = 0; x < width; ++x) {
            const unsigned char *src = &px[((size_t)fileRow * width + (rightToLeft ? width - 1 - x : x)) * bpp];
            (*out)[(size_t)row * width + x] = grey ? Rgb{src[0] / 255.f, src[0] / 255.f, src[0] / 255.f} : Rgb{src[2] / 255.f, src[1] / 255.f, src[0] / 255.f};
        }
    }
    *w = width; *h = height;
    return true;
}

// ReadImage (core/imageio.cpp:54-75): the file type is the name's last four characters
bool ReadImageFile(const std::string &path, int *w, int *h, std::vector<Rgb> *out, std::string *err) {
    const std::string suffix = path.size() >= 5 ? path.substr(path.size() - 4) : std::string();
    if (suffix == ".tga" || suffix == ".TGA") return ReadTga(path, w, h, out, err);
    if (suffix == ".pfm" || suffix == ".PFM") return ReadPfm(path, w, h, out, err);
    *err = "Can't determine image file type from suffix of filename \"" + path + "\" (.pfm and .tga are read; no OpenEXR on this path)";
    return false;
}

}  // namespace

bool LoadAlphaTexture(const std::string &file, float shift, float scale, float gamma, AlphaTexture *out, std::string *err) {
    int w = 0, h = 0;
    std::vector<Rgb> px;
    if (!ReadImageFile(file, &w, &h, &px, err)) return false;
    for (Rgb &c : px) {
        const float y = 0.212671f * c.r + 0.715160f * c.g + 0.072169f * c.b;                // RGBSpectrum::y()
        c.r = c.g = c.b = scale * (powf(y, gamma) + shift);                                 // convertIn (textures/imagemap.h:94-97)
    }
    Pyramid::ZoomToPow2(&w, &h, &px);                 // MIPMap<float>'s level 0 (three equal channels resample like one)
    out->present = true; out->w = w; out->h = h; out->texels.resize(px.size());
    for (size_t i = 0; i < px.size(); ++i) out->texels[i] = px[i].r;
    return true;
}

// ImageTexture<RGBSpectrum, Spectrum> (textures/imagemap.cpp:56-86, imagemap.h:90-93): texels scale * (Pow(rgb, gamma) + shift), the
// MIPMap pyramid, and the EWA weight table exp(-2 r2) - exp(-2) (core/mipmap.h:186-193)
bool LoadAlbedoTexture(const std::string &file, float shift, float scale, float gamma, AlbedoTexture *out, std::string *err) {
    int w = 0, h = 0;
    std::vector<Rgb> px;
    if (!ReadImageFile(file, &w, &h, &px, err)) return false;
    for (Rgb &c : px) { c.r = scale * (powf(c.r, gamma) + shift); c.g = scale * (powf(c.g, gamma) + shift); c.b = scale * (powf(c.b, gamma) + shift); }
    Pyramid mip;
    mip.Build(w, h, px);
    *out = AlbedoTexture();
    out->present = true;
    for (const Pyramid::Level &l : mip.levels) {
        out->w.push_back(l.w); out->h.push_back(l.h);
        for (const Rgb &c : l.px) { out->rgb.push_back(c.r); out->rgb.push_back(c.g); out->rgb.push_back(c.b); }
    }
    for (int i = 0; i < 128; ++i) {
        const float alpha = 2, r2 = float(i) / float(128 - 1);
        out->weightLut[i] = expf(-alpha * r2) - expf(-alpha);
    }
    return true;
}

// ImageTexture<float, float> (textures/imagemap.cpp:56-86, imagemap.h:94-97): texels scale * (powf(luminance, gamma) + shift) and their
// MIPMap<float> pyramid -- the box filter of three equal channels is the float box filter
bool LoadBumpTexture(const std::string &file, float shift, float scale, float gamma, AlbedoTexture *out, std::string *err) {
    AlphaTexture level0;
    if (!LoadAlphaTexture(file, shift, scale, gamma, &level0, err)) return false;
    std::vector<Rgb> px(level0.texels.size());
    for (size_t i = 0; i < px.size(); ++i) px[i] = Rgb{level0.texels[i], level0.texels[i], level0.texels[i]};
    Pyramid mip;
    mip.Build(level0.w, level0.h, px);
    *out = AlbedoTexture();
    out->present = true; out->channels = 1;
    for (const Pyramid::Level &l : mip.levels) {
        out->w.push_back(l.w); out->h.push_back(l.h);
        for (const Rgb &c : l.px) out->rgb.push_back(c.r);
    }
    for (int i = 0; i < 128; ++i) {
        const float alpha = 2, r2 = float(i) / float(128 - 1);
        out->weightLut[i] = expf(-alpha * r2) - expf(-alpha);
    }
    return true;
}

// SampledSpectrum::ToRGB (core/spectrum.h:374-398): ToXYZ's running sums over the 30 samples, then XYZToRGB
void SpectrumToRGB(const float c[30], float rgb[3]) {
    float xyz[3] = {0.f, 0.f, 0.f};
    for (int i = 0; i < 30; ++i) { xyz[0] += kCieXYZ30[0][i] * c[i]; xyz[1] += kCieXYZ30[1][i] * c[i]; xyz[2] += kCieXYZ30[2][i] * c[i]; }
    const float scale = float(700 - 400) / float(kCieYIntegral * 30);
    xyz[0] *= scale; xyz[1] *= scale; xyz[2] *= scale;
    rgb[0] = 3.240479f * xyz[0] - 1.537150f * xyz[1] - 0.498535f * xyz[2];
    rgb[1] = -0.969256f * xyz[0] + 1.875991f * xyz[1] + 0.041556f * xyz[2];
    rgb[2] = 0.055648f * xyz[0] - 0.204043f * xyz[1] + 1.057311f * xyz[2];
}

// CreateInfiniteLight + the InfiniteAreaLight constructor (lights/infinite.cpp:67-103,185-193)
bool MakeInfiniteLight(const Transform &l2w, const float Lrgb[3], const float *scaleRgb, int nSamples, const std::string &mapname,
                       bool quickRender, SphereLight *light, EnvMap *env, std::string *err) {
    float L[30], sc[30], rgbL[3];
    RGBToSpectrum(Lrgb, L);                                                       // FindOneSpectrum("L"): FromRGB, reflectance
    if (scaleRgb) { RGBToSpectrum(scaleRgb, sc); for (int i = 0; i < 30; ++i) L[i] *= sc[i]; }      // L * sc; the default scale is Spectrum(1.0)
    SpectrumToRGB(L, rgbL);                                                       // L.ToRGBSpectrum()
    int w = 1, h = 1;
    std::vector<Rgb> texels;
    if (!mapname.empty()) {
        if (!ReadImageFile(mapname, &w, &h, &texels, err)) return false;
        if ((w & (w - 1)) || (h & (h - 1))) {
            *err = "LightSource \"infinite\": environment map \"" + mapname + "\" is not a power of two in both directions (its sampling distribution keeps the file's size while the MIPMap is resampled: the device holds one size)";
            return false;
        }
        for (Rgb &t : texels) { t.r *= rgbL[0]; t.g *= rgbL[1]; t.b *= rgbL[2]; }
    } else texels.assign(1, Rgb{rgbL[0], rgbL[1], rgbL[2]});
    Pyramid mip;
    mip.Build(w, h, texels);
    // scalar image for the sampling distribution
    const float filter = 1.f / std::max(w, h);
    const float pi = 3.14159265358979323846f;                                     // M_PI is a float literal in the reference (core/pbrt.h)
    std::vector<float> img((size_t)w * h);
    for (int v = 0; v < h; ++v) {
        const float vp = (float)v / (float)h;
        const float sinTheta = sinf(pi * float(v + .5f) / float(h));
        for (int u = 0; u < w; ++u) {
            const float up = (float)u / (float)w;
            const Rgb c = mip.Lookup(up, vp, filter);
            float y = 0.212671f * c.r + 0.715160f * c.g + 0.072169f * c.b;       // RGBSpectrum::y()
            y *= sinTheta;
            img[(size_t)u + (size_t)v * w] = y;
        }
    }
    env->w = w; env->h = h;
    env->rgb.resize((size_t)3 * w * h);
    for (size_t i = 0; i < texels.size(); ++i) { env->rgb[3 * i] = texels[i].r; env->rgb[3 * i + 1] = texels[i].g; env->rgb[3 * i + 2] = texels[i].b; }
    env->condFunc = img;
    env->condCdf.resize((size_t)h * (w + 1)); env->condInt.resize(h);
    for (int v = 0; v < h; ++v) env->condInt[v] = StepCdf(&img[(size_t)v * w], w, &env->condCdf[(size_t)v * (w + 1)]);
    env->margFunc = env->condInt;
    env->margCdf.resize(h + 1);
    env->margInt = StepCdf(env->margFunc.data(), h, env->margCdf.data());
    env->present = true;
    SphereLight l; l.kind = 4; l.radius = 0.f;
    l.nSamples = quickRender ? std::max(1, nSamples / 4) : nSamples;
    l.ObjectToWorld = l2w;
    l.envW = w; l.envH = h;
    memcpy(l.Lrgb, Lrgb, 12);
    *light = l;
    return true;
}

}  // namespace skin

/* mpc_oracle.c -- CPU restatement of the multipole diffusion-profile calculator.
 *
 * TEST INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke() and bench.py's checker legs may load this; the product
 * path (pbrt-v2-skin_b200/) never does.  Pinned against the real reference: tests/test_oracle.py compares it with
 * tests/golden/mpc_kat.npz, the output of the reference's own MPC_ComputeDiffusionProfile for its known-answer
 * configurations (src/multipole/test/test.cpp:84-116).
 *
 * Follows, function by function:
 *   DipoleCalculator ctor / Rd / Td        multipole/MultipoleProfileCalculator/DipoleCalculator.cpp:38-91
 *   ComputeLayerProfile                     MultipoleProfileCalculator.cpp:151-230
 *   ToFrequencyDomain / ToTimeDomain        :233-250 (ScaleAndShift / ScaleAndShiftReversed, numutil.h:192-217)
 *   CombineLayerProfiles                    :253-280 (complex Matrix ops, numutil.h:115-169)
 *   MPC_ComputeDiffusionProfile             :294-346 (unique radii through an ordered set, Kahan Sum numutil.h:228-239)
 *   resample / MPC_ResampleFor...           :355-426
 * The reference transforms with kissfft 1.3.0 (real 2-D, fp64); here a plain radix-2 complex FFT in fp64 is used --
 * the transform is the same linear map, its rounding noise (1e-16 of the grid peak) is not.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { float ior, thickness, mua, musp, g_HG; } orc_layer_spec;          /* MPC_LayerSpec */
typedef struct { float desiredStepSize; uint32_t desiredLength; int32_t lerpOnThinSlab; } orc_mpc_options;
typedef struct { double re, im; } cpx;

static uint32_t round_up_pow2(uint32_t v) { v--; v |= v >> 1; v |= v >> 2; v |= v >> 4; v |= v >> 8; v |= v >> 16; return v + 1; }

/* ---- DipoleCalculator ------------------------------------------------------------------- */
typedef struct { float d, sigma_tr, alphap, zpos, zneg; } dipole;
static const float PI_F = 3.141592654f;
#define INV_FOURPI (0.25f / PI_F)

static float fresnel_diffuse_reflectance(float eta) {
    if (eta >= 1.f) return -1.4399f / (eta * eta) + 0.7099f / eta + 0.6681f + 0.0636f * eta;
    float eta2 = eta * eta;
    return -0.4399f + 0.7099f / eta - 0.3319f / eta2 + 0.0636f / (eta2 * eta);
}
static float A_of(float Fdr) { return (1.f + Fdr) / (1.f - Fdr); }

static dipole dipole_make(float eta_0, float eta_d, float d, float sigma_a, float sigmap_s, int zi, int lerpOnThinSlab) {
    dipole c;
    c.d = d;
    float sigmap_t = sigma_a + sigmap_s;
    c.sigma_tr = sqrtf(3 * sigma_a * sigmap_t);
    c.alphap = sigmap_s / sigmap_t;
    float A_0 = A_of(fresnel_diffuse_reflectance(eta_0)), A_d = A_of(fresnel_diffuse_reflectance(eta_d));
    float D = 1.f / (3.f * sigmap_t);
    float zb_0 = 2.f * A_0 * D, zb_d = 2.f * A_d * D;
    float l = 1.f / sigmap_t;
    if (lerpOnThinSlab && l > d * .5f) l = d * .5f;
    c.zpos = 2.f * (float)zi * (d + zb_0 + zb_d) + l;
    c.zneg = c.zpos - 2.f * (l + zb_0);
    return c;
}
static float dipole_Rd(const dipole *c, float dsq) {
    float dpos = sqrtf(dsq + c->zpos * c->zpos), dneg = sqrtf(dsq + c->zneg * c->zneg);
    float dpos3 = dpos * dpos * dpos, dneg3 = dneg * dneg * dneg;
    return c->alphap * INV_FOURPI * (c->zpos * (1 + c->sigma_tr * dpos) * expf(-c->sigma_tr * dpos) / dpos3 -
                                     c->zneg * (1 + c->sigma_tr * dneg) * expf(-c->sigma_tr * dneg) / dneg3);
}
static float dipole_Td(const dipole *c, float dsq) {
    float d = c->d;
    float dpos = sqrtf(dsq + (d - c->zpos) * (d - c->zpos)), dneg = sqrtf(dsq + (d - c->zneg) * (d - c->zneg));
    float dpos3 = dpos * dpos * dpos, dneg3 = dneg * dneg * dneg;
    return c->alphap * INV_FOURPI * ((d - c->zpos) * (1 + c->sigma_tr * dpos) * expf(-c->sigma_tr * dpos) / dpos3 -
                                     (d - c->zneg) * (1 + c->sigma_tr * dneg) * expf(-c->sigma_tr * dneg) / dneg3);
}

/* ---- ComputeLayerProfile: R and T grids, G x G doubles, row-major ------------------------- */
static void layer_profile(const orc_layer_spec *spec, float iorUpper, float iorLower, float stepSize, int lerpOnThinSlab,
                          uint32_t G, double *R, double *T) {
    memset(R, 0, (size_t)G * G * sizeof(double));
    memset(T, 0, (size_t)G * G * sizeof(double));
    float thickness = spec->thickness;
    double mfp2 = 2. / (spec->mua + spec->musp);
    double lerp = 1.;
    if (lerpOnThinSlab) {
        lerp = (thickness < mfp2) ? (1. - exp(-thickness * 2. / mfp2)) / (1. - exp(-2.)) : 1.;
        if (thickness < 0.01 * mfp2) thickness = (float)(0.01 * mfp2);
    }
    uint32_t center = (G - 1) / 2, extent = center;
    dipole dcs[11];
    for (int pair = -5; pair <= 5; ++pair) dcs[pair + 5] = dipole_make(iorUpper, iorLower, thickness, spec->mua, spec->musp, pair, lerpOnThinSlab);
    double normalizeFactor = stepSize * stepSize;
    for (uint32_t row = 0; row <= extent; ++row)
        for (uint32_t col = row; col <= extent; ++col) {
            double drow2 = row * row, dcol2 = col * col;
            double r2 = (drow2 + dcol2) * (stepSize * stepSize);
            for (int k = 0; k < 11; ++k) {
                double rd = dipole_Rd(&dcs[k], (float)r2) * normalizeFactor;
                double td = dipole_Td(&dcs[k], (float)r2) * normalizeFactor;
                R[(size_t)(center + row) * G + center + col] += rd;
                T[(size_t)(center + row) * G + center + col] += td;
            }
        }
    if (lerp < 1.) {
        for (uint32_t row = 0; row <= extent; ++row)
            for (uint32_t col = row; col <= extent; ++col) {
                R[(size_t)(center + row) * G + center + col] *= lerp;
                T[(size_t)(center + row) * G + center + col] *= lerp;
            }
        T[(size_t)center * G + center] += 1. - lerp;
    }
    for (uint32_t row = 1; row <= extent; ++row)
        for (uint32_t col = 0; col < row; ++col) {
            R[(size_t)(center + row) * G + center + col] = R[(size_t)(center + col) * G + center + row];
            T[(size_t)(center + row) * G + center + col] = T[(size_t)(center + col) * G + center + row];
        }
    for (uint32_t row = 0; row <= extent; ++row)
        for (uint32_t col = 0; col <= extent; ++col) {
            double rd = R[(size_t)(center + row) * G + center + col], td = T[(size_t)(center + row) * G + center + col];
            R[(size_t)(center - row) * G + center + col] = rd; R[(size_t)(center + row) * G + center - col] = rd; R[(size_t)(center - row) * G + center - col] = rd;
            T[(size_t)(center - row) * G + center + col] = td; T[(size_t)(center + row) * G + center - col] = td; T[(size_t)(center - row) * G + center - col] = td;
        }
}

/* ---- 2-D complex FFT, N x N, N a power of two ------------------------------------------------ */
static void fft1d(cpx *a, uint32_t n, int stride, int inverse, const cpx *tw) {
    uint32_t logn = 0; while ((1u << logn) < n) ++logn;
    for (uint32_t i = 0; i < n; ++i) {
        uint32_t j = 0; for (uint32_t b = 0; b < logn; ++b) if (i & (1u << b)) j |= 1u << (logn - 1 - b);
        if (j > i) { cpx t = a[(size_t)i * stride]; a[(size_t)i * stride] = a[(size_t)j * stride]; a[(size_t)j * stride] = t; }
    }
    for (uint32_t len = 2; len <= n; len <<= 1) {
        uint32_t half = len >> 1, step = n / len;
        for (uint32_t s = 0; s < n; s += len)
            for (uint32_t k = 0; k < half; ++k) {
                cpx w = tw[k * step]; if (inverse) w.im = -w.im;
                cpx *x = &a[(size_t)(s + k) * stride], *y = &a[(size_t)(s + k + half) * stride];
                cpx t = {w.re * y->re - w.im * y->im, w.re * y->im + w.im * y->re};
                y->re = x->re - t.re; y->im = x->im - t.im;
                x->re += t.re; x->im += t.im;
            }
    }
}
static void fft2d(cpx *a, uint32_t n, int inverse) {
    cpx *tw = (cpx *)malloc(sizeof(cpx) * (n / 2 ? n / 2 : 1));
    for (uint32_t k = 0; k < n / 2; ++k) { long double ang = -2.0L * 3.14159265358979323846264338327950288L * k / n; tw[k].re = (double)cosl(ang); tw[k].im = (double)sinl(ang); }
    for (uint32_t r = 0; r < n; ++r) fft1d(a + (size_t)r * n, n, 1, inverse, tw);
    for (uint32_t c = 0; c < n; ++c) fft1d(a + c, n, (int)n, inverse, tw);
    free(tw);
}
/* ToFrequencyDomain: zero-pad to 2G and shift the centre to the origin (ScaleAndShift) */
static cpx *to_frequency(const double *g, uint32_t G) {
    uint32_t N = 2 * G, c = (G - 1) / 2;
    cpx *f = (cpx *)calloc((size_t)N * N, sizeof(cpx));
    for (uint32_t i = 0; i < G; ++i) {
        uint32_t ii = N - c + i; if (ii >= N) ii -= N;
        for (uint32_t j = 0; j < G; ++j) { uint32_t jj = N - c + j; if (jj >= N) jj -= N; f[(size_t)ii * N + jj].re = g[(size_t)i * G + j]; }
    }
    fft2d(f, N, 0);
    return f;
}
/* ToTimeDomain: inverse, scale by 1/N^2, ScaleAndShiftReversed back into a G x G window; the reference's inverse is a
 * complex-to-real transform, i.e. it keeps the real part */
static void to_time(cpx *f, uint32_t G, double *g) {
    uint32_t N = 2 * G, c = (G - 1) / 2;
    fft2d(f, N, 1);
    double scale = (double)1 / ((double)N * N);
    for (uint32_t i = 0; i < G; ++i) {
        uint32_t ii = N - c + i; if (ii >= N) ii -= N;
        for (uint32_t j = 0; j < G; ++j) { uint32_t jj = N - c + j; if (jj >= N) jj -= N; g[(size_t)i * G + j] = f[(size_t)ii * N + jj].re * scale; }
    }
}
static cpx cmul(cpx a, cpx b) { cpx r = {a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; return r; }
static cpx cdiv(cpx a, cpx b) { double d = b.re * b.re + b.im * b.im; cpx r = {(a.re * b.re + a.im * b.im) / d, (a.im * b.re - a.re * b.im) / d}; return r; }

/* CombineLayerProfiles: T12 = T1*T2/(1-R2*R1), R12 = R1 + T1*R2*T1/(1-R2*R1) in the frequency domain */
static void combine(const double *R1, const double *T1, const double *R2, const double *T2, uint32_t G, double *R12, double *T12) {
    uint32_t N = 2 * G; size_t nn = (size_t)N * N;
    cpx *fR1 = to_frequency(R1, G), *fR2 = to_frequency(R2, G), *fT1 = to_frequency(T1, G), *fT2 = to_frequency(T2, G);
    cpx *fR = (cpx *)malloc(nn * sizeof(cpx)), *fT = (cpx *)malloc(nn * sizeof(cpx));
    for (size_t i = 0; i < nn; ++i) {
        cpx one = cmul(fR2[i], fR1[i]); one.re = 1. - one.re; one.im = -one.im;
        cpx r = cdiv(cmul(cmul(fT1[i], fR2[i]), fT1[i]), one); r.re += fR1[i].re; r.im += fR1[i].im;
        fR[i] = r;
        fT[i] = cdiv(cmul(fT1[i], fT2[i]), one);
    }
    to_time(fR, G, R12); to_time(fT, G, T12);
    free(fR1); free(fR2); free(fT1); free(fT2); free(fR); free(fT);
}

static double kahan_sum(const double *v, size_t n) {     /* numutil.h:228-239 */
    double sum = 0., c = 0.;
    for (size_t i = 0; i < n; ++i) { double y = v[i] - c; double t = sum + y; c = (t - sum) - y; sum = t; }
    return sum;
}

typedef struct { float d2, r, t; uint32_t order; } entry;
static int entry_cmp(const void *a, const void *b) {
    const entry *x = (const entry *)a, *y = (const entry *)b;
    if (x->d2 < y->d2) return -1;
    if (x->d2 > y->d2) return 1;
    return x->order < y->order ? -1 : (x->order > y->order);
}

static void resample_one(const float *pDsq, const float *pRef, const float *pTrans, uint32_t length, float dsq, float *r, float *t) {
    float extent = pDsq[length - 1];
    if (dsq > extent) { *r = *t = 0.f; return; }
    uint32_t lo = 0, hi = length - 1;
    while (lo < hi) { uint32_t mid = (lo + hi) / 2; if (dsq > pDsq[mid]) lo = mid + 1; else hi = mid; }   /* first entry >= dsq */
    if (lo) {
        float la = (dsq - pDsq[lo - 1]) / (pDsq[lo] - pDsq[lo - 1]);
        if (la < 0.f) la = 0.f; else if (la > 1.f) la = 1.f;
        if (la != la) la = 0.5f;
        *r = (1.f - la) * pRef[lo - 1] + la * pRef[lo];
        *t = (1.f - la) * pTrans[lo - 1] + la * pTrans[lo];
    } else { *r = pRef[0]; *t = pTrans[0]; }
}

/* MPC_ComputeDiffusionProfile (+ MPC_ResampleForUniformDistanceSquaredDistribution when resample != 0).
 * Returns the table length; with d2 == NULL only the length is computed.  totals = {totalReflectance, totalTransmittance}. */
uint32_t orc_mpc_profile(uint32_t numLayers, const orc_layer_spec *specs, const orc_mpc_options *opt, int resample,
                         float *d2, float *R, float *T, float *totals) {
    uint32_t length = round_up_pow2(opt->desiredLength), G = 2 * length;
    uint32_t center = length - 1, extent = center;
    uint32_t cap = 0;
    for (uint32_t i = 0; i <= extent; ++i) for (uint32_t j = i; i * i + j * j <= extent * extent; ++j) ++cap;
    entry *e = (entry *)malloc(sizeof(entry) * cap);
    float stepSize = opt->desiredStepSize;
    size_t gg = (size_t)G * G;
    double *R0 = NULL, *T0 = NULL;
    if (d2) {
        R0 = (double *)malloc(gg * sizeof(double)); T0 = (double *)malloc(gg * sizeof(double));
        float iorLower = numLayers > 1 ? specs[0].ior / specs[1].ior : specs[0].ior;
        layer_profile(&specs[0], specs[0].ior, iorLower, stepSize, opt->lerpOnThinSlab, G, R0, T0);
        for (uint32_t i = 1; i < numLayers; ++i) {
            iorLower = numLayers > i + 1 ? specs[i].ior / specs[i + 1].ior : specs[i].ior;
            double *R1 = (double *)malloc(gg * sizeof(double)), *T1 = (double *)malloc(gg * sizeof(double));
            double *Rc = (double *)malloc(gg * sizeof(double)), *Tc = (double *)malloc(gg * sizeof(double));
            layer_profile(&specs[i], specs[i].ior / specs[i - 1].ior, iorLower, stepSize, opt->lerpOnThinSlab, G, R1, T1);
            combine(R0, T0, R1, T1, G, Rc, Tc);
            free(R0); free(T0); free(R1); free(T1);
            R0 = Rc; T0 = Tc;
        }
    }
    /* std::set<OutEntry> ordered by distanceSquared, first inserted wins */
    float denorm = 1.f / (stepSize * stepSize);
    uint32_t n = 0;
    for (uint32_t i = 0; i <= extent; ++i)
        for (uint32_t j = i; i * i + j * j <= extent * extent; ++j) {
            e[n].d2 = (float)(i * i + j * j) * stepSize * stepSize;
            e[n].r = R0 ? (float)R0[(size_t)(center + i) * G + center + j] * denorm : 0.f;
            e[n].t = T0 ? (float)T0[(size_t)(center + i) * G + center + j] * denorm : 0.f;
            e[n].order = n; ++n;
        }
    qsort(e, n, sizeof(entry), entry_cmp);
    uint32_t m = 0;
    for (uint32_t k = 0; k < n; ++k) if (m == 0 || e[k].d2 != e[m - 1].d2) e[m++] = e[k];
    uint32_t outLen = resample ? 2 * m : m;
    if (d2) {
        if (totals) { totals[0] = (float)kahan_sum(R0, gg); totals[1] = (float)kahan_sum(T0, gg); }
        if (!resample) for (uint32_t k = 0; k < m; ++k) { d2[k] = e[k].d2; R[k] = e[k].r; T[k] = e[k].t; }
        else {
            float *pd = (float *)malloc(3 * sizeof(float) * m), *pr = pd + m, *pt = pd + 2 * m;
            for (uint32_t k = 0; k < m; ++k) { pd[k] = e[k].d2; pr[k] = e[k].r; pt[k] = e[k].t; }
            float ext = pd[m - 1];
            for (uint32_t i = 0; i < outLen; ++i) {
                float dsq = (float)i * ext / (float)(outLen - 1);
                d2[i] = dsq;
                resample_one(pd, pr, pt, m, dsq, &R[i], &T[i]);
            }
            free(pd);
        }
        free(R0); free(T0);
    }
    free(e);
    return outLen;
}

// skc_common.cuh -- context object and arithmetic helpers shared by the kernels.
//
// Arithmetic convention (matters for parity, SURVEY.md section 7 "hard parts" #1):
// the reference is plain x86-64 SSE2 code (no FMA contraction, IEEE div/sqrt, no FTZ),
// so the whole library is compiled with -fmad=false -prec-div=true -prec-sqrt=true and
// every helper below spells the reference's operation order.  Where the reference
// promotes to double (Cross: core/geometry.h:529-536) so do we; M_PI is a float literal there (core/pbrt.h:196).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>
#include <map>
#include "../../include/skin_cuda.h"

#define SKC_NS SKC_NSPECTRAL
#define SKC_EPAD 32            // spectral rows are padded to 32 floats = one 128 B line
#define SKC_NSTAGES 12

// InfiniteAreaLight's radiance map and Distribution2D on the device (skc_upload_envmap; lights/infinite.cpp, core/montecarlo.h)
struct EnvMapDev {
    const float *rgb = nullptr;                      // [h][w][3] level-0 texels
    const float *condFunc = nullptr, *condCdf = nullptr, *condInt = nullptr;      // [h][w], [h][w + 1], [h]
    const float *margFunc = nullptr, *margCdf = nullptr;                          // [h], [h + 1]
    const float *illum = nullptr;                    // [7][32]: the illuminant basis spectra of Spectrum(RGBSpectrum, SPECTRUM_ILLUMINANT)
    float margInt = 0.f;
    int w = 0, h = 0;
};

// LayeredSkin's "texture albedo" on the device (skc_upload_albedo_texture): the MIPMap pyramid, the EWA weight table, the
// UVMapping2D and the reflectance basis spectra of Spectrum::FromRGB (textures/imagemap.h:98-102)
#define SKC_TEX_MAX_LEVELS 14
struct AlbedoTexDev {
    const float *rgb = nullptr;                    // levels one after the other, [h][w][3] each; nullptr: no texture
    const float *lut = nullptr;                    // [128]
    const float *refl = nullptr;                   // [7][32] white, cyan, magenta, yellow, red, green, blue
    int nLevels = 0;
    int w[SKC_TEX_MAX_LEVELS] = {0}, h[SKC_TEX_MAX_LEVELS] = {0};
    uint32_t off[SKC_TEX_MAX_LEVELS] = {0};        // first float of each level
    float su = 1.f, sv = 1.f, du = 0.f, dv = 0.f;
};

struct skc_ctx {
    int device = 0;
    int smCount = 148;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t timer0 = nullptr, timer1 = nullptr;   // skc_timer_begin / skc_timer_end_ms
    cudaStream_t prefetchStream = nullptr;   // skc_frame_prefetch: sampler stream, closest hits, direct-lighting terms beside Preprocess (lowest priority)
    cudaStream_t auxStream = nullptr;        // octree topology while the irradiance pass runs on `stream` (skc_preprocess*)
    cudaEvent_t defEv0 = nullptr, defEv1 = nullptr;   // deferred stage timing of the irradiance pass
    cudaStream_t copyStream = nullptr;       // host<->device copies of the host-buffer entry points, overlapped with compute
    std::vector<cudaEvent_t> evPool;         // untimed events ordering copyStream against stream
    std::map<void *, size_t> uploadCap;      // bytes allocated behind each uploaded scene array (skc_upload_bvh reuses them)
    bool stageOff = false;                   // skc_stage_begin/end are no-ops (work enqueued on auxStream without host synchronisation: skc_frame_prefetch)
    bool stageAccumulate = false;            // skc_stage_end adds to stageMs instead of overwriting (chunked passes)
    std::string err;
    uint64_t launches = 0;
    float stageMs[SKC_NSTAGES] = {0};

    // --- BVH (skc_bvh.cu)
    float4 *bvhNodes = nullptr;    // 2 x float4 per node
    uint32_t nNodes = 0;
    float4 *bvhWide = nullptr;     // 4-wide form for any-hit rays, 8 x float4 per node (skc_bvh.cuh bvhOccludedWide), or nullptr
    uint32_t nWideNodes = 0;
    bool bvhWideOn = true;         // env SKC_BVH_WIDE=0 keeps the binary walk for any-hit rays too
    float4 *bvhPrims = nullptr;    // 3 x float4 per primitive (triangle: p1,e1,e2 | sphere tag)
    uint32_t nPrims = 0;
    int4 *primInfo = nullptr;      // per primitive: v0,v1,v2 (mesh vertex ids), type
    float *meshP = nullptr, *meshN = nullptr, *meshUV = nullptr;
    int32_t *meshIdx = nullptr;    // vertex indices in MESH triangle order (surface-point tessellation)
    uint32_t meshTris = 0;
    uint32_t nVerts = 0;
    float meshO2W[32];
    int meshFlip = 0;              // ReverseOrientation ^ TransformSwapsHandedness
    skc_light *spheres = nullptr;  // quadric primitives in the BVH
    uint32_t nSpheres = 0;

    // --- material tables (skc_mo.cu)
    float2 *profPairs = nullptr;   // [30][len] {data[i], data[i+1]}
    float *profRcp = nullptr;      // [32]
    float profMaxD2 = 0.f;         // beyond this d^2 every wavelength returns 0
    uint32_t profLen = 0;
    float *rhoHd = nullptr;        // [1025][32]
    // compact shared-abscissa form of the profile tables, copied into shared memory by k_mo_lambda (skc_mo_table.h)
    float *mocRows = nullptr;      // [mocRowCount][32]
    float *mocHeads = nullptr;     // table heads of the wavelengths whose compact window starts above 0
    float4 *mocLam = nullptr;      // [32] {rcp, fLo, fHi, headLen} per wavelength
    uint32_t *mocHeadOff = nullptr;// [32]
    uint32_t mocRowCount = 0, mocHeadFloats = 0;
    float mocInvA = 0.f, mocM = 0.f;
    int moTable = 1;               // 1 = k_mo_lambda (compact table in shared memory + exact fallback) when the table's windows cover enough of the tables (MOC_MIN_COVERAGE), else k_mo_packet; 0 = always k_mo_packet, 2 = always k_mo_lambda; env SKC_MO_TABLE=exact|auto|compact
    float mocCoverage = 0.f, mocMaxErr = 0.f;
    AlbedoTexDev albedoTex;           // "texture albedo" of the skin material; albedoTex.rgb == nullptr: none
    float *albedoTexStore = nullptr;
    AlbedoTexDev bumpTex;             // "texture bumpmap" (r = g = b = the displacement); bumpTex.rgb == nullptr: none
    float *bumpTexStore = nullptr;
    const float4 *hitAlbedo = nullptr;   // per camera hit: the texture's RGB (k_hit_textures), set by skc_frame_render around its Mo pass
    float *alphaTex = nullptr, *alphaUV = nullptr;     // the mesh's alpha texture (skc_upload_alpha) and per-primitive vertex uvs; null: none
    int alphaW = 0, alphaH = 0;
    float alphaMap[4] = {1.f, 1.f, 0.f, 0.f};          // UVMapping2D su, sv, du, dv
    EnvMapDev env;                    // of the scene's infinite light (light record with radius -4); env.rgb == nullptr: none
    float *envStore = nullptr;        // one allocation behind env's pointers
    float *albedoFactors = nullptr;   // [2][32]: Pow(albedo, mix), Pow(albedo, 1 - mix) of a constant "color albedo" (skc_set_albedo_factors); null = Spectrum(1)
    uint64_t moHist[8] = {0};
    int moPacketHits = 0;        // SKC_MO_PACKET_HITS: 0 = by launch size, else 8 / 16 / 32 hits per packet of k_mo_lambda
    uint64_t moGroups[2] = {0, 0};       // last counted pass: (cluster, packet) and (leaf point, packet) accept groups
    uint64_t moLookups[3] = {0, 0, 0};   // last counted pass: profile lookups in range / from shared-memory heads / from global tables
    int smemOptin = 0;             // cudaDevAttrMaxSharedMemoryPerBlockOptin
    int moGroup = 4;               // contributions per round of k_mo_lambda's summation loop; env SKC_MO_GROUP=2|4|8
    int moWarps = 0;               // cap on warps per CTA of k_mo_lambda; env SKC_MO_WARPS
    float mocTol = 1e-5f;          // window tolerance of the compact table; env SKC_MO_TOL

    // --- lights (skc_irradiance.cu)
    skc_light *lights = nullptr;
    uint32_t nLights = 0;

    // --- irradiance point set
    float4 *ipPosArea = nullptr;   // {p.xyz, area}
    float *ipE = nullptr;          // [n][32]
    uint64_t ipCount = 0, ipCap = 0;

    // --- octree (skc_octree.cu)
    float4 *otGeo = nullptr;       // 3 x float4 per node: {p.xyz,sumArea} {min.xyz,max.x} {max.y,max.z,w0,w1}
    float *otEt = nullptr;         // [nNodes][32]
    uint32_t *otPointIdx = nullptr;// leaf-ordered permutation of irradiance points
    float4 *otLeafPosArea = nullptr; // irradiance points gathered in leaf order
    uint32_t otNodes = 0, otCap = 0, otLevels = 0;
    uint32_t lightNS[16] = {0};   // RoundUpPow2(nSamples) per light (host copy for launch shapes)
    int lightKinds[16] = {0};     // 0 sphere emitter, 1 point, 2 spot, 3 distant, 4 infinite (host copy)
    std::vector<uint32_t> otLevelStart;   // first node of each level (+ sentinel), kept between topology and finalize
    bool otTopologyReady = false;
    int moSplit = 0;         // warps per Mo packet (each sums 1/moSplit of the spectrum): 0 = by launch size; env SKC_MO_SPLIT=1|2
    bool hasDeltaLights = false;   // any point / spot / distant light among the uploaded lights (selects the DELTA kernel instances)
    uint32_t bvhDepth = 0;      // depth of the uploaded BVH (longest root-to-leaf path), bounds the traversal stacks
    int moSort = -1;         // spatial sort of the hit points before the packet kernel: -1 = where the caller fixes the order (skc_mo, skc_li_sss, skc_camera_pass*), 0 never, 1 also in the frame path; env SKC_MO_SORT
    int moPacket = 1;        // Mo kernel: 1 = packet (lane = hit), 0 = warp per hit (lane = wavelength); env SKC_MO_KERNEL=warp|packet
    uint64_t otPoints = 0;
    float otBoundsHost[6];

    // --- resident inputs (skc_set_surface_points / skc_set_camera_rays)
    skc_surface_point *resPts = nullptr;
    uint64_t resPtsCount = 0, resPtsCap = 0;
    skc_ray *resRays = nullptr;
    uint64_t resRaysCount = 0;

    // --- scratch
    void *scratch[24] = {nullptr};     // 20 skc_comm host all-gather; 0-7 irradiance / camera pass, 8-10 octree build (the two may run concurrently), 11 tessellation, 12-15 frame, 16 Mo hit sort
    size_t scratchBytes[24] = {0};

    // --- whole-frame state (skc_frame.cu)
    struct skc_frame_state *frame = nullptr;
};

// topology flags word: bits 0-7 child mask (interior) or point count (leaf),
// bit 8 leaf, bit 9 Et black
#define SKC_OT_LEAF 0x100u
#define SKC_OT_BLACK 0x200u

int skc_fail(skc_ctx *ctx, const char *fmt, ...);
int skc_check(skc_ctx *ctx, cudaError_t e, const char *what);
void *skc_scratch(skc_ctx *ctx, int slot, size_t bytes);
void skc_stage_begin(skc_ctx *ctx);
void skc_stage_end(skc_ctx *ctx, int stage);
void skc_frame_release(skc_ctx *ctx);             // frees ctx->frame (skc_frame.cu)
cudaEvent_t skc_event(skc_ctx *ctx, size_t i);   // pooled cudaEventDisableTiming event, nullptr on failure

#define SKC_CUDA(call) do { if (skc_check(ctx, (call), #call)) return 1; } while (0)
#define SKC_LAUNCHED(what) do { ctx->launches++; if (skc_check(ctx, cudaGetLastError(), what)) return 1; } while (0)

// ------------------------------------------------------------------------------------
// device helpers
// ------------------------------------------------------------------------------------
struct f3 { float x, y, z; };
__host__ __device__ inline f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
__host__ __device__ inline f3 sub3(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__host__ __device__ inline f3 add3(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
__host__ __device__ inline f3 mul3(f3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
__host__ __device__ inline float dot3(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// core/geometry.h:529-536: products and differences in double, rounded once to float
__host__ __device__ inline f3 cross3(f3 a, f3 b) {
    double ax = a.x, ay = a.y, az = a.z, bx = b.x, by = b.y, bz = b.z;
    return mk3((float)((ay * bz) - (az * by)), (float)((az * bx) - (ax * bz)), (float)((ax * by) - (ay * bx)));
}
__host__ __device__ inline float len3(f3 a) { return sqrtf(a.x * a.x + a.y * a.y + a.z * a.z); }
// core/geometry.h:102-106,562: v / len == v * (1/len)
__host__ __device__ inline f3 normalize3(f3 a) { float inv = 1.f / len3(a); return mk3(a.x * inv, a.y * inv, a.z * inv); }
__host__ __device__ inline float dist2(f3 a, f3 b) { f3 d = sub3(a, b); return d.x * d.x + d.y * d.y + d.z * d.z; }

/* core/pbrt.h:193-196 redefines M_PI as a FLOAT literal: everything "times pi" is fp32 */
#define SKC_PI_F 3.14159265358979323846f

// ------------------------------------------------------------------------------------
// sinf / cosf exactly as the reference's libm computes them.  The reference calls glibc's float sinf/cosf (UniformSampleCone,
// core/montecarlo.cpp; Beckmann::Sample_f, core/reflection.cpp), which are NOT correctly rounded (0.56 ULP): against the rounded
// fp64 result they differ in 0.09 % (sin) / 0.04 % (cos) of the arguments in [0, 2pi] -- with 16 evaluations per irradiance point
// that made 6 % of the irradiance rows differ from the reference's in the last bit, and at 24.6 M points such one-ulp
// differences flip enough accept/refine decisions of the hierarchical Mo to move 1 % of the hit points by 1e-4..3e-3.  So the
// device evaluates glibc's own algorithm (sysdeps/ieee754/flt-32/s_sincosf.h: fp64 range reduction by pi/2 and two fp64
// polynomials, one rounding to float), operation for operation, without contraction.  Verified against libm's sinf/cosf on
// every one of the 1 086 918 620 floats in [0, 2pi] (tests/test_sincosf.py restates the check on a sample).  |x| >= 120 (never
// reached by the callers: phi in [0, 2pi], theta in [0, pi/2]) falls back to the rounded fp64 result.
__device__ __forceinline__ float skcSinCosPoly(double x, double x2, bool negate, int n) {
    const double sg = negate ? -1.0 : 1.0;
    if ((n & 1) == 0) {
        const double s1c = -0x1.555545995a603p-3, s2c = 0x1.1107605230bc4p-7, s3c = -0x1.994eb3774cf24p-13;
        double x3 = __dmul_rn(x, x2);
        double s1 = __dadd_rn(s2c, __dmul_rn(x2, s3c));
        double x7 = __dmul_rn(x3, x2);
        double s = __dadd_rn(x, __dmul_rn(x3, s1c));
        return (float)__dadd_rn(s, __dmul_rn(x7, s1));
    }
    const double c0 = sg * 0x1p0, c1c = sg * -0x1.ffffffd0c621cp-2, c2c = sg * 0x1.55553e1068f19p-5, c3c = sg * -0x1.6c087e89a359dp-10, c4c = sg * 0x1.99343027bf8c3p-16;
    double x4 = __dmul_rn(x2, x2);
    double c2 = __dadd_rn(c3c, __dmul_rn(x2, c4c));
    double c1 = __dadd_rn(c0, __dmul_rn(x2, c1c));
    double x6 = __dmul_rn(x4, x2);
    double c = __dadd_rn(c1, __dmul_rn(x4, c2c));
    return (float)__dadd_rn(c, __dmul_rn(x6, c2));
}
template <bool COS>
__device__ __forceinline__ float skcSinCosGlibc(float y) {
    const uint32_t top = (__float_as_uint(y) >> 20) & 0x7ffu;
    double x = (double)y;
    if (top < ((0x3f490fdbu >> 20) & 0x7ffu)) {                     // |y| < pi/4 (abstop12 compare)
        if (top < ((0x39800000u >> 20) & 0x7ffu)) return COS ? 1.0f : y;   // |y| < 2^-12
        return skcSinCosPoly(x, __dmul_rn(x, x), false, COS ? 1 : 0);
    }
    if (top < ((0x42f00000u >> 20) & 0x7ffu)) {                     // |y| < 120: reduce_fast
        double r = __dmul_rn(x, 0x1.45F306DC9C883p+23);
        int n = (__double2int_rz(r) + 0x800000) >> 24;
        x = __dsub_rn(x, __dmul_rn((double)n, 0x1.921FB54442D18p0));
        const double sign = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
        return skcSinCosPoly(__dmul_rn(x, sign), __dmul_rn(x, x), (n & 2) != 0, COS ? (n ^ 1) : n);
    }
    return COS ? (float)cos(x) : (float)sin(x);
}
__device__ __forceinline__ float skc_sinf(float x) { return skcSinCosGlibc<false>(x); }
__device__ __forceinline__ float skc_cosf(float x) { return skcSinCosGlibc<true>(x); }

// ---- logf, expf, atanf of the reference's libm (glibc 2.39 on x86-64), restated the same way and verified against libm on
// every float of (0, 1] (logf), [-100, 0] (expf) and [0, inf] (atanf).  They feed Beckmann::Sample_f / D (core/reflection.cpp
// :548-569, reflection.h:516-523), i.e. the rho_hd table (core/multipole.cpp:488-549) and the BSDF sampling of direct lighting.
// A rho_hd table that differs from the reference's in the last bit of some entries makes 6 % of the irradiance rows differ in
// their last bit, which at 24.6 M points is enough to flip accept/refine decisions of Mo (see the note at skc_sinf).
__device__ __forceinline__ float skc_logf(float x) {       // sysdeps/ieee754/flt-32/e_logf.c (N = 16 table, cubic in fp64)
    const double invc[16] = {0x1.661ec79f8f3bep+0, 0x1.571ed4aaf883dp+0, 0x1.49539f0f010bp+0, 0x1.3c995b0b80385p+0, 0x1.30d190c8864a5p+0, 0x1.25e227b0b8eap+0,
                             0x1.1bb4a4a1a343fp+0, 0x1.12358f08ae5bap+0, 0x1.0953f419900a7p+0, 0x1p+0, 0x1.e608cfd9a47acp-1, 0x1.ca4b31f026aap-1,
                             0x1.b2036576afce6p-1, 0x1.9c2d163a1aa2dp-1, 0x1.886e6037841edp-1, 0x1.767dcf5534862p-1};
    const double logc[16] = {-0x1.57bf7808caadep-2, -0x1.2bef0a7c06ddbp-2, -0x1.01eae7f513a67p-2, -0x1.b31d8a68224e9p-3, -0x1.6574f0ac07758p-3, -0x1.1aa2bc79c81p-3,
                             -0x1.a4e76ce8c0e5ep-4, -0x1.1973c5a611cccp-4, -0x1.252f438e10c1ep-5, 0x0p+0, 0x1.aa5aa5df25984p-5, 0x1.c5e53aa362eb4p-4,
                             0x1.526e57720db08p-3, 0x1.bc2860d22477p-3, 0x1.1058bc8a07ee1p-2, 0x1.4043057b6ee09p-2};
    uint32_t ix = __float_as_uint(x);
    if (ix == 0x3f800000u) return 0.f;
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) {
        if (ix * 2 == 0) return -INFINITY;
        if (ix == 0x7f800000u) return x;
        if ((ix & 0x80000000u) || ix * 2 >= 0xff000000u) return NAN;
        ix = __float_as_uint(__fmul_rn(x, 0x1p23f)); ix -= 23u << 23;
    }
    const uint32_t tmp = ix - 0x3f330000u;
    const int i = (tmp >> 19) & 15, k = (int32_t)tmp >> 23;
    const double z = (double)__uint_as_float(ix - (tmp & 0xff800000u));
    const double r = __dsub_rn(__dmul_rn(z, invc[i]), 1.0);
    const double y0 = __dadd_rn(logc[i], __dmul_rn((double)k, 0x1.62e42fefa39efp-1));
    const double r2 = __dmul_rn(r, r);
    double y = __dadd_rn(__dmul_rn(0x1.5575b0be00b6ap-2, r), -0x1.ffffef20a4123p-2);
    y = __dadd_rn(__dmul_rn(-0x1.00ea348b88334p-2, r2), y);
    y = __dadd_rn(__dmul_rn(y, r2), __dadd_rn(y0, r));
    return (float)y;
}
static __device__ const unsigned long long skcExp2Tab[32] = {0x3ff0000000000000ull, 0x3fefd9b0d3158574ull, 0x3fefb5586cf9890full, 0x3fef9301d0125b51ull, 0x3fef72b83c7d517bull, 0x3fef54873168b9aaull, 0x3fef387a6e756238ull, 0x3fef1e9df51fdee1ull, 0x3fef06fe0a31b715ull, 0x3feef1a7373aa9cbull, 0x3feedea64c123422ull, 0x3feece086061892dull, 0x3feebfdad5362a27ull, 0x3feeb42b569d4f82ull, 0x3feeab07dd485429ull, 0x3feea47eb03a5585ull, 0x3feea09e667f3bcdull, 0x3fee9f75e8ec5f74ull, 0x3feea11473eb0187ull, 0x3feea589994cce13ull, 0x3feeace5422aa0dbull, 0x3feeb737b0cdc5e5ull, 0x3feec49182a3f090ull, 0x3feed503b23e255dull, 0x3feee89f995ad3adull, 0x3feeff76f2fb5e47ull, 0x3fef199bdd85529cull, 0x3fef3720dcef9069ull, 0x3fef5818dcfba487ull, 0x3fef7c97337b9b5full, 0x3fefa4afa2a490daull, 0x3fefd0765b6e4540ull};   // asuint64(2^(i/32)) - (i << 52) / 32
// sysdeps/ieee754/flt-32/e_expf.c; libm selects its FMA build on this CPU family (sysdeps/x86_64/fpu/multiarch/e_expf.c): the
// scaled argument enters the rounding-to-integer step and the remainder through fused multiply-adds
__device__ __forceinline__ float skc_expf(float x) {
    const uint32_t abstop = (__float_as_uint(x) >> 20) & 0x7ffu;
    if (abstop >= 0x42bu) {                       // |x| >= 88
        if (__float_as_uint(x) == 0xff800000u) return 0.f;
        if (abstop >= 0x7f8u) return x + x;
        if (x > 0x1.62e42ep6f) return INFINITY;
        if (x < -0x1.9fe368p6f) return 0.f;
    }
    const double inv = 0x1.71547652b82fep+0 * 32, shift = 0x1.8p52, xd = (double)x;
    const double z = __dmul_rn(inv, xd);
    double kd = __fma_rn(inv, xd, shift);
    const unsigned long long ki = (unsigned long long)__double_as_longlong(kd);
    kd = __dsub_rn(kd, shift);
    const double r = __fma_rn(inv, xd, -kd);
    (void)z;
    const double s = __longlong_as_double((long long)(skcExp2Tab[ki & 31] + (ki << 47)));
    const double C0 = 0x1.c6af84b912394p-5 / 32 / 32 / 32, C1 = 0x1.ebfce50fac4f3p-3 / 32 / 32, C2 = 0x1.62e42ff0c52d6p-1 / 32;
    const double zz = __dadd_rn(__dmul_rn(C0, r), C1), r2 = __dmul_rn(r, r);
    double y = __dadd_rn(__dmul_rn(C2, r), 1.0);
    y = __dadd_rn(__dmul_rn(zz, r2), y);
    return (float)__dmul_rn(y, s);
}
__device__ __forceinline__ float skc_atanf(float x) {      // sysdeps/ieee754/flt-32/s_atanf.c (fdlibm, fp32 arithmetic, no contraction)
    const float atanhi[4] = {4.6364760399e-01f, 7.8539812565e-01f, 9.8279368877e-01f, 1.5707962513e+00f};
    const float atanlo[4] = {5.0121582440e-09f, 3.7748947079e-08f, 3.4473217170e-08f, 7.5497894159e-08f};
    const float aT[11] = {3.3333334327e-01f, -2.0000000298e-01f, 1.4285714924e-01f, -1.1111110449e-01f, 9.0908870101e-02f, -7.6918758452e-02f,
                          6.6610731184e-02f, -5.8335702866e-02f, 4.9768779427e-02f, -3.6531571299e-02f, 1.6285819933e-02f};
    const int32_t hx = __float_as_int(x), ix = hx & 0x7fffffff;
    int id;
    if (ix >= 0x4c000000) { if (ix > 0x7f800000) return x + x; return hx > 0 ? __fadd_rn(atanhi[3], atanlo[3]) : __fsub_rn(-atanhi[3], atanlo[3]); }
    if (ix < 0x3ee00000) { if (ix < 0x31000000) return x; id = -1; }
    else {
        x = fabsf(x);
        if (ix < 0x3f980000) {
            if (ix < 0x3f300000) { id = 0; x = __fdiv_rn(__fsub_rn(__fmul_rn(2.0f, x), 1.0f), __fadd_rn(2.0f, x)); }
            else { id = 1; x = __fdiv_rn(__fsub_rn(x, 1.0f), __fadd_rn(x, 1.0f)); }
        } else {
            if (ix < 0x401c0000) { id = 2; x = __fdiv_rn(__fsub_rn(x, 1.5f), __fadd_rn(1.0f, __fmul_rn(1.5f, x))); }
            else { id = 3; x = __fdiv_rn(-1.0f, x); }
        }
    }
    const float z = __fmul_rn(x, x), w = __fmul_rn(z, z);
    float a = __fmul_rn(w, aT[10]);
    a = __fmul_rn(w, __fadd_rn(aT[8], a)); a = __fmul_rn(w, __fadd_rn(aT[6], a)); a = __fmul_rn(w, __fadd_rn(aT[4], a)); a = __fmul_rn(w, __fadd_rn(aT[2], a));
    const float s1 = __fmul_rn(z, __fadd_rn(aT[0], a));
    float b = __fmul_rn(w, aT[9]);
    b = __fmul_rn(w, __fadd_rn(aT[7], b)); b = __fmul_rn(w, __fadd_rn(aT[5], b)); b = __fmul_rn(w, __fadd_rn(aT[3], b));
    const float s2 = __fmul_rn(w, __fadd_rn(aT[1], b));
    if (id < 0) return __fsub_rn(x, __fmul_rn(x, __fadd_rn(s1, s2)));
    const float zr = __fsub_rn(atanhi[id], __fsub_rn(__fsub_rn(__fmul_rn(x, __fadd_rn(s1, s2)), atanlo[id]), x));
    return hx < 0 ? -zr : zr;
}

// powf as glibc 2.39 computes it (sysdeps/ieee754/flt-32/e_powf.c, the FMA build): log2(x) from a 16-entry table and a degree-5
// polynomial in double, y * log2(x), then exp2 from skc_expf's 2^(i/32) table and a cubic, one rounding to float.  The tables are the
// system libm's own (__powf_log2_data); checked against libm's powf on 3e8 arguments on the CPU, 0 differences.  x >= 0, y >= 0
// finite (Pow(Spectrum, float) of reflectances: core/spectrum.h).
static __device__ const double skcPowLog2Tab[16][2] = {
  {0x1.661ec79f8f3bep+0, -0x1.efec65b963019p-2},
  {0x1.571ed4aaf883dp+0, -0x1.b0b6832d4fca4p-2},
  {0x1.49539f0f010b0p+0, -0x1.7418b0a1fb77bp-2},
  {0x1.3c995b0b80385p+0, -0x1.39de91a6dcf7bp-2},
  {0x1.30d190c8864a5p+0, -0x1.01d9bf3f2b631p-2},
  {0x1.25e227b0b8ea0p+0, -0x1.97c1d1b3b7af0p-3},
  {0x1.1bb4a4a1a343fp+0, -0x1.2f9e393af3c9fp-3},
  {0x1.12358f08ae5bap+0, -0x1.960cbbf788d5cp-4},
  {0x1.0953f419900a7p+0, -0x1.a6f9db6475fcep-5},
  {0x1.0000000000000p+0, 0x0.0p+0},
  {0x1.e608cfd9a47acp-1, 0x1.338ca9f24f53dp-4},
  {0x1.ca4b31f026aa0p-1, 0x1.476a9543891bap-3},
  {0x1.b2036576afce6p-1, 0x1.e840b4ac4e4d2p-3},
  {0x1.9c2d163a1aa2dp-1, 0x1.40645f0c6651cp-2},
  {0x1.886e6037841edp-1, 0x1.88e9c2c1b9ff8p-2},
  {0x1.767dcf5534862p-1, 0x1.ce0a44eb17bccp-2}
};
__device__ __forceinline__ float skc_powf(float x, float y) {
    if (x == 0.f) return y == 0.f ? 1.f : 0.f;
    if (y == 0.f) return 1.f;
    uint32_t ix = __float_as_uint(x);
    if (ix < 0x00800000u) { ix = __float_as_uint(__fmul_rn(x, 8388608.f)) & 0x7fffffffu; ix -= 23u << 23; }      // subnormal x
    const uint32_t tmp = ix - 0x3f330000u, top = tmp & 0xff800000u;
    const int i = (int)((tmp >> 19) & 15u), k = (int)top >> 23;
    const double invc = skcPowLog2Tab[i][0], logc = skcPowLog2Tab[i][1];
    const double z = (double)__uint_as_float(ix - top);
    const double r = __fma_rn(z, invc, -1.0);
    const double y0 = __dadd_rn(logc, (double)k);
    const double A0 = 0x1.27616c9496e0bp-2, A1 = -0x1.71969a075c67ap-2, A2 = 0x1.ec70a6ca7baddp-2, A3 = -0x1.7154748bef6c8p-1, A4 = 0x1.71547652ab82bp+0;
    const double r2 = __dmul_rn(r, r);
    double yy = __fma_rn(A0, r, A1);
    const double p = __fma_rn(A2, r, A3);
    const double r4 = __dmul_rn(r2, r2);
    double q = __fma_rn(A4, r, y0);
    q = __fma_rn(p, r2, q);
    yy = __fma_rn(yy, r4, q);
    const double ylogx = __dmul_rn((double)y, yy);
    if (fabs(ylogx) >= 126.0) return ylogx > 0 ? __int_as_float(0x7f800000) : 0.f;      // out of the callers' range; not the reference's edge handling
    const double SHIFT = 0x1.8p+52 / 32, C0 = 0x1.c6af84b912394p-5, C1 = 0x1.ebfce50fac4f3p-3, C2 = 0x1.62e42ff0c52d6p-1;
    double kd = __dadd_rn(ylogx, SHIFT);
    const unsigned long long ki = (unsigned long long)__double_as_longlong(kd);
    kd = __dsub_rn(kd, SHIFT);
    const double rr = __dsub_rn(ylogx, kd);
    const double s = __longlong_as_double((long long)(skcExp2Tab[ki & 31] + (ki << 47)));
    const double zz = __fma_rn(C0, rr, C1), rr2 = __dmul_rn(rr, rr);
    double w = __fma_rn(C2, rr, 1.0);
    w = __fma_rn(zz, rr2, w);
    return (float)__dmul_rn(w, s);
}

__device__ __forceinline__ float skc_acosf(float x) {      // sysdeps/ieee754/flt-32/e_acosf.c (fdlibm, fp32, no contraction); tools/libm_check.c
    const float pi = 3.1415925026e+00f, pio2_hi = 1.5707962513e+00f, pio2_lo = 7.5497894159e-08f;
    const float pS0 = 1.6666667163e-01f, pS1 = -3.2556581497e-01f, pS2 = 2.0121252537e-01f, pS3 = -4.0055535734e-02f, pS4 = 7.9153501429e-04f, pS5 = 3.4793309169e-05f;
    const float qS1 = -2.4033949375e+00f, qS2 = 2.0209457874e+00f, qS3 = -6.8828397989e-01f, qS4 = 7.7038154006e-02f;
    const int32_t hx = __float_as_int(x), ix = hx & 0x7fffffff;
    if (ix == 0x3f800000) return hx > 0 ? 0.0f : __fadd_rn(pi, __fmul_rn(2.0f, pio2_lo));
    if (ix > 0x3f800000) return __int_as_float(0x7fc00000);
    auto P = [&](float z) { return __fmul_rn(z, __fadd_rn(pS0, __fmul_rn(z, __fadd_rn(pS1, __fmul_rn(z, __fadd_rn(pS2, __fmul_rn(z, __fadd_rn(pS3, __fmul_rn(z, __fadd_rn(pS4, __fmul_rn(z, pS5))))))))))); };
    auto Q = [&](float z) { return __fadd_rn(1.0f, __fmul_rn(z, __fadd_rn(qS1, __fmul_rn(z, __fadd_rn(qS2, __fmul_rn(z, __fadd_rn(qS3, __fmul_rn(z, qS4)))))))); };
    if (ix < 0x3f000000) {
        if (ix <= 0x23000000) return __fadd_rn(pio2_hi, pio2_lo);
        const float z = __fmul_rn(x, x), r = __fdiv_rn(P(z), Q(z));
        return __fsub_rn(pio2_hi, __fsub_rn(x, __fsub_rn(pio2_lo, __fmul_rn(x, r))));
    }
    if (hx < 0) {
        const float z = __fmul_rn(__fadd_rn(1.0f, x), 0.5f), s = __fsqrt_rn(z), r = __fdiv_rn(P(z), Q(z));
        const float w = __fsub_rn(__fmul_rn(r, s), pio2_lo);
        return __fsub_rn(pi, __fmul_rn(2.0f, __fadd_rn(s, w)));
    }
    const float z = __fmul_rn(__fsub_rn(1.0f, x), 0.5f), s = __fsqrt_rn(z);
    const float df = __int_as_float(__float_as_int(s) & 0xfffff000);
    const float c = __fdiv_rn(__fsub_rn(z, __fmul_rn(df, df)), __fadd_rn(s, df));
    const float r = __fdiv_rn(P(z), Q(z)), w = __fadd_rn(__fmul_rn(r, s), c);
    return __fmul_rn(2.0f, __fadd_rn(df, w));
}
__device__ __forceinline__ float skc_atan2f(float y, float x) {      // sysdeps/ieee754/flt-32/e_atan2f.c (fdlibm); tools/libm_check.c
    const float tiny = 1.0e-30f, pi_o_4 = 7.8539818525e-01f, pi_o_2 = 1.5707963705e+00f, pi = 3.1415927410e+00f, pi_lo = -8.7422776573e-08f;
    const int32_t hx = __float_as_int(x), hy = __float_as_int(y), ix = hx & 0x7fffffff, iy = hy & 0x7fffffff;
    if (ix > 0x7f800000 || iy > 0x7f800000) return x + y;
    if (hx == 0x3f800000) return skc_atanf(y);
    const int m = ((hy >> 31) & 1) | ((hx >> 30) & 2);
    if (iy == 0) return m < 2 ? y : (m == 2 ? __fadd_rn(pi, tiny) : __fsub_rn(-pi, tiny));
    if (ix == 0) return hy < 0 ? __fsub_rn(-pi_o_2, tiny) : __fadd_rn(pi_o_2, tiny);
    if (ix == 0x7f800000) {
        if (iy == 0x7f800000) return m == 0 ? __fadd_rn(pi_o_4, tiny) : m == 1 ? __fsub_rn(-pi_o_4, tiny) : m == 2 ? __fadd_rn(__fmul_rn(3.0f, pi_o_4), tiny) : __fsub_rn(__fmul_rn(-3.0f, pi_o_4), tiny);
        return m == 0 ? 0.0f : m == 1 ? -0.0f : m == 2 ? __fadd_rn(pi, tiny) : __fsub_rn(-pi, tiny);
    }
    if (iy == 0x7f800000) return hy < 0 ? __fsub_rn(-pi_o_2, tiny) : __fadd_rn(pi_o_2, tiny);
    const int k = (iy - ix) >> 23;
    float z;
    if (k > 60) z = __fadd_rn(pi_o_2, __fmul_rn(0.5f, pi_lo));
    else if (hx < 0 && k < -60) z = 0.0f;
    else z = skc_atanf(fabsf(__fdiv_rn(y, x)));
    if (m == 0) return z;
    if (m == 1) return __int_as_float(__float_as_int(z) ^ 0x80000000);
    if (m == 2) return __fsub_rn(pi, __fsub_rn(z, pi_lo));
    return __fsub_rn(__fsub_rn(z, pi_lo), pi);
}

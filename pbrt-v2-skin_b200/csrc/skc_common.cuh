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
    int moTable = 1;               // 1 = compact table in shared memory + exact fallback (k_mo_lambda), 0 = exact tables only (k_mo_packet); env SKC_MO_TABLE=exact|compact
    float mocCoverage = 0.f, mocMaxErr = 0.f;
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
    float *otLeafE = nullptr;        // [n][32] in leaf order
    uint32_t otNodes = 0, otCap = 0, otLevels = 0;
    uint32_t lightNS[16] = {0};   // RoundUpPow2(nSamples) per light (host copy for launch shapes)
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
    void *scratch[20] = {nullptr};     // 0-7 irradiance / camera pass, 8-10 octree build (the two may run concurrently), 11 tessellation, 12-15 frame, 16 Mo hit sort
    size_t scratchBytes[20] = {0};

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

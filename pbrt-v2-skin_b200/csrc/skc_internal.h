// skc_internal.h -- device-pointer entry points shared between translation units.
#pragma once
#include "skc_common.cuh"
#include "skc_bvh.cuh"

BvhView skc_bvh_view(const skc_ctx *ctx);
MeshView skc_mesh_view(const skc_ctx *ctx);
int skc_trace_any_dev(skc_ctx *ctx, const skc_ray *raysDev, uint64_t n, uint8_t *outDev, unsigned long long *cntDev);
int skc_trace_closest_dev(skc_ctx *ctx, const skc_ray *raysDev, uint64_t n, skc_hit *outDev, unsigned long long *cntDev, float4 *frameDev = nullptr);
int skc_mo_dev(skc_ctx *ctx, const float *p3Dev, uint64_t strideFloats, uint64_t n, float maxError, float *MoDev /*[n][32]*/);
int skc_ensure_points(skc_ctx *ctx, uint64_t n);
// sort-based octree topology (skc_octree_sort.cu): 0 = done, 2 = tree deeper than the path keys (fall back), 1 = error
int skc_octree_topology_sorted(skc_ctx *ctx, uint32_t n, const float *hb, uint32_t cap, std::vector<uint32_t> &levelStart, uint32_t *totalOut);
// fused Mo + Li subsurface term for device hits: L rows [n][30] into dL (skc_mo.cu)
int skc_li_fused_dev(skc_ctx *ctx, const skc_hit *dHits, uint64_t n, float maxError, float mix, float *dL);
// device film {Lxyz, weightSum} per pixel as a flat float array (skc_frame.cu), nullptr before skc_frame_setup
void *skc_frame_film_floats(skc_ctx *ctx, uint64_t *floats);

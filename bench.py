#!/usr/bin/env python3
"""bench.py -- the subsurface hot path of pbrt-v2-skin on B200 (contract: see DESIGN.md section 6).

One "step" = one FRAME of the reconstructed skin.scene (BASELINE.json configs[1]), i.e. SamplerRenderer::Render
without the material construction:
    MultipoleSubsurfaceIntegrator::Preprocess  (surface points by tessellation of the mesh; irradiance pass
        over every point: light sampling + shadow rays through the BVH; octree build + InitHierarchy)
  + the camera pass                            (LDSampler stream, camera rays, closest hit, octree Mo + the Li
        subsurface term, emission, direct lighting with the LayeredSkin BSDF and MIS, film accumulation)
metric  = shading points per second = (irradiance points + camera samples) / step time
value   = scene (mesh, BVH, lights, material tables) resident in HBM before the timed region
e2e     = the same from HOST buffers: the scene (BVH nodes, primitives, mesh, lights) is uploaded from host memory
          every step, the film comes back to the host and is converted to RGB there (ImageFilm::WriteImage's
          arithmetic) -- what `pbrt scene.pbrt` does per frame after parsing
N > 1   = points sharded over ranks, in-place NCCL all-gather of the irradiance set, replicated octree build,
          render tasks (tiles) dealt round-robin over the ranks, film reduced (sum) to rank 0 over NCCL -- both
          collectives are C-ABI calls into libskincuda.so (skc_comm_*); torch.distributed (gloo) only carries the
          communicator id at start-up (strong scaling: the scene is fixed).

`--impl reference` times the reference's own CPU code (oracle/_ref/ref_driver bench: Preprocess + its
SamplerRendererTask loop with the full Li) on a bounded sample of the same scene (1/16 of the points and of the
film), all host threads.
"""
import argparse
import json
import os
import re
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "skin.pbrt subsurface hot path: shading points/s (irradiance points + camera samples per second of Preprocess [surface points + irradiance + octree] + Li)"
UNIT = "shading points/s"
DATA = "scene file: the reference's own head mesh (scenes/geometry/head.pbrt) in the reconstructed skin.pbrt (SURVEY.md F1/F2); no external dataset, material tables generated at set-up"
WORKLOAD = "scenes/skin.scene (reconstructed skin.pbrt: head mesh Scale 10, 2 sphere lights x 4 samples, minsampledistance 0.0015, 1280x720x4spp, maxerror 0.1)"


def clocks_sampler(stop, out):
    q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    try:
        p = subprocess.Popen(["nvidia-smi", "--query-gpu=" + q, "--format=csv,noheader,nounits", "-lms", "200", "-i", os.environ.get("LOCAL_RANK", "0")],
                             stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
    except Exception:
        return
    def reader():
        for line in p.stdout:
            out.append(line.strip())
    t = threading.Thread(target=reader, daemon=True); t.start()
    stop.wait()
    p.terminate()
    t.join(timeout=2)


def summarise_clocks(lines):
    sm, mx, reasons = [], 0.0, set()
    for l in lines:
        f = [x.strip() for x in l.split(",")]
        if len(f) < 7:
            continue
        try:
            sm.append(float(f[0])); mx = max(mx, float(f[1]))
        except ValueError:
            continue
        for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
            if v.lower().startswith("active"):
                reasons.add(name)
    return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def build_material(scene, dev):
    """LayeredSkin construction on the device: 30 two-layer diffusion profiles (skc_profile_batch) and the rho_hd table
    (skc_rho_table), uploaded into the context.  Untimed set-up (once per material, not per frame); its device time is
    reported as material_ms.  When the checker's export of the same scene is around (container runs), compare."""
    t0 = time.time()
    hdr, rd = scene.build_profile(dev, None, upload=True)
    info = {"source": "generated on the device (skc_profile_batch + skc_rho_table)", "profile_len": int(rd.shape[1]),
            "profile_ms": round(dev.stage_ms("profile"), 3), "rho_ms": round(dev.stage_ms("rho"), 3), "wall_s": round(time.time() - t0, 3)}
    path = os.path.join(ROOT, "scenes", "_generated", "skin_tables.npz")
    if os.path.exists(path) and os.path.basename(scene.path) == "skin.scene":
        z = np.load(path)
        if z["profile_rd"].shape == rd.shape:
            ref = z["profile_rd"].astype(np.float64)
            floor = np.abs(ref).max(axis=1, keepdims=True) * 1e-9
            info["vs_reference_tables_max_rel"] = float((np.abs(rd - ref) / np.maximum(np.abs(ref), floor)).max())
    return info


def reference_arm(args):
    """The reference's own CPU implementation on the host cores (rank 0 only)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    drv = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
    if not os.path.exists(drv):
        emit({"impl": "reference", "unavailable": "oracle/_ref/ref_driver not built (oracle/build_ref.sh needs /root/reference)"})
        return
    # Same config as the B200 arm: whole frames of scenes/skin.scene.  One costs the reference 24.5 s on the box's 16 host cores
    # (preprocess 14.8 s, render 9.6 s: profiles/r2_reference_full_step.json), so K timed steps take K x 25 s; warm-up does
    # nothing for this CPU code beyond paging the binary in, and is capped at one step to keep the driver's 20 + 5 run under
    # nine minutes.  --sample16 times the bounded 1/16-scale sample of the same scene instead (1.35 s per step).
    full = not args.sample16
    sample = os.path.join(ROOT, "scenes", "skin.scene" if full else "skin_sample16.scene")
    cores = os.cpu_count() or 1
    warm = min(args.warmup, 1) if full else args.warmup
    out = subprocess.run([drv, "bench", sample, "--ncores", str(cores), "--steps", str(args.steps), "--warmup", str(warm)],
                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, check=True).stdout
    steps = [l.split() for l in out.splitlines() if l.startswith("step ") and int(l.split()[1]) >= 0]
    units = sum(int(s[3]) + int(s[5]) for s in steps)
    secs = sum(float(s[7]) + float(s[9]) for s in steps)
    value = units / secs
    what = "reference Preprocess + its SamplerRendererTask loop (sampler, camera rays, Scene::Intersect, full Li, film AddSample), %d steps" % len(steps)
    if full:
        desc = "skin.scene at FULL scale (%s irradiance points, 1280x720x4spp -> %s camera samples; preprocess %.1f s + render %.1f s per step), %s" % (
            steps[0][3], steps[0][5], sum(float(s[7]) for s in steps) / len(steps), sum(float(s[9]) for s in steps) / len(steps), what)
    else:
        desc = "skin.scene at 1/16 scale (minsampledistance 0.006 -> %s irradiance points, 320x180x4spp -> %s camera samples), %s" % (steps[0][3], steps[0][5], what)
    emit({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": warm,
        "ms_per_step": 1e3 * secs / max(1, len(steps)), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": DATA, "config": {"workload": WORKLOAD, "sample": desc},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference", "sample": desc},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


def claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries loaded later write there too (NCCL prints its version line to
    stdout when NCCL_DEBUG is set on the box), so file descriptor 1 is pointed at stderr for the run and the JSON line goes
    to the saved descriptor."""
    global _OUT
    sys.stdout.flush()
    _OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def emit(obj):
    _OUT.write(json.dumps(obj) + "\n")
    _OUT.flush()


_OUT = sys.stdout


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--scene", default=os.path.join(ROOT, "scenes", "skin.scene"))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--full-scale", action="store_true", help="--impl reference: whole frames of scenes/skin.scene (the default; kept for old command lines)")
    ap.add_argument("--sample16", action="store_true", help="--impl reference: time the bounded 1/16-scale sample of the scene instead of whole frames")
    ap.add_argument("--overlap-topology", action="store_true", help="N=1 experiment: octree topology on the auxiliary stream beside the irradiance pass (skc_preprocess_resident); measured 95.4 vs 94.4 ms, so not the default")
    ap.add_argument("--no-prefetch", action="store_true", help="staged flow: sampler stream, closest hits and direct lighting after Preprocess instead of beside it")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    import __graft_entry__ as ge
    pkg = ge.load_package()
    dist = comm = None
    if world > 1:
        # the launcher's rendezvous (torchrun env) on a CPU backend: it carries the 128-byte NCCL id of rank 0 to the other ranks
        # and nothing else -- every exchange of the data path is an NCCL call inside libskincuda.so (csrc/skc_comm.cu)
        import torch.distributed as dist
        dist.init_process_group("gloo")
    dev = pkg.SkinDevice(local)
    if world > 1:
        comm = pkg.SkinComm(dev, rank, world, pkg.exchange_unique_id(dist, rank, pkg.SkinComm.unique_id))

    # ---- workload (host side, untimed): scene, BVH, surface points, camera rays, tables
    scene = pkg.HostScene(args.scene)
    scene.upload(dev)
    material = build_material(scene, dev)
    n_pts = dev.tessellate_resident(scene.minsampledistance, material_id=0)     # also done inside every step
    n_tasks = scene.irradiance_tasks(n_pts, 1)
    # rank shards (contiguous, equal padded size so the all-gather can run in place)
    p_first, p_count, shard = pkg.shard_range(n_pts, rank, world)
    # camera samples: the reference's render tasks (sub-windows of the film), dealt round-robin over the ranks
    r_count = dev.frame_setup(scene.frame_desc(n_cores=1, rank=rank, world=world))
    r_first = 0
    n_rays = r_count
    if world > 1:
        dev.irradiance_dev(shard * world)          # the point arrays sized for the padded exchange
        import torch
        t = torch.tensor([r_count], dtype=torch.int64); dist.all_reduce(t); n_rays = int(t.item())      # (a count for the report)

    def exchange_and_build():
        """Irradiance exchange + octree.  N > 1: in-place ncclAllGather of the shards; the positions (16 B/point) go first on the
        main stream, then the irradiance rows (128 B/point) travel on the copy stream while every rank already builds the octree
        topology from the positions; the finalize pass (leaf gather + InitHierarchy) waits for the rows on the device."""
        if world == 1:
            dev.build_octree()
            return
        comm.allgather_positions(shard)
        comm.allgather_rows_begin(shard)
        # padded tail (only when n_pts % world != 0): shards are contiguous and only the LAST one is short,
        # so the valid points are exactly [0, n_pts)
        dev.set_point_count(n_pts)
        dev.build_octree_topology()
        comm.allgather_rows_wait()
        dev.build_octree_finalize()

    def step_resident():
        if not args.no_prefetch:
            dev.frame_prefetch()                # sampler stream, camera rays, closest hits, direct-lighting terms: second stream, beside Preprocess
        if world == 1:
            dev.tessellate_resident(scene.minsampledistance, material_id=0)
        else:
            dev.tessellate_range(scene.minsampledistance, p_first, p_count, material_id=0)      # this rank's shard of the points
        if world == 1 and args.overlap_topology:
            dev.preprocess_resident(p_first, p_count, n_tasks, scene.mix)       # irradiance pass on the main stream, octree topology (positions only) beside it
            dev.build_octree_finalize()
        else:
            dev.irradiance_resident(p_first, p_count, n_tasks, scene.mix)
            exchange_and_build()
        if args.no_prefetch:
            dev.frame_samples()
        dev.frame_film_clear()
        dev.frame_render(r_first, r_count, scene.maxerror, scene.mix)
        if world > 1:
            comm.reduce_film(0)

    h2d_bytes = sum(a.nbytes for a in (scene.nodes, scene.prims, scene.P, scene.idx, scene.lights, scene.lights) if a is not None) \
        + sum(a.nbytes for a in (scene.N, scene.uv) if a is not None)
    d2h_bytes = scene.xres * scene.yres * 12 if rank == 0 else 0
    e2e_image = {}
    rgb_host = dev.pinned((scene.yres, scene.xres, 3), np.float32) if rank == 0 else None       # pinned: the image comes back at full PCIe rate

    def step_e2e():
        scene.upload(dev)                       # BVH nodes, primitives, mesh, lights from host memory
        step_resident()
        if rank == 0:
            e2e_image["rgb"] = dev.frame_rgb(out=rgb_host)      # ImageFilm::WriteImage's pixel arithmetic on the device, RGB image to the host

    def barrier():
        dev.synchronize()
        if world > 1:
            comm.barrier()

    def timed(fn, k):
        barrier()
        dev.timer_begin()
        for _ in range(k):
            fn()
        ms = dev.timer_end_ms()
        barrier()
        return comm.max_f32(ms) if world > 1 else ms

    for _ in range(args.warmup):
        step_resident()
    stop, lines = threading.Event(), []
    th = threading.Thread(target=clocks_sampler, args=(stop, lines), daemon=True); th.start()
    launches0 = dev.launch_count()
    stage_acc = {}
    barrier()
    dev.timer_begin()
    for _ in range(args.steps):
        step_resident()
        for s in ("tessellate", "scramble", "irradiance", "octree", "samples", "closest", "mo", "direct", "film"):
            stage_acc[s] = stage_acc.get(s, 0.0) + dev.stage_ms(s)
    ms_total = dev.timer_end_ms()
    barrier()
    launches = dev.launch_count() - launches0
    if world > 1:
        ms_total = comm.max_f32(ms_total)
    for _ in range(min(2, args.warmup)):
        step_e2e()
    ms_e2e = timed(step_e2e, args.steps)
    stop.set(); th.join(timeout=3)

    units = n_pts + n_rays
    ms_step = ms_total / args.steps
    value = units / (ms_step * 1e-3)
    e2e_value = units / (ms_e2e / args.steps * 1e-3)
    stage_ms = {k: v / args.steps for k, v in stage_acc.items()}

    # ---- roofline of the dominant kernel (one extra instrumented pass, untimed)
    c_irr = dev.irradiance_resident(p_first, p_count, n_tasks, scene.mix, counters=True)
    exchange_and_build()
    c_cam = dev.camera_pass_resident(r_first, r_count, scene.maxerror, scene.mix, counters=True)
    n_lights = scene.n_lights
    mo_stats = dev.mo_table_stats()
    mo_kernel = "k_mo_lambda" if mo_stats["rows"] else "k_mo_packet"
    # ALGORITHMIC bytes, SURVEY.md 8(d): Mo per hit 12 (p in) + 120 (Mo out) + 16 per node visited + 120 per accepted cluster (Et)
    #   + 136 per contributing leaf point (p, area, E) + 240 per contribution (30 wavelengths x 2 table floats);
    #   irradiance per point 44 in + 136 out + 12*nLights scramble + per shadow ray 40 + 32 per slab test + 48 per primitive test + 1 out
    V, leaf_pts, C_int, C_pt = int(c_cam[2]), int(c_cam[3]), int(c_cam[4]), int(c_cam[5])
    mo_bytes = r_count * (12 + 120) + 16 * V + 120 * C_int + 136 * C_pt + 240 * (C_int + C_pt)
    irr_bytes = p_count * (44 + 136 + 12 * n_lights) + 41 * int(c_irr[2]) + 32 * int(c_irr[0]) + 48 * int(c_irr[1])
    # the same in this repo's own layout: 48 B per node visit, 16 B per leaf point tested, 128 B rows,
    # and 8 B per profile lookup that leaves shared memory (the compact table serves the rest)
    mo_layout_bytes = r_count * (12 + 128) + 48 * V + 16 * leaf_pts + 128 * (C_int + C_pt) + 8 * (mo_stats["from_global"] if mo_stats["rows"] else int(c_cam[6]))
    peak, peak_kind = measured_peak()
    dom = "irradiance" if stage_ms["irradiance"] >= stage_ms["mo"] else "mo"
    dom_bytes = irr_bytes if dom == "irradiance" else mo_bytes
    achieved = dom_bytes / (stage_ms[dom] * 1e-3) / 1e9
    # DRAM traffic of the dominant kernel per launch: from the ncu --set full capture of this same single-GPU step that is
    # committed under profiles/ (a profiler cannot run inside the timed run); a shard of the launch at N > 1 has no capture: null
    traffic, traffic_src = None, None
    if world == 1 and dom == "mo":
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "r2_dram_traffic.json")))
            traffic, traffic_src = tj.get(mo_kernel), tj.get("source")
        except Exception:
            pass
    roofline = {"bound": "hbm", "kernel": "k_irr_sample + k_irr_shadow + k_irr_spectral" if dom == "irradiance" else mo_kernel, "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": peak_kind + " (MEASURED_PEAKS.json hbm_gbs, copy bandwidth)" if peak_kind == "measured" else "fallback 6650 GB/s (B200_PROFILING.md)",
                "algorithmic_bytes_per_launch": dom_bytes, "formula": "SURVEY.md 8(d)", "launch_ms": stage_ms[dom],
                "note": "algorithmic bytes give no credit for cache reuse: the octree top, rows shared by neighbouring samples and the profile tables (shared memory) never reach DRAM, so achieved/peak describes work rate, not DRAM load",
                "other": {"k_irradiance": {"GBps": irr_bytes / (stage_ms["irradiance"] * 1e-3) / 1e9, "ms": stage_ms["irradiance"],
                                           "shadow_rays": int(c_irr[2]), "slab_tests": int(c_irr[0]), "prim_tests": int(c_irr[1])},
                          "k_mo": {"GBps": mo_bytes / (stage_ms["mo"] * 1e-3) / 1e9, "ms": stage_ms["mo"], "nodes_visited": V, "leaf_points_tested": leaf_pts,
                                   "cluster_contribs": C_int, "point_contribs": C_pt, "profile_lookups": int(c_cam[6]),
                                   "layout_bytes_per_launch": mo_layout_bytes, "layout_GBps": mo_layout_bytes / (stage_ms["mo"] * 1e-3) / 1e9,
                                   "compact_table": mo_stats}}}
    # Scene::IntersectP / Scene::Intersect alone over the frame's camera rays (untimed extra passes; SURVEY.md 8d asks for both)
    rays = {"irradiance_pass_shadow_rays_any_hit": int(c_irr[2]) * world / (stage_ms["irradiance"] * 1e-3)}
    try:
        dev.trace_resident(0, r_count, any_hit=True); t_any = dev.stage_ms("any")
        dev.trace_resident(0, r_count, any_hit=False); t_cl = dev.stage_ms("closest")
        rays["camera_rays_any_hit"] = r_count / (t_any * 1e-3); rays["camera_rays_closest_hit"] = r_count / (t_cl * 1e-3)
        rays["camera_rays"] = int(r_count)
    except Exception as ex:       # resident camera rays are a by-product of the frame path; never fatal for the line
        rays["error"] = str(ex)

    # ---- multi-GPU self-check (untimed): the sharded + all-gathered octree must equal the one built
    # from a full local irradiance pass, bit for bit (root p / sumArea / Et) on every rank
    mg_check = None
    if world > 1:
        import torch
        _, root_sharded, _ = dev.build_octree()
        dev.irradiance_resident(0, n_pts, n_tasks, scene.mix)
        dev.set_point_count(n_pts)
        _, root_full, _ = dev.build_octree()
        same = bool(np.array_equal(root_sharded.view(np.uint32), root_full.view(np.uint32)))
        t = torch.tensor([1 if same else 0]); dist.all_reduce(t, op=dist.ReduceOp.MIN)
        mg_check = "octree root bit-exact vs single-GPU pass on all ranks" if int(t.item()) == 1 else "MISMATCH"
        # ... and the film: the N films reduced to rank 0 against rank 0 rendering the whole frame alone (same octree)
        dev.frame_film_clear()
        dev.frame_render(r_first, r_count, scene.maxerror, scene.mix)
        comm.reduce_film(0)
        dev.synchronize()
        if rank == 0:
            film_n = dev.frame_film().copy()
            n_all = dev.frame_setup(scene.frame_desc(n_cores=1, rank=0, world=1))
            dev.frame_samples()
            dev.frame_film_clear()
            dev.frame_render(0, n_all, scene.maxerror, scene.mix)
            film_1 = dev.frame_film()
            err = float(np.abs(film_n - film_1).max() / np.abs(film_1).max())
            mg_check += "; film of %d ranks (ncclReduce) vs single-rank film: max |diff| / max = %.3g (%s)" % (world, err, "ok" if err < 1e-5 else "MISMATCH")
        comm.barrier()
    if rank != 0:
        return
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": DATA,
        "config": {"workload": WORKLOAD, "irradiance_points": n_pts, "camera_samples": n_rays, "irradiance_tasks": n_tasks,
                   "l2": "working set larger than L2 (surface points %.2f GB generated per step, octree + points > 4 GB)" % (n_pts * 44 / 1e9),
                   "camera_sample_pattern": "the reference's LDSampler stream, regenerated on the device every step (render tasks for 1 host core: %d)" % int(scene.frame_desc(1)[12:16].view(np.int32)[0]),
                   "li": "full Li: emission + subsurface term + direct lighting (2 lights x 4 samples, MIS) + film",
                   "material_tables": material},
        "stage_ms": stage_ms,
        "hot_path_seconds": ms_step * 1e-3,
        # shadow rays of the irradiance pass (+ the camera rays where their closest-hit stage is timed on its own, i.e. --no-prefetch)
        "rays_per_s": rays,
        "irradiance_points_per_s": n_pts / ((stage_ms["scramble"] + stage_ms["irradiance"]) * 1e-3),
        "mo_points_per_s": n_rays / (stage_ms["mo"] * 1e-3),
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d_bytes), "d2h_bytes_per_step": int(d2h_bytes),
                "ms_per_step": ms_e2e / args.steps, "render_seconds": ms_e2e / args.steps * 1e-3,
                "what": "scene upload from host + Preprocess + camera pass + RGB image to host"},
        "gpu_launches": int(launches),
        "clocks": summarise_clocks(lines),
        "roofline": roofline,
    }
    if mg_check:
        out["multi_gpu_check"] = mg_check
    if world == 1 and not args.no_cpu_baseline:
        drv = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
        if os.path.exists(drv):
            cores = os.cpu_count() or 1
            o = subprocess.run([drv, "bench", os.path.join(ROOT, "scenes", "skin_sample16.scene"), "--ncores", str(cores), "--steps", "2", "--warmup", "0"],
                               stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
            st = [l.split() for l in o.splitlines() if l.startswith("step ")]
            if st:
                u = sum(int(s[3]) + int(s[5]) for s in st); sec = sum(float(s[7]) + float(s[9]) for s in st)
                out["cpu_baseline"] = {"value": u / sec, "unit": UNIT, "cores": cores, "kind": "reference",
                                       "sample": "skin.scene at 1/16 scale (%s points, %s camera samples), reference Preprocess + its SamplerRendererTask loop with the full Li, 2 steps, %.1f s" % (st[0][3], st[0][5], sec)}
                try:     # one whole frame of the full scene through the reference, measured once on a GPU box's host (committed)
                    out["cpu_baseline"]["full_scale_step"] = json.load(open(os.path.join(ROOT, "profiles", "r2_reference_full_step.json")))
                except Exception:
                    pass
    emit(out)
    if comm:
        comm.close()
    dev.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

#!/usr/bin/env python3
"""BASELINE.json configs[4] at full size on the B200: 256 thin-epidermis parameter sets x 30 wavelengths =
7 680 two-layer diffusion profiles at desiredLength 512 (1024^2 grids, 2048^2 transforms, 119 766-entry tables), generated
by skc_profile_batch in chunks of `--chunk` thicknesses (a chunk's tables come back to the host: 345 MB per 8 thicknesses).
Device time = the library's "profile" stage timer; algorithmic bytes per profile as SURVEY.md 8d: three complex-double 2048^2
transforms, each two passes over 64 MB read + written, plus the layer grids and the tables.  With --reference N the real
reference's MPC_ComputeDiffusionProfile + resample (oracle/_ref/ref_driver mpc) is timed on N of the same parameter sets."""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def c5_parameter_sets(pkg, n_thick=256, desired_length=512):
    """Epidermis thickness swept over [0.02, 2] mean free paths (all thin: the lerp branch), fixed dermis 2.0, eta 1.4/1.4,
    mua/musp = the 30-bin spectra of SkinCoefficients(0.15, 1, 0.002, 0.3) (mirrors renderers/batchmcprofile.cpp:52-56)."""
    sp = pkg.skin_spectra(0.15, 1.0, 0.002, 0.3)
    f = np.float32
    specs, lps, opts = [], [], []
    for k in range(n_thick):
        for w in range(30):
            mfp0 = f(1) / f(sp["mua"][0][w] + sp["musp"][0][w])
            t = f(0.02 + k * (2.0 - 0.02) / max(n_thick - 1, 1)) * mfp0
            layers = [(1.4, t, sp["mua"][0][w], sp["musp"][0][w], 0.0), (1.4, 2.0, sp["mua"][1][w], sp["musp"][1][w], 0.0)]
            mfp = f((f(1) / f(layers[0][2] + layers[0][3]) + f(1) / f(layers[1][2] + layers[1][3])) / f(2))
            specs += layers; lps.append(2)
            opts.append((f(f(12) * mfp) / f(desired_length), desired_length, 1))
    return specs, lps, opts


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--thicknesses", type=int, default=256)
    ap.add_argument("--chunk", type=int, default=8)
    ap.add_argument("--length", type=int, default=512)
    ap.add_argument("--reference", type=int, default=0, help="time the real reference on this many of the sets (CPU, one core each)")
    args = ap.parse_args()
    # one JSON document on stdout: NCCL prints its version line there, so descriptor 1 is pointed at stderr for the run
    sys.stdout.flush()
    out_stream = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    import __graft_entry__ as ge
    pkg = ge.load_package()
    # several GPUs (first row): the parameter sets are independent, so the chunks are dealt round-robin over the
    # ranks of a torchrun launch and nothing is exchanged but the timing (max over ranks) and the checksum
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = pkg.SkinDevice(local)
    specs, lps, opts = c5_parameter_sets(pkg, args.thicknesses, args.length)
    n = len(lps)
    ms, entries, wall0 = 0.0, 0, time.time()
    checksum = 0.0
    for ci, c0 in enumerate(range(0, args.thicknesses, args.chunk)):
        if ci % world != rank:
            continue
        a, b = 30 * c0, 30 * min(c0 + args.chunk, args.thicknesses)
        d2, R, T, tot = dev.profile_batch(specs[2 * a:2 * b], lps[a:b], opts[a:b], resample=True)
        ms += dev.stage_ms("profile")
        entries = R.shape[1]
        checksum += float(tot.sum())
        assert np.all(np.isfinite(R)) and np.all(np.isfinite(T))
    if world > 1:
        import torch
        t = torch.tensor([ms], device="cuda", dtype=torch.float64); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t.item())
        c = torch.tensor([checksum, float(entries)], device="cuda", dtype=torch.float64); dist.all_reduce(c, op=dist.ReduceOp.SUM)
        checksum, entries = float(c[0].item()), int(round(float(c[1].item()) / world))
        dist.destroy_process_group()
        if rank != 0:
            dev.close()
            return
    grid = 2 * (2 * args.length)                      # transform size per side (zero-padded 2 x desiredLength x 2)
    fft_bytes = 3 * 2 * 2 * grid * grid * 16          # 3 transforms x 2 passes x (read + write) x complex double
    layer_bytes = 2 * 2 * (grid // 2) ** 2 * 8         # two layers, R and T grids, fp64
    bytes_per_profile = fft_bytes + layer_bytes + 3 * entries * 4
    out = {"n_gpus": world, "profiles": n, "desiredLength": args.length, "table_entries": entries, "device_ms": ms, "ms_per_profile": ms / n,
           "profiles_per_s": n / (ms * 1e-3), "algorithmic_GB_per_profile": bytes_per_profile / 1e9,
           "algorithmic_GBps": bytes_per_profile * n / (ms * 1e-3) / 1e9, "wall_s_incl_host_copies": time.time() - wall0,
           "sum_of_totals": checksum}
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        out["frac_of_hbm_peak"] = out["algorithmic_GBps"] / peak
    except Exception:
        pass
    drv = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
    if args.reference and os.path.exists(drv):
        with tempfile.TemporaryDirectory() as d:
            inp, outp = os.path.join(d, "in.txt"), os.path.join(d, "out.bin")
            pick = np.linspace(0, n - 1, args.reference).astype(int)
            with open(inp, "w") as f:
                for s in pick:
                    l0, l1 = specs[2 * s], specs[2 * s + 1]
                    f.write("2 %d 1 1 %r %r %r %r %r %r %r %r %r\n" % (args.length, float(opts[s][0]), l0[0], float(l0[1]), float(l0[2]), float(l0[3]),
                                                                      l1[0], float(l1[1]), float(l1[2]), float(l1[3])))
            t0 = time.time()
            subprocess.run([drv, "mpc", inp, outp], check=True, stdout=subprocess.DEVNULL)
            sec = time.time() - t0
            # the same sets again on the device, table against table (north_star: 1e-5 of each entry)
            sel_specs = [specs[2 * s + j] for s in pick for j in (0, 1)]
            d2, R, T, tot = dev.profile_batch(sel_specs, [2] * len(pick), [opts[s] for s in pick], resample=True)
            raw = np.fromfile(outp, np.uint8)
            worst, off, d2_equal = 0.0, 0, True
            for i in range(len(pick)):
                ln = int(raw[off:off + 4].view(np.uint32)[0]); tt = raw[off + 4:off + 12].view(np.float32); off += 12
                body = raw[off:off + 12 * ln].view(np.float32).reshape(3, ln); off += 12 * ln
                assert ln == R.shape[1]
                d2_equal = d2_equal and bool(np.array_equal(body[0], d2[i]))
                floor = max(float(np.abs(body[1]).max()), float(np.abs(body[2]).max())) * 1e-30      # 1e-30 of the profile's maximum
                for got, ref in ((R[i], body[1]), (T[i], body[2])):
                    ref64 = ref.astype(np.float64)
                    worst = max(worst, float((np.abs(got - ref64) / np.maximum(np.abs(ref64), floor)).max()))
                worst = max(worst, float(np.abs(tot[i] - tt).max() / np.abs(tt).max()))
        out["reference_cpu"] = {"profiles": int(args.reference), "seconds": sec, "s_per_profile_one_core": sec / args.reference,
                                "speedup_vs_one_core": (sec / args.reference) / (ms / n * 1e-3),
                                "tables_max_rel_err_vs_device": worst, "d2_grids_bit_equal": d2_equal, "within_1e-5": bool(worst < 1e-5)}
    out_stream.write(json.dumps(out, indent=1) + "\n")
    out_stream.flush()
    dev.close()


if __name__ == "__main__":
    main()

"""ctypes wrapper of oracle/_build/libskinoracle.so (the C restatement).  TEST INFRASTRUCTURE:
imported only from tests/, __graft_entry__.smoke() and checker legs."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(ROOT, "oracle", "_build", "libskinoracle.so")
        if not os.path.exists(path):
            subprocess.run(["make", "-C", os.path.join(ROOT, "oracle")], check=True, stdout=subprocess.DEVNULL)
        _lib = C.CDLL(path)
        _lib.orc_octree_build.restype = C.c_void_p
        _lib.orc_octree_info.restype = C.c_uint64
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f(a):
    return None if a is None else np.ascontiguousarray(a, np.float32)


def _scene_args(g):
    return (np.ascontiguousarray(g.bvh_nodes), np.ascontiguousarray(g.bvh_prims, np.int32), _f(g.mesh_P), _f(g.mesh_N), _f(g.mesh_uv),
            np.ascontiguousarray(g.mesh_idx, np.int32), _f(g.lights), _f(g.mesh_o2w))


def trace_closest(g, rays, counters=False):
    nodes, prims, P, N, UV, idx, sph, o2w = _scene_args(g)
    rays = _f(rays).reshape(-1, 8)
    from golden_util import DT_HIT
    hits = np.zeros(len(rays), DT_HIT)
    cnt = np.zeros(2, np.uint64) if counters else None
    flip = int(g.meta["mesh0_reverse"]) ^ int(g.meta["mesh0_swaps"])
    lib().orc_trace_closest(_p(nodes), _p(prims), _p(P), _p(N), _p(UV), _p(idx), _p(sph), _p(o2w), C.c_int(flip), _p(rays), C.c_uint64(len(rays)), _p(hits), _p(cnt))
    return (hits, cnt) if counters else hits


def trace_any(g, rays):
    nodes, prims, P, N, UV, idx, sph, o2w = _scene_args(g)
    rays = _f(rays).reshape(-1, 8)
    out = np.zeros(len(rays), np.uint8)
    lib().orc_trace_any(_p(nodes), _p(prims), _p(P), _p(idx), _p(sph), _p(rays), C.c_uint64(len(rays)), _p(out), None)
    return out


def irradiance(g, points=None, n_tasks=None, mix=None, rho=None, counters=False):
    from golden_util import DT_IP
    nodes, prims, P, N, UV, idx, sph, o2w = _scene_args(g)
    pts = np.ascontiguousarray(g.surface_points if points is None else points)
    out = np.zeros(len(pts), DT_IP)
    rho = _f(g.rho_hd[:1025] if rho is None else rho)
    cnt = np.zeros(3, np.uint64) if counters else None
    lib().orc_irradiance(_p(nodes), _p(prims), _p(P), _p(idx), _p(sph), C.c_uint32(len(g.lights)), _p(pts), C.c_uint64(len(pts)),
                         C.c_uint32(g.n_tasks if n_tasks is None else n_tasks), C.c_float(g.mix if mix is None else mix), _p(rho), None, _p(out), _p(cnt))
    return (out, cnt) if counters else out


def scrambles(n_pts, n_tasks, n_lights):
    out = np.zeros((n_pts, n_lights, 3), np.uint32)
    lib().orc_scrambles(C.c_uint64(n_pts), C.c_uint32(n_tasks), C.c_uint32(n_lights), _p(out))
    return out


class Octree:
    def __init__(self, ips, cie_y):
        self.ips = np.ascontiguousarray(ips)          # must outlive the tree (leaves point into it)
        self.cie = _f(cie_y)
        self.h = C.c_void_p(lib().orc_octree_build(_p(self.ips), C.c_uint64(len(self.ips)), _p(self.cie)))

    def info(self):
        root = np.zeros(34, np.float32); b = np.zeros(6, np.float32)
        n = lib().orc_octree_info(self.h, _p(root), _p(b))
        return int(n), root, b

    def mo(self, p, table, rcp, max_error, counters=False):
        p = _f(p).reshape(-1, 3); table = _f(table); rcp = _f(rcp)
        out = np.zeros((len(p), 30), np.float32)
        cnt = np.zeros(3, np.uint64) if counters else None
        lib().orc_mo(self.h, _p(p), C.c_uint64(len(p)), _p(table), _p(rcp), C.c_uint32(table.shape[1]), C.c_float(max_error), _p(out), _p(cnt))
        return (out, cnt) if counters else out

    def close(self):
        if self.h:
            lib().orc_octree_free(self.h); self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def li_sss(hits, mo, rho, mix):
    hits = np.ascontiguousarray(hits); mo = _f(mo); rho = _f(rho)
    out = np.zeros((len(hits), 30), np.float32)
    lib().orc_li_sss(_p(hits), C.c_uint64(len(hits)), _p(mo), _p(rho), None, C.c_float(mix), _p(out))
    return out


DT_ORC_LAYER = np.dtype([("ior", "<f4"), ("thickness", "<f4"), ("mua", "<f4"), ("musp", "<f4"), ("g_HG", "<f4")])
DT_ORC_MPC_OPTIONS = np.dtype([("desiredStepSize", "<f4"), ("desiredLength", "<u4"), ("lerpOnThinSlab", "<i4")])


def mpc_profile(specs, opt, resample=True):
    """oracle/mpc_oracle.c orc_mpc_profile: (d2, R, T, totals) of one layer stack."""
    L = lib()
    L.orc_mpc_profile.restype = C.c_uint32
    specs = np.ascontiguousarray(np.array([tuple(s) for s in specs], DT_ORC_LAYER))
    o = np.ascontiguousarray(np.array([tuple(opt)], DT_ORC_MPC_OPTIONS))
    n = L.orc_mpc_profile(len(specs), _p(specs), _p(o), int(resample), None, None, None, None)
    d2 = np.empty(n, np.float32); R = np.empty(n, np.float32); T = np.empty(n, np.float32); tot = np.empty(2, np.float32)
    L.orc_mpc_profile(len(specs), _p(specs), _p(o), int(resample), _p(d2), _p(R), _p(T), _p(tot))
    return d2, R, T, tot

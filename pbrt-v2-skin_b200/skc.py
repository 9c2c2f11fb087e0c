"""ctypes binding of include/skin_cuda.h (libskincuda.so).  Plumbing only -- see __init__."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

# record layouts of include/skin_cuda.h
DT_BVH_NODE = np.dtype([("pmin", "<f4", 3), ("pmax", "<f4", 3), ("offset", "<u4"), ("nPrimitives", "u1"), ("axis", "u1"), ("pad", "u1", 2)])
DT_PRIM = np.dtype([("type", "<i4"), ("mesh", "<i4"), ("index", "<i4"), ("primitiveId", "<i4")])
DT_LIGHT = np.dtype([("o2w", "<f4", 16), ("w2o", "<f4", 16), ("radius", "<f4"), ("zmin", "<f4"), ("zmax", "<f4"), ("phiMax", "<f4"),
                     ("nSamples", "<f4"), ("reverseOrientation", "<f4"), ("thetaMin", "<f4"), ("thetaMax", "<f4"), ("Lemit", "<f4", 30)])
DT_RAY = np.dtype([("o", "<f4", 3), ("d", "<f4", 3), ("mint", "<f4"), ("maxt", "<f4")])
DT_HIT = np.dtype([("hit", "<i4"), ("prim", "<i4"), ("t", "<f4"), ("rayEpsilon", "<f4"), ("p", "<f4", 3), ("n", "<f4", 3), ("wo", "<f4", 3), ("hasBssrdf", "<i4")])
DT_SURFACE_POINT = np.dtype([("p", "<f4", 3), ("n", "<f4", 3), ("u", "<f4"), ("v", "<f4"), ("materialId", "<u4"), ("area", "<f4"), ("rayEpsilon", "<f4")])
DT_IRR_POINT = np.dtype([("p", "<f4", 3), ("n", "<f4", 3), ("E", "<f4", 30), ("area", "<f4"), ("rayEpsilon", "<f4")])
DT_LAYER = np.dtype([("ior", "<f4"), ("thickness", "<f4"), ("mua", "<f4"), ("musp", "<f4"), ("g_HG", "<f4")])
DT_MPC_OPTIONS = np.dtype([("desiredStepSize", "<f4"), ("desiredLength", "<u4"), ("lerpOnThinSlab", "<i4")])
assert DT_BVH_NODE.itemsize == 32 and DT_LIGHT.itemsize == 280 and DT_RAY.itemsize == 32 and DT_HIT.itemsize == 56
assert DT_SURFACE_POINT.itemsize == 44 and DT_IRR_POINT.itemsize == 152

STAGES = {"scramble": 0, "irradiance": 1, "octree": 2, "closest": 3, "mo": 4, "any": 5, "profile": 6, "rho": 7, "tessellate": 8,
          "samples": 9, "direct": 10, "film": 11}

# every symbol include/skin_cuda.h declares
SYMBOLS = ["skc_create", "skc_destroy", "skc_last_error", "skc_synchronize", "skc_launch_count", "skc_sm_count", "skc_stage_ms",
           "skc_upload_bvh", "skc_trace_any", "skc_trace_closest", "skc_upload_profile", "skc_upload_lights",
           "skc_irradiance", "skc_trace_resident", "skc_set_irradiance_points", "skc_irradiance_dev", "skc_build_octree", "skc_mo",
           "skc_li_sss", "skc_camera_pass", "skc_profile_batch", "skc_profile_resample", "skc_set_surface_points", "skc_irradiance_resident",
           "skc_set_camera_rays", "skc_camera_pass_resident", "skc_alloc_pinned", "skc_free_pinned",
           "skc_timer_begin", "skc_timer_end_ms", "skc_set_point_count", "skc_rho_table",
           "skc_build_octree_topology", "skc_build_octree_finalize", "skc_preprocess", "skc_preprocess_resident",
           "skc_tessellate_resident", "skc_tessellate_host", "skc_tessellate_range", "skc_get_surface_points",
           "skc_frame_setup", "skc_frame_prefetch", "skc_frame_samples", "skc_frame_get_samples", "skc_frame_sample_floats", "skc_frame_render",
           "skc_frame_film_clear", "skc_frame_film_dev", "skc_frame_film_get", "skc_frame_film_rgb", "skc_frame_get_order",
           "skc_comm_unique_id", "skc_comm_create", "skc_comm_create_local", "skc_comm_destroy", "skc_comm_rank", "skc_comm_world",
           "skc_comm_allgather_positions", "skc_comm_allgather_rows_begin", "skc_comm_allgather_rows_wait", "skc_comm_reduce_film",
           "skc_comm_max_f32", "skc_comm_barrier", "skc_comm_allgather_host", "skc_mo_static", "skc_test_sincosf", "skc_test_libmf", "skc_get_irradiance_rows", "skc_set_irradiance_rows", "skc_mo_table_stats", "skc_mo_group_histogram", "skc_set_albedo_factors", "skc_upload_envmap", "skc_upload_alpha", "skc_upload_albedo_texture", "skc_upload_bump_texture", "skc_poisson_resident", "skc_moc_build", "skc_moc_info", "skc_moc_eval", "skc_moc_free"]


class SkcError(RuntimeError):
    pass


def lib_path():
    return os.path.join(_HERE, "libskincuda.so")


_lib = None


def load_library():
    """dlopen libskincuda.so and declare prototypes.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise SkcError("%s not built: run `python -c 'import __graft_entry__ as g; g.build()'` (nvcc, sm_100a). "
                       "There is no CPU fallback." % path)
    lib = C.CDLL(path)
    vp, u64, u32, i32, f32 = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int, C.c_float
    lib.skc_create.restype = vp; lib.skc_create.argtypes = [i32]
    lib.skc_destroy.restype = None; lib.skc_destroy.argtypes = [vp]
    lib.skc_last_error.restype = C.c_char_p; lib.skc_last_error.argtypes = [vp]
    lib.skc_synchronize.restype = i32; lib.skc_synchronize.argtypes = [vp]
    lib.skc_launch_count.restype = u64; lib.skc_launch_count.argtypes = [vp]
    lib.skc_sm_count.restype = i32; lib.skc_sm_count.argtypes = [vp]
    lib.skc_stage_ms.restype = f32; lib.skc_stage_ms.argtypes = [vp, i32]
    lib.skc_upload_bvh.restype = i32
    lib.skc_upload_bvh.argtypes = [vp, vp, u32, vp, u32, vp, vp, vp, u32, vp, u32, vp, i32, i32, vp, u32]
    lib.skc_trace_any.restype = i32; lib.skc_trace_any.argtypes = [vp, vp, u64, vp]
    lib.skc_trace_closest.restype = i32; lib.skc_trace_closest.argtypes = [vp, vp, u64, vp]
    lib.skc_upload_profile.restype = i32; lib.skc_upload_profile.argtypes = [vp, vp, vp, u32, vp]
    lib.skc_upload_lights.restype = i32; lib.skc_upload_lights.argtypes = [vp, vp, u32]
    lib.skc_irradiance.restype = i32; lib.skc_irradiance.argtypes = [vp, vp, u64, u64, u64, u32, f32, vp, vp]
    lib.skc_trace_resident.restype = i32; lib.skc_trace_resident.argtypes = [vp, u64, u64, i32, vp, vp]
    lib.skc_set_irradiance_points.restype = i32; lib.skc_set_irradiance_points.argtypes = [vp, vp, u64]
    lib.skc_irradiance_dev.restype = i32; lib.skc_irradiance_dev.argtypes = [vp, u64, C.POINTER(vp), C.POINTER(vp)]
    lib.skc_build_octree.restype = i32; lib.skc_build_octree.argtypes = [vp, C.POINTER(u32), vp, vp]
    lib.skc_tessellate_resident.restype = i32; lib.skc_tessellate_resident.argtypes = [vp, f32, i32, u32, C.POINTER(u64)]
    lib.skc_get_surface_points.restype = i32; lib.skc_get_surface_points.argtypes = [vp, u64, u64, vp]
    lib.skc_preprocess.restype = i32; lib.skc_preprocess.argtypes = [vp, vp, u64, u64, u64, u32, f32]
    lib.skc_preprocess_resident.restype = i32; lib.skc_preprocess_resident.argtypes = [vp, u64, u64, u32, f32]
    lib.skc_build_octree_topology.restype = i32; lib.skc_build_octree_topology.argtypes = [vp]
    lib.skc_build_octree_finalize.restype = i32; lib.skc_build_octree_finalize.argtypes = [vp, C.POINTER(u32), vp, vp]
    lib.skc_mo.restype = i32; lib.skc_mo.argtypes = [vp, vp, u64, f32, vp]
    lib.skc_li_sss.restype = i32; lib.skc_li_sss.argtypes = [vp, vp, u64, f32, f32, vp, vp, vp]
    lib.skc_camera_pass.restype = i32; lib.skc_camera_pass.argtypes = [vp, vp, u64, f32, f32, vp, vp]
    lib.skc_rho_table.restype = i32
    lib.skc_rho_table.argtypes = [vp, C.c_float, C.c_float, C.c_float, i32, vp]
    lib.skc_profile_batch.restype = i32
    lib.skc_profile_batch.argtypes = [vp, vp, vp, vp, u32, i32, C.POINTER(u32), vp, vp, vp, vp]
    lib.skc_set_surface_points.restype = i32; lib.skc_set_surface_points.argtypes = [vp, vp, u64]
    lib.skc_irradiance_resident.restype = i32; lib.skc_irradiance_resident.argtypes = [vp, u64, u64, u32, f32, vp]
    lib.skc_set_camera_rays.restype = i32; lib.skc_set_camera_rays.argtypes = [vp, vp, u64]
    lib.skc_camera_pass_resident.restype = i32; lib.skc_camera_pass_resident.argtypes = [vp, u64, u64, f32, f32, vp, vp]
    lib.skc_alloc_pinned.restype = vp; lib.skc_alloc_pinned.argtypes = [u64]
    lib.skc_free_pinned.restype = None; lib.skc_free_pinned.argtypes = [vp]
    lib.skc_timer_begin.restype = i32; lib.skc_timer_begin.argtypes = [vp]
    lib.skc_timer_end_ms.restype = f32; lib.skc_timer_end_ms.argtypes = [vp]
    lib.skc_set_point_count.restype = i32; lib.skc_set_point_count.argtypes = [vp, u64]
    lib.skc_tessellate_range.restype = i32; lib.skc_tessellate_range.argtypes = [vp, f32, i32, u32, u64, u64, C.POINTER(u64)]
    lib.skc_frame_setup.restype = i32; lib.skc_frame_setup.argtypes = [vp, vp, C.POINTER(u64)]
    lib.skc_frame_samples.restype = i32; lib.skc_frame_samples.argtypes = [vp]
    lib.skc_frame_prefetch.restype = i32; lib.skc_frame_prefetch.argtypes = [vp]
    lib.skc_frame_get_samples.restype = i32; lib.skc_frame_get_samples.argtypes = [vp, u64, u64, vp, vp, vp]
    lib.skc_frame_sample_floats.restype = u32; lib.skc_frame_sample_floats.argtypes = [vp]
    lib.skc_frame_render.restype = i32; lib.skc_frame_render.argtypes = [vp, u64, u64, f32, f32, vp]
    lib.skc_frame_get_order.restype = i32; lib.skc_frame_get_order.argtypes = [vp, u64, u64, vp]
    lib.skc_frame_film_clear.restype = i32; lib.skc_frame_film_clear.argtypes = [vp]
    lib.skc_frame_film_dev.restype = vp; lib.skc_frame_film_dev.argtypes = [vp]
    lib.skc_frame_film_get.restype = i32; lib.skc_frame_film_get.argtypes = [vp, vp]
    lib.skc_frame_film_rgb.restype = i32; lib.skc_frame_film_rgb.argtypes = [vp, vp]
    lib.skc_get_irradiance_rows.restype = i32; lib.skc_get_irradiance_rows.argtypes = [vp, u64, u64, u64, vp, vp]
    lib.skc_set_irradiance_rows.restype = i32; lib.skc_set_irradiance_rows.argtypes = [vp, u64, u64, vp, vp]
    lib.skc_test_libmf.restype = i32; lib.skc_test_libmf.argtypes = [vp, i32, vp, u64, vp]
    lib.skc_test_sincosf.restype = i32; lib.skc_test_sincosf.argtypes = [vp, vp, u64, vp, vp]
    lib.skc_comm_unique_id.restype = i32; lib.skc_comm_unique_id.argtypes = [vp]
    lib.skc_comm_create.restype = vp; lib.skc_comm_create.argtypes = [vp, i32, i32, vp]
    lib.skc_comm_destroy.restype = None; lib.skc_comm_destroy.argtypes = [vp]
    lib.skc_comm_rank.restype = i32; lib.skc_comm_rank.argtypes = [vp]
    lib.skc_comm_world.restype = i32; lib.skc_comm_world.argtypes = [vp]
    for f in ("skc_comm_allgather_positions", "skc_comm_allgather_rows_begin"):
        getattr(lib, f).restype = i32; getattr(lib, f).argtypes = [vp, u64]
    lib.skc_comm_allgather_rows_wait.restype = i32; lib.skc_comm_allgather_rows_wait.argtypes = [vp]
    lib.skc_comm_reduce_film.restype = i32; lib.skc_comm_reduce_film.argtypes = [vp, i32]
    lib.skc_comm_max_f32.restype = i32; lib.skc_comm_max_f32.argtypes = [vp, C.POINTER(f32)]
    lib.skc_comm_barrier.restype = i32; lib.skc_comm_barrier.argtypes = [vp]
    lib.skc_comm_allgather_host.restype = i32; lib.skc_comm_allgather_host.argtypes = [vp, vp, u64]
    lib.skc_mo_static.restype = i32; lib.skc_mo_static.argtypes = [vp, vp, u64, f32, vp]
    lib.skc_upload_albedo_texture.restype = i32; lib.skc_upload_albedo_texture.argtypes = [vp, u32, vp, vp, vp, vp, f32, f32, f32, f32]
    lib.skc_poisson_resident.restype = i32; lib.skc_poisson_resident.argtypes = [vp, vp, f32, u32, i32, vp, vp]
    lib.skc_upload_bump_texture.restype = i32; lib.skc_upload_bump_texture.argtypes = [vp, u32, vp, vp, vp, vp, f32, f32, f32, f32]
    lib.skc_upload_alpha.restype = i32; lib.skc_upload_alpha.argtypes = [vp, u32, u32, vp, f32, f32, f32, f32]
    lib.skc_upload_envmap.restype = i32; lib.skc_upload_envmap.argtypes = [vp, u32, u32, vp, vp, vp, vp, vp, vp, f32]
    lib.skc_set_albedo_factors.restype = i32; lib.skc_set_albedo_factors.argtypes = [vp, vp, vp]
    lib.skc_mo_group_histogram.restype = i32; lib.skc_mo_group_histogram.argtypes = [vp, vp]
    lib.skc_mo_table_stats.restype = i32; lib.skc_mo_table_stats.argtypes = [vp, vp]
    lib.skc_moc_build.restype = vp; lib.skc_moc_build.argtypes = [vp, vp, u32, f32, i32, i32, u32]
    lib.skc_moc_info.restype = C.c_double; lib.skc_moc_info.argtypes = [vp, vp, C.POINTER(u32), C.POINTER(u32)]
    lib.skc_moc_eval.restype = None; lib.skc_moc_eval.argtypes = [vp, i32, vp, u64, vp, vp]
    lib.skc_moc_free.restype = None; lib.skc_moc_free.argtypes = [vp]
    _lib = lib
    return lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _c(a, dtype):
    return None if a is None else np.ascontiguousarray(a, dtype=dtype)


class SkinDevice:
    """One skc_ctx.  Methods mirror the C entry points one to one, with numpy arrays."""

    def __init__(self, device=-1):
        self.lib = load_library()
        self.ctx = self.lib.skc_create(device)
        if not self.ctx:
            raise SkcError(self.lib.skc_last_error(None).decode())

    def close(self):
        if self.ctx:
            for p in getattr(self, "_pinned", []):
                self.lib.skc_free_pinned(p)
            self._pinned = []
            self.lib.skc_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise SkcError(self.lib.skc_last_error(self.ctx).decode())

    def synchronize(self):
        self._ck(self.lib.skc_synchronize(self.ctx))

    def timer_begin(self):
        self._ck(self.lib.skc_timer_begin(self.ctx))

    def timer_end_ms(self):
        return float(self.lib.skc_timer_end_ms(self.ctx))

    def set_point_count(self, n):
        self._ck(self.lib.skc_set_point_count(self.ctx, n))

    def launch_count(self):
        return int(self.lib.skc_launch_count(self.ctx))

    def sm_count(self):
        return int(self.lib.skc_sm_count(self.ctx))

    def stage_ms(self, name):
        return float(self.lib.skc_stage_ms(self.ctx, STAGES[name]))

    # --- BVHAccel
    def upload_bvh(self, nodes, prims, P, N, uv, indices, mesh_o2w=None, reverse=0, swaps=0, spheres=None):
        nodes = _c(nodes.view(DT_BVH_NODE) if nodes.dtype != DT_BVH_NODE else nodes, DT_BVH_NODE)
        prims = _c(prims, np.int32).reshape(-1, 4)
        P = _c(P, np.float32).reshape(-1, 3); N = _c(N, np.float32); uv = _c(uv, np.float32)
        indices = _c(indices, np.int32).reshape(-1, 3)
        mesh_o2w = _c(mesh_o2w, np.float32)
        spheres = _c(spheres, np.float32)
        ns = 0 if spheres is None else spheres.size // 70
        self._keep = (nodes, prims, P, N, uv, indices, mesh_o2w, spheres)
        self._ck(self.lib.skc_upload_bvh(self.ctx, _ptr(nodes), nodes.size, _ptr(prims), len(prims), _ptr(P), _ptr(N), _ptr(uv),
                                         len(P), _ptr(indices), len(indices), _ptr(mesh_o2w), int(reverse), int(swaps), _ptr(spheres), ns))

    def trace_any(self, rays):
        rays = _c(rays, np.float32).reshape(-1, 8)
        out = np.empty(len(rays), np.uint8)
        self._ck(self.lib.skc_trace_any(self.ctx, _ptr(rays), len(rays), _ptr(out)))
        return out

    def trace_closest(self, rays):
        rays = _c(rays, np.float32).reshape(-1, 8)
        out = np.empty(len(rays), DT_HIT)
        self._ck(self.lib.skc_trace_closest(self.ctx, _ptr(rays), len(rays), _ptr(out)))
        return out

    # --- tables, lights
    def upload_profile(self, table=None, rcp=None, rho_hd=None):
        table = _c(table, np.float32); rcp = _c(rcp, np.float32); rho_hd = _c(rho_hd, np.float32)
        ln = 0 if table is None else table.shape[-1]
        self._ck(self.lib.skc_upload_profile(self.ctx, _ptr(table), _ptr(rcp), ln, _ptr(rho_hd)))

    def upload_lights(self, lights):
        lights = _c(lights, np.float32).reshape(-1, 70)
        self._ck(self.lib.skc_upload_lights(self.ctx, _ptr(lights), len(lights)))

    # --- Preprocess
    def irradiance(self, points, n_tasks, mix=0.5, first=0, count=None, albedo=None, want_out=True):
        pts = points if points.dtype == DT_SURFACE_POINT else points.view(DT_SURFACE_POINT)
        pts = np.ascontiguousarray(pts)
        n = len(pts)
        count = n - first if count is None else count
        out = np.empty(count, DT_IRR_POINT) if want_out else None
        albedo = _c(albedo, np.float32)
        self._ck(self.lib.skc_irradiance(self.ctx, _ptr(pts), n, first, count, n_tasks, mix, _ptr(albedo), _ptr(out)))
        return out

    def set_irradiance_points(self, pts):
        pts = np.ascontiguousarray(pts if pts.dtype == DT_IRR_POINT else pts.view(DT_IRR_POINT).reshape(-1))
        self._ck(self.lib.skc_set_irradiance_points(self.ctx, _ptr(pts), len(pts)))

    def irradiance_rows(self, first, count, stride=1):
        """({p, area} [count][4], E [count][30]) of device irradiance points first, first+stride, ..."""
        pa = np.empty((count, 4), np.float32); E = np.empty((count, 30), np.float32)
        self._ck(self.lib.skc_get_irradiance_rows(self.ctx, first, count, stride, _ptr(pa), _ptr(E)))
        return pa, E

    def test_libmf(self, fn, x):
        """The device's restatement of libm's float function fn (0 sinf, 1 cosf, 2 logf, 3 expf, 4 atanf) on x."""
        x = _c(x, np.float32).reshape(-1)
        out = np.empty_like(x)
        self._ck(self.lib.skc_test_libmf(self.ctx, int(fn), _ptr(x), len(x), _ptr(out)))
        return out

    def test_sincosf(self, x):
        x = _c(x, np.float32).reshape(-1)
        s = np.empty_like(x); c = np.empty_like(x)
        self._ck(self.lib.skc_test_sincosf(self.ctx, _ptr(x), len(x), _ptr(s), _ptr(c)))
        return s, c

    def set_irradiance_rows(self, first, E, pos_area=None):
        E = _c(E, np.float32).reshape(-1, 30); pos_area = _c(pos_area, np.float32)
        self._ck(self.lib.skc_set_irradiance_rows(self.ctx, first, len(E), _ptr(pos_area), _ptr(E)))

    def irradiance_dev(self, n_total):
        a, b = C.c_void_p(), C.c_void_p()
        self._ck(self.lib.skc_irradiance_dev(self.ctx, n_total, C.byref(a), C.byref(b)))
        return a.value, b.value

    def build_octree(self):
        nn = C.c_uint32(0)
        root = np.zeros(34, np.float32); bounds = np.zeros(6, np.float32)
        self._ck(self.lib.skc_build_octree(self.ctx, C.byref(nn), _ptr(root), _ptr(bounds)))
        return int(nn.value), root, bounds

    def tessellate_resident(self, min_dist, incenter=False, material_id=0):
        """Surface points of the uploaded mesh, generated on the device into the resident set; returns the count."""
        n = C.c_uint64(0)
        self._ck(self.lib.skc_tessellate_resident(self.ctx, min_dist, int(incenter), material_id, C.byref(n)))
        return int(n.value)

    def tessellate_range(self, min_dist, first, count, incenter=False, material_id=0):
        """Surface points [first, first+count) of the tessellated set only (a rank's shard); returns the size of the whole set."""
        n = C.c_uint64(0)
        self._ck(self.lib.skc_tessellate_range(self.ctx, float(min_dist), int(incenter), int(material_id), int(first), int(count), C.byref(n)))
        return int(n.value)

    def poisson_resident(self, origin, min_sample_dist, material_id=0, max_fails=2000):
        """FindPoissonPointDistribution as the reference runs it on one core (skc_poisson_resident); returns (count, stats dict)."""
        o = np.ascontiguousarray(origin, np.float32); n = C.c_uint64(0); st = np.zeros(4, np.uint64)
        self._ck(self.lib.skc_poisson_resident(self.ctx, _ptr(o), float(min_sample_dist), int(material_id), int(max_fails), C.byref(n), _ptr(st)))
        return int(n.value), dict(points=int(st[0]), paths=int(st[1]), rays=int(st[2]), rounds=int(st[3]))

    def get_surface_points(self, first, count):
        out = np.empty(count, DT_SURFACE_POINT)
        self._ck(self.lib.skc_get_surface_points(self.ctx, first, count, _ptr(out)))
        return out

    def preprocess(self, points, n_tasks, mix=0.5, first=0, count=None):
        """Irradiance pass over [first, first+count) + octree topology of all points, overlapped (skc_preprocess)."""
        pts = np.ascontiguousarray(points if points.dtype == DT_SURFACE_POINT else points.view(DT_SURFACE_POINT))
        count = len(pts) - first if count is None else count
        self._ck(self.lib.skc_preprocess(self.ctx, _ptr(pts), len(pts), first, count, n_tasks, mix))

    def preprocess_resident(self, first, count, n_tasks, mix=0.5):
        self._ck(self.lib.skc_preprocess_resident(self.ctx, first, count, n_tasks, mix))

    def build_octree_topology(self):
        self._ck(self.lib.skc_build_octree_topology(self.ctx))

    def build_octree_finalize(self):
        nn = C.c_uint32(0)
        root = np.zeros(34, np.float32); bounds = np.zeros(6, np.float32)
        self._ck(self.lib.skc_build_octree_finalize(self.ctx, C.byref(nn), _ptr(root), _ptr(bounds)))
        return int(nn.value), root, bounds

    # --- Li
    def mo(self, p, max_error):
        p = _c(p, np.float32).reshape(-1, 3)
        out = np.empty((len(p), 30), np.float32)
        self._ck(self.lib.skc_mo(self.ctx, _ptr(p), len(p), max_error, _ptr(out)))
        return out

    MO_STATIC_MAX_HITS = 16384

    def mo_static(self, p, max_error):
        """skc_mo_static: Mo through the file-scope device scratch (no allocation inside the call); at most MO_STATIC_MAX_HITS points."""
        p = _c(p, np.float32).reshape(-1, 3)
        out = np.empty((len(p), 30), np.float32)
        self._ck(self.lib.skc_mo_static(self.ctx, _ptr(p), len(p), max_error, _ptr(out)))
        return out

    def mo_table_stats(self):
        """Compact profile table of the Mo kernel: dict(rows, smem_bytes, coverage, max_err, lookups, from_heads, from_global)."""
        st = np.zeros(8, np.float64)
        self._ck(self.lib.skc_mo_table_stats(self.ctx, _ptr(st)))
        return dict(rows=int(st[0]), smem_bytes=int(st[1]), coverage=float(st[2]), max_err=float(st[3]),
                    lookups=int(st[4]), from_heads=int(st[5]), from_global=int(st[6]), accept_groups=int(st[7]))

    def upload_albedo_texture(self, tex):
        """tex: dict with w, h (per level), rgb, lut, map = (su, sv, du, dv) (HostScene.albedo_texture()), or None to remove it."""
        if tex is None:
            return self._ck(self.lib.skc_upload_albedo_texture(self.ctx, 0, None, None, None, None, 1.0, 1.0, 0.0, 0.0))
        w = np.ascontiguousarray(tex["w"], np.int32); h = np.ascontiguousarray(tex["h"], np.int32)
        rgb = np.ascontiguousarray(tex["rgb"], np.float32); lut = np.ascontiguousarray(tex["lut"], np.float32)
        self._ck(self.lib.skc_upload_albedo_texture(self.ctx, len(w), _ptr(w), _ptr(h), _ptr(rgb), _ptr(lut), *[float(x) for x in tex["map"]]))

    def upload_bump_texture(self, tex):
        """tex: dict with w, h (per level), texels (one displacement each), lut, map = (su, sv, du, dv) (HostScene.bump_texture()), or None to remove it."""
        if tex is None:
            return self._ck(self.lib.skc_upload_bump_texture(self.ctx, 0, None, None, None, None, 1.0, 1.0, 0.0, 0.0))
        w = np.ascontiguousarray(tex["w"], np.int32); h = np.ascontiguousarray(tex["h"], np.int32)
        t = np.ascontiguousarray(tex["texels"], np.float32); lut = np.ascontiguousarray(tex["lut"], np.float32)
        self._ck(self.lib.skc_upload_bump_texture(self.ctx, len(w), _ptr(w), _ptr(h), _ptr(t), _ptr(lut), *[float(x) for x in tex["map"]]))

    def upload_alpha(self, alpha):
        """alpha: dict with w, h, texels, map = (su, sv, du, dv) (HostScene.alpha_texture()), or None to remove it."""
        if alpha is None:
            return self._ck(self.lib.skc_upload_alpha(self.ctx, 0, 0, None, 1.0, 1.0, 0.0, 0.0))
        t = np.ascontiguousarray(alpha["texels"], np.float32)
        self._ck(self.lib.skc_upload_alpha(self.ctx, int(alpha["w"]), int(alpha["h"]), _ptr(t), *[float(x) for x in alpha["map"]]))

    def upload_envmap(self, env):
        """env: dict with w, h, rgb, condFunc, condCdf, condInt, margFunc, margCdf, margInt (HostScene.envmap()), or None to remove it."""
        if env is None:
            return self._ck(self.lib.skc_upload_envmap(self.ctx, 0, 0, None, None, None, None, None, None, 0.0))
        a = [np.ascontiguousarray(env[k], np.float32) for k in ("rgb", "condFunc", "condCdf", "condInt", "margFunc", "margCdf")]
        self._ck(self.lib.skc_upload_envmap(self.ctx, int(env["w"]), int(env["h"]), *[_ptr(x) for x in a], float(env["margInt"])))

    def set_albedo_factors(self, pow_mix=None, pow_one_minus_mix=None):
        a = None if pow_mix is None else np.ascontiguousarray(pow_mix, np.float32); b = None if pow_one_minus_mix is None else np.ascontiguousarray(pow_one_minus_mix, np.float32)
        self._ck(self.lib.skc_set_albedo_factors(self.ctx, None if a is None else _ptr(a), None if b is None else _ptr(b)))

    def mo_group_histogram(self):
        h = np.zeros(8, np.uint64)
        self._ck(self.lib.skc_mo_group_histogram(self.ctx, _ptr(h)))
        return [int(x) for x in h]

    def li_sss(self, hits, max_error, mix=0.5, albedo=None):
        hits = np.ascontiguousarray(hits if hits.dtype == DT_HIT else hits.view(DT_HIT).reshape(-1))
        mo = np.empty((len(hits), 30), np.float32); L = np.empty((len(hits), 30), np.float32)
        albedo = _c(albedo, np.float32)
        self._ck(self.lib.skc_li_sss(self.ctx, _ptr(hits), len(hits), max_error, mix, _ptr(albedo), _ptr(mo), _ptr(L)))
        return mo, L

    def camera_pass(self, rays, max_error, mix=0.5, want_hits=False, out=None):
        rays = _c(rays, np.float32).reshape(-1, 8)
        L = np.empty((len(rays), 30), np.float32) if out is None else out
        hits = np.empty(len(rays), DT_HIT) if want_hits else None
        self._ck(self.lib.skc_camera_pass(self.ctx, _ptr(rays), len(rays), max_error, mix, _ptr(hits), _ptr(L)))
        return (L, hits) if want_hits else L

    # --- resident inputs
    def set_surface_points(self, points):
        pts = np.ascontiguousarray(points if points.dtype == DT_SURFACE_POINT else points.view(DT_SURFACE_POINT))
        self._ck(self.lib.skc_set_surface_points(self.ctx, _ptr(pts), len(pts)))

    def irradiance_resident(self, first, count, n_tasks, mix=0.5, counters=False):
        c = np.zeros(3, np.uint64) if counters else None
        self._ck(self.lib.skc_irradiance_resident(self.ctx, first, count, n_tasks, mix, _ptr(c)))
        return c

    def set_camera_rays(self, rays):
        rays = _c(rays, np.float32).reshape(-1, 8)
        self._ck(self.lib.skc_set_camera_rays(self.ctx, _ptr(rays), len(rays)))

    def trace_resident(self, first, count, any_hit=False, want_out=False, counters=False):
        """BVH traversal alone over the resident rays; returns (out or None, counters2 or None)."""
        out = (np.empty(count, np.uint8) if any_hit else np.empty(count, DT_HIT)) if want_out else None
        c = np.zeros(2, np.uint64) if counters else None
        self._ck(self.lib.skc_trace_resident(self.ctx, first, count, int(bool(any_hit)), _ptr(out), _ptr(c)))
        return out, c

    def camera_pass_resident(self, first, count, max_error, mix=0.5, out=None, counters=False):
        c = np.zeros(7, np.uint64) if counters else None
        self._ck(self.lib.skc_camera_pass_resident(self.ctx, first, count, max_error, mix, _ptr(out), _ptr(c)))
        return c

    # --- whole frame (SamplerRenderer::Render around the subsurface term)
    def frame_setup(self, desc):
        """desc: bytes of a skc_frame_desc (HostScene.frame_desc).  Returns the number of camera samples."""
        self._frame_desc = np.frombuffer(bytes(desc), np.uint8).copy()
        n = C.c_uint64(0)
        self._ck(self.lib.skc_frame_setup(self.ctx, _ptr(self._frame_desc), C.byref(n)))
        self._frame_res = tuple(int(v) for v in self._frame_desc[:8].view(np.int32))
        return int(n.value)

    def frame_prefetch(self):
        """Sampler stream + camera rays + closest hits + direct-lighting terms on the second stream (no wait)."""
        self._ck(self.lib.skc_frame_prefetch(self.ctx))

    def frame_samples(self):
        self._ck(self.lib.skc_frame_samples(self.ctx))

    def frame_get_samples(self, first, count):
        fl = int(self.lib.skc_frame_sample_floats(self.ctx))
        xy = np.empty((count, 2), np.float32); arr = np.empty((count, fl), np.float32); rays = np.empty((count, 8), np.float32)
        self._ck(self.lib.skc_frame_get_samples(self.ctx, first, count, _ptr(xy), _ptr(arr), _ptr(rays)))
        return xy, arr, rays

    def frame_render(self, first, count, max_error, mix=0.5, want_L=False):
        """Samples [first, first+count) of the render order; L rows (optional) come back in render order."""
        L = np.empty((count, 30), np.float32) if want_L else None
        self._ck(self.lib.skc_frame_render(self.ctx, first, count, max_error, mix, _ptr(L)))
        return L

    def frame_order(self, first, count):
        """Sampler-stream index of each render-order position in [first, first+count)."""
        idx = np.empty(count, np.uint64)
        self._ck(self.lib.skc_frame_get_order(self.ctx, first, count, _ptr(idx)))
        return idx.astype(np.int64)

    def frame_film_clear(self):
        self._ck(self.lib.skc_frame_film_clear(self.ctx))

    def frame_film_dev(self):
        return self.lib.skc_frame_film_dev(self.ctx)

    def frame_film(self, out=None):
        """Film pixels [yres][xres]{Lxyz, weightSum}; out: a preallocated (e.g. pinned) array to copy into."""
        xres, yres = self._frame_res
        px = np.empty((yres, xres, 4), np.float32) if out is None else out
        self._ck(self.lib.skc_frame_film_get(self.ctx, _ptr(px)))
        return px

    def frame_rgb(self, out=None):
        """The frame as ImageFilm::WriteImage computes it, converted on the device: RGB [yres][xres][3]."""
        xres, yres = self._frame_res
        rgb = np.empty((yres, xres, 3), np.float32) if out is None else out
        self._ck(self.lib.skc_frame_film_rgb(self.ctx, _ptr(rgb)))
        return rgb

    def pinned(self, shape, dtype):
        """numpy array over cudaHostAlloc'ed memory (freed with the device object)."""
        dtype = np.dtype(dtype)
        n = int(np.prod(shape)) * dtype.itemsize
        p = self.lib.skc_alloc_pinned(n)
        if not p:
            raise SkcError("cudaHostAlloc(%d) failed" % n)
        self._pinned = getattr(self, "_pinned", []) + [p]
        buf = (C.c_uint8 * max(n, 1)).from_address(p)
        return np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    # --- MPC
    def profile_batch(self, specs, layers_per_set, opts, resample=True):
        specs = np.ascontiguousarray(specs, DT_LAYER); lps = _c(layers_per_set, np.uint32); opts = np.ascontiguousarray(opts, DT_MPC_OPTIONS)
        n = len(lps)
        ln = C.c_uint32(0)
        self._ck(self.lib.skc_profile_batch(self.ctx, _ptr(specs), _ptr(lps), _ptr(opts), n, int(resample), C.byref(ln), None, None, None, None))
        L = int(ln.value)
        d2 = np.empty((n, L), np.float32); R = np.empty((n, L), np.float32); T = np.empty((n, L), np.float32); tot = np.empty((n, 2), np.float32)
        self._ck(self.lib.skc_profile_batch(self.ctx, _ptr(specs), _ptr(lps), _ptr(opts), n, int(resample), C.byref(ln), _ptr(d2), _ptr(R), _ptr(T), _ptr(tot)))
        return d2, R, T, tot

    def rho_table(self, roughness, eta_t, eta_i=1.0, fixed_fresnel=False):
        """rho_hd [1025][30] of Microfacet(Beckmann(roughness), FresnelDielectric(eta_i, eta_t)) (core/multipole.cpp:518-549)."""
        hd = np.empty((1025, 30), np.float32)
        self._ck(self.lib.skc_rho_table(self.ctx, float(roughness), float(eta_i), float(eta_t), int(fixed_fresnel), _ptr(hd)))
        return hd


class SkinComm:
    """One skc_comm: the NCCL communicator of a SkinDevice among `world` ranks (csrc/skc_comm.cu)."""

    @staticmethod
    def unique_id():
        lib = load_library()
        buf = (C.c_uint8 * 128)()
        if lib.skc_comm_unique_id(buf):
            raise SkcError("ncclGetUniqueId failed")
        return bytes(buf)

    def __init__(self, dev, rank, world, id128):
        self.dev, self.lib = dev, dev.lib
        self._id = (C.c_uint8 * 128).from_buffer_copy(id128)
        self.h = self.lib.skc_comm_create(dev.ctx, rank, world, self._id)
        if not self.h:
            raise SkcError(self.lib.skc_last_error(dev.ctx).decode())
        self.rank, self.world = rank, world

    def _ck(self, rc):
        if rc != 0:
            raise SkcError(self.lib.skc_last_error(self.dev.ctx).decode())

    def allgather_positions(self, shard):
        self._ck(self.lib.skc_comm_allgather_positions(self.h, shard))

    def allgather_rows_begin(self, shard):
        self._ck(self.lib.skc_comm_allgather_rows_begin(self.h, shard))

    def allgather_rows_wait(self):
        self._ck(self.lib.skc_comm_allgather_rows_wait(self.h))

    def reduce_film(self, root=0):
        self._ck(self.lib.skc_comm_reduce_film(self.h, root))

    def max_f32(self, value):
        v = C.c_float(value)
        self._ck(self.lib.skc_comm_max_f32(self.h, C.byref(v)))
        return float(v.value)

    def barrier(self):
        self._ck(self.lib.skc_comm_barrier(self.h))

    def close(self):
        if self.h:
            self.lib.skc_comm_destroy(self.h); self.h = None


class CompactProfile:
    """Host-only view of the compact Rd table skc_upload_profile builds (csrc/skc_mo_table.h); no GPU needed."""

    def __init__(self, table, rcp, tol=1e-5, K=256, M=16, head_cap=16384):
        self.lib = load_library()
        self.table = np.ascontiguousarray(table, np.float32); self.rcp = np.ascontiguousarray(rcp, np.float32)
        self.h = self.lib.skc_moc_build(_ptr(self.table), _ptr(self.rcp), self.table.shape[1], tol, K, M, head_cap)
        if self.h:
            info = np.zeros((30, 4), np.float32); rows = C.c_uint32(0); hf = C.c_uint32(0)
            self.coverage = float(self.lib.skc_moc_info(self.h, _ptr(info), C.byref(rows), C.byref(hf)))
            self.f_lo, self.f_hi, self.head_len, self.max_err = info[:, 0].copy(), info[:, 1].copy(), info[:, 2].astype(int), info[:, 3].copy()
            self.rows, self.head_floats = int(rows.value), int(hf.value)

    def eval(self, lam, d2):
        """(kernel value, sampleProfile value) of wavelength lam at the fp32 distances d2."""
        d2 = np.ascontiguousarray(d2, np.float32)
        out = np.empty(len(d2), np.float32); ex = np.empty(len(d2), np.float32)
        self.lib.skc_moc_eval(self.h, int(lam), _ptr(d2), len(d2), _ptr(out), _ptr(ex))
        return out, ex

    def close(self):
        if self.h:
            self.lib.skc_moc_free(self.h); self.h = None

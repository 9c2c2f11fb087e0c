"""ctypes binding of the C++ host mirror (libskinhost.so, host/skin_integrator.cpp skh_*).  Plumbing."""
import ctypes as C
import os

import numpy as np

from .skc import DT_BVH_NODE, DT_SURFACE_POINT, SkcError

_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None


def load_host_library():
    global _lib
    if _lib is not None:
        return _lib
    path = os.path.join(_HERE, "libskinhost.so")
    if not os.path.exists(path):
        raise SkcError("%s not built (make -C pbrt-v2-skin_b200)" % path)
    lib = C.CDLL(path)
    vp = C.c_void_p
    lib.skh_load.restype = vp; lib.skh_load.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
    lib.skh_load_opts.restype = vp; lib.skh_load_opts.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_char_p, C.c_int]
    lib.skh_free.restype = None; lib.skh_free.argtypes = [vp]
    lib.skh_warning_count.restype = C.c_int; lib.skh_warning_count.argtypes = [vp]
    lib.skh_warning.restype = C.c_char_p; lib.skh_warning.argtypes = [vp, C.c_int]
    lib.skh_film_filename.restype = C.c_char_p; lib.skh_film_filename.argtypes = [vp]
    lib.skh_filter.restype = None; lib.skh_filter.argtypes = [vp, vp, vp]
    lib.skh_accelerator.restype = C.c_int; lib.skh_accelerator.argtypes = [vp, C.c_char_p, C.c_int]
    lib.skh_light_samples.restype = None; lib.skh_light_samples.argtypes = [vp, vp]
    lib.skh_pbrt_render.restype = C.c_int
    lib.skh_pbrt_render.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_int, vp, vp, vp, C.c_char_p, C.c_int]
    lib.skh_sizes.restype = None; lib.skh_sizes.argtypes = [vp, vp]
    for f in ("skh_nodes", "skh_prims", "skh_lights", "skh_mesh_P", "skh_mesh_N", "skh_mesh_uv", "skh_mesh_idx"):
        getattr(lib, f).restype = vp; getattr(lib, f).argtypes = [vp]
    lib.skh_mesh_o2w.restype = None; lib.skh_mesh_o2w.argtypes = [vp, vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.skh_integrator_params.restype = None; lib.skh_integrator_params.argtypes = [vp, vp]
    lib.skh_points_write.restype = C.c_int; lib.skh_points_write.argtypes = [C.c_char_p, vp, C.c_uint64, C.c_int, C.c_char_p, C.c_int]
    lib.skh_points_read.restype = C.c_int; lib.skh_points_read.argtypes = [C.c_char_p, vp, C.c_uint64, C.POINTER(C.c_uint64), C.c_char_p, C.c_int]
    lib.skh_poisson_params.restype = None; lib.skh_poisson_params.argtypes = [vp, vp, vp]
    lib.skh_skin_params.restype = None; lib.skh_skin_params.argtypes = [vp, vp]
    lib.skh_albedo_tex_info.restype = C.c_int; lib.skh_albedo_tex_info.argtypes = [vp, vp, vp, vp]
    lib.skh_albedo_tex_rgb.restype = vp; lib.skh_albedo_tex_rgb.argtypes = [vp]
    lib.skh_albedo_tex_lut.restype = vp; lib.skh_albedo_tex_lut.argtypes = [vp]
    lib.skh_bump_tex_info.restype = C.c_int; lib.skh_bump_tex_info.argtypes = [vp, vp, vp, vp]
    lib.skh_bump_tex_texels.restype = vp; lib.skh_bump_tex_texels.argtypes = [vp]
    lib.skh_bump_tex_lut.restype = vp; lib.skh_bump_tex_lut.argtypes = [vp]
    lib.skh_alpha_info.restype = None; lib.skh_alpha_info.argtypes = [vp, vp, vp]
    lib.skh_alpha_texels.restype = vp; lib.skh_alpha_texels.argtypes = [vp]
    lib.skh_env_sizes.restype = C.c_float; lib.skh_env_sizes.argtypes = [vp, vp]
    lib.skh_env_array.restype = vp; lib.skh_env_array.argtypes = [vp, C.c_int]
    lib.skh_skin_albedo.restype = C.c_int; lib.skh_skin_albedo.argtypes = [vp, vp]
    lib.skh_tessellate.restype = C.c_uint64; lib.skh_tessellate.argtypes = [vp, C.c_float, C.c_int, C.POINTER(vp)]
    lib.skh_count_points.restype = C.c_uint64; lib.skh_count_points.argtypes = [vp, C.c_float]
    lib.skh_camera_rays.restype = None; lib.skh_camera_rays.argtypes = [vp, vp, C.c_uint64, vp]
    lib.skh_bench_rays.restype = C.c_uint64; lib.skh_bench_rays.argtypes = [vp, C.POINTER(vp)]
    lib.skh_irradiance_tasks.restype = C.c_uint32; lib.skh_irradiance_tasks.argtypes = [C.c_uint64, C.c_int]
    lib.skh_skin_coefficients.restype = None; lib.skh_skin_coefficients.argtypes = [vp, vp]
    lib.skh_skin_spectra.restype = None; lib.skh_skin_spectra.argtypes = [vp, vp]
    lib.skh_build_profile.restype = C.c_int; lib.skh_build_profile.argtypes = [vp, vp, vp, C.POINTER(C.c_uint32), vp, vp, C.c_int]
    lib.skh_frame_desc.restype = None; lib.skh_frame_desc.argtypes = [vp, C.c_int, vp]
    lib.skh_film_to_rgb.restype = None; lib.skh_film_to_rgb.argtypes = [vp, C.c_int, C.c_int, vp]
    lib.skh_write_pfm.restype = C.c_int; lib.skh_write_pfm.argtypes = [C.c_char_p, vp, C.c_int, C.c_int]
    _lib = lib
    return lib


def _view(ptr, nbytes, dtype):
    if not ptr or nbytes == 0:
        return None
    buf = (C.c_uint8 * nbytes).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype)


def skin_spectra(f_mel=0.15, f_eu=1.0, f_blood=0.002, f_ohg=0.3, nmperunit=100e6, layers=((0.25e6, 1.4), (20e6, 1.4))):
    """LayeredSkin layer coefficients from the pigment parameters alone (defaults = materials/layeredskin.cpp:234-257):
    dict of mua/musp [2][30] in scene units, eta[2], thickness[2]."""
    lib = load_host_library()
    p = np.array([f_mel, f_eu, f_blood, f_ohg, nmperunit, layers[0][0], layers[0][1], layers[1][0], layers[1][1]], np.float32)
    v = np.zeros(124, np.float32)
    lib.skh_skin_spectra(p.ctypes.data_as(C.c_void_p), v.ctypes.data_as(C.c_void_p))
    c = v[:120].reshape(2, 2, 30)
    return {"flat": v, "mua": c[:, 0], "musp": c[:, 1], "eta": v[120:122], "thickness": v[122:124]}


def write_points_file(path, points, versioned=False):
    """The integrator's `pointsfile`: the reference's raw SurfacePoint[44 B] dump, or with the 32-byte tag in front (host/skin_scene.h)."""
    from .skc import DT_SURFACE_POINT
    pts = np.ascontiguousarray(points if points.dtype == DT_SURFACE_POINT else points.view(DT_SURFACE_POINT))
    err = C.create_string_buffer(512)
    if load_host_library().skh_points_write(os.fsencode(path), pts.ctypes.data_as(C.c_void_p), len(pts), int(versioned), err, 512):
        raise RuntimeError(err.value.decode())


def read_points_file(path):
    """Either form of the points file (the tag decides); raises on a tagged file that is truncated, foreign or of another version."""
    from .skc import DT_SURFACE_POINT
    lib = load_host_library()
    err = C.create_string_buffer(512); n = C.c_uint64(0)
    if lib.skh_points_read(os.fsencode(path), None, 0, C.byref(n), err, 512):
        raise RuntimeError(err.value.decode())
    out = np.zeros(n.value, DT_SURFACE_POINT)
    if n.value and lib.skh_points_read(os.fsencode(path), out.ctypes.data_as(C.c_void_p), n.value, C.byref(n), err, 512):
        raise RuntimeError(err.value.decode())
    return out


class HostScene:
    """A scene file loaded and built by the C++ host mirror (BVH, world-space mesh, lights)."""

    def __init__(self, path, quick=False, outfile=None):
        """path: a `.pbrt` scene description (host/skin_pbrt.cpp) or the line format of scenes/*.scene.  quick / outfile are
        the reference's --quick / --outfile command-line options (they reach the Film, Sampler and light factories)."""
        self.lib = load_host_library()
        self.path = path
        err = C.create_string_buffer(1024)
        self.h = self.lib.skh_load_opts(path.encode(), int(bool(quick)), None if outfile is None else outfile.encode(), err, 1024)
        if not self.h:
            raise SkcError(err.value.decode())
        self.warnings = [self.lib.skh_warning(self.h, i).decode() for i in range(self.lib.skh_warning_count(self.h))]
        self.film_filename = self.lib.skh_film_filename(self.h).decode()
        sz = np.zeros(10, np.int64)
        self.lib.skh_sizes(self.h, sz.ctypes.data_as(C.c_void_p))
        (self.n_nodes, self.n_prims, self.n_verts, self.n_tris, self.n_lights, has_n, has_uv, self.xres, self.yres, self.spp) = [int(x) for x in sz]
        L = self.lib
        self.nodes = _view(L.skh_nodes(self.h), 32 * self.n_nodes, DT_BVH_NODE)
        self.prims = _view(L.skh_prims(self.h), 16 * self.n_prims, np.int32).reshape(-1, 4)
        self.lights = _view(L.skh_lights(self.h), 280 * self.n_lights, np.float32).reshape(-1, 70)
        self.P = _view(L.skh_mesh_P(self.h), 12 * self.n_verts, np.float32).reshape(-1, 3)
        self.N = _view(L.skh_mesh_N(self.h), 12 * self.n_verts, np.float32).reshape(-1, 3) if has_n else None
        self.uv = _view(L.skh_mesh_uv(self.h), 8 * self.n_verts, np.float32).reshape(-1, 2) if has_uv else None
        self.idx = _view(L.skh_mesh_idx(self.h), 12 * self.n_tris, np.int32).reshape(-1, 3)
        self.o2w = np.zeros(32, np.float32)
        r, s = C.c_int(0), C.c_int(0)
        L.skh_mesh_o2w(self.h, self.o2w.ctypes.data_as(C.c_void_p), C.byref(r), C.byref(s))
        self.reverse, self.swaps = r.value, s.value
        ip = np.zeros(4, np.float32); L.skh_integrator_params(self.h, ip.ctypes.data_as(C.c_void_p))
        self.maxerror, self.minsampledistance, self.mix, self.maxdepth = float(ip[0]), float(ip[1]), float(ip[2]), int(ip[3])
        pu = np.zeros(2, np.int32); self.camera_position = np.zeros(3, np.float32)
        L.skh_poisson_params(self.h, pu.ctypes.data_as(C.c_void_p), self.camera_position.ctypes.data_as(C.c_void_p))
        self.use_poisson, self.poisson_max_fails = bool(pu[0]), int(pu[1])
        self.skin = np.zeros(12, np.float32); L.skh_skin_params(self.h, self.skin.ctypes.data_as(C.c_void_p))

    def pixel_filter(self):
        """(xwidth, ywidth, ImageFilm::filterTable [16][16])"""
        w = np.zeros(2, np.float32); t = np.zeros(256, np.float32)
        self.lib.skh_filter(self.h, w.ctypes.data_as(C.c_void_p), t.ctypes.data_as(C.c_void_p))
        return float(w[0]), float(w[1]), t.reshape(16, 16)

    def accelerator(self):
        buf = C.create_string_buffer(64)
        n = self.lib.skh_accelerator(self.h, buf, 64)
        return buf.value.decode(), int(n)

    def light_samples(self):
        out = np.zeros(max(1, self.n_lights), np.int32)
        self.lib.skh_light_samples(self.h, out.ctypes.data_as(C.c_void_p))
        return out[:self.n_lights]

    def close(self):
        if self.h:
            self.lib.skh_free(self.h); self.h = None

    def tessellate(self, min_dist=None, incenter=False):
        p = C.c_void_p()
        n = self.lib.skh_tessellate(self.h, self.minsampledistance if min_dist is None else min_dist, int(incenter), C.byref(p))
        return _view(p.value, 44 * n, DT_SURFACE_POINT) if n else np.zeros(0, DT_SURFACE_POINT)

    def count_points(self, min_dist=None):
        return int(self.lib.skh_count_points(self.h, self.minsampledistance if min_dist is None else min_dist))

    def camera_rays(self, xy):
        xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
        out = np.empty((len(xy), 8), np.float32)
        self.lib.skh_camera_rays(self.h, xy.ctypes.data_as(C.c_void_p), len(xy), out.ctypes.data_as(C.c_void_p))
        return out

    def bench_rays(self):
        p = C.c_void_p()
        n = self.lib.skh_bench_rays(self.h, C.byref(p))
        return _view(p.value, 32 * n, np.float32).reshape(-1, 8)

    def albedo_rgb(self):
        """The material's \"color albedo\" (rgb), or None when the scene leaves it at Spectrum(1)."""
        rgb = np.zeros(3, np.float32)
        return rgb if self.lib.skh_skin_albedo(self.h, rgb.ctypes.data_as(C.c_void_p)) else None

    def irradiance_tasks(self, n_points, n_cores=1):
        return int(self.lib.skh_irradiance_tasks(n_points, n_cores))

    def envmap(self):
        """The infinite light's radiance map and sampling distribution (host/skin_envmap.cpp), or None."""
        wh = np.zeros(2, np.int32)
        marg_int = float(self.lib.skh_env_sizes(self.h, wh.ctypes.data_as(C.c_void_p)))
        w, h = int(wh[0]), int(wh[1])
        if not w:
            return None
        n = {"rgb": 3 * w * h, "condFunc": w * h, "condCdf": h * (w + 1), "condInt": h, "margFunc": h, "margCdf": h + 1}
        out = {"w": w, "h": h, "margInt": np.float32(marg_int)}
        for i, k in enumerate(n):
            out[k] = _view(self.lib.skh_env_array(self.h, i), 4 * n[k], np.float32).copy()
        return out

    def albedo_texture(self):
        """The material's "texture albedo": the MIPMap pyramid (levels' sizes and rgb texels), the EWA weight table and the uv mapping; or None."""
        w = np.zeros(12, np.int32); h = np.zeros(12, np.int32); m = np.zeros(4, np.float32)
        n = self.lib.skh_albedo_tex_info(self.h, w.ctypes.data_as(C.c_void_p), h.ctypes.data_as(C.c_void_p), m.ctypes.data_as(C.c_void_p))
        if not n:
            return None
        total = int(sum(3 * int(w[i]) * int(h[i]) for i in range(n)))
        return {"w": w[:n].copy(), "h": h[:n].copy(), "map": tuple(float(x) for x in m),
                "rgb": _view(self.lib.skh_albedo_tex_rgb(self.h), 4 * total, np.float32).copy(), "lut": _view(self.lib.skh_albedo_tex_lut(self.h), 512, np.float32).copy()}

    def bump_texture(self):
        """The material's "texture bumpmap": the MIPMap<float> pyramid (levels' sizes, one displacement per texel), the EWA weight table and the uv mapping; or None."""
        w = np.zeros(12, np.int32); h = np.zeros(12, np.int32); m = np.zeros(4, np.float32)
        n = self.lib.skh_bump_tex_info(self.h, w.ctypes.data_as(C.c_void_p), h.ctypes.data_as(C.c_void_p), m.ctypes.data_as(C.c_void_p))
        if not n:
            return None
        total = int(sum(int(w[i]) * int(h[i]) for i in range(n)))
        return {"w": w[:n].copy(), "h": h[:n].copy(), "map": tuple(float(x) for x in m),
                "texels": _view(self.lib.skh_bump_tex_texels(self.h), 4 * total, np.float32).copy(), "lut": _view(self.lib.skh_bump_tex_lut(self.h), 512, np.float32).copy()}

    def alpha_texture(self):
        """The mesh's "texture alpha" (level-0 float texels and its uv mapping), or None."""
        wh = np.zeros(2, np.int32); m = np.zeros(4, np.float32)
        self.lib.skh_alpha_info(self.h, wh.ctypes.data_as(C.c_void_p), m.ctypes.data_as(C.c_void_p))
        if not wh[0]:
            return None
        w, h = int(wh[0]), int(wh[1])
        return {"w": w, "h": h, "map": tuple(float(x) for x in m), "texels": _view(self.lib.skh_alpha_texels(self.h), 4 * w * h, np.float32).reshape(h, w).copy()}

    def upload(self, dev):
        dev.upload_bvh(self.nodes, self.prims, self.P, self.N, self.uv, self.idx, self.o2w, self.reverse, self.swaps, self.lights)
        dev.upload_lights(self.lights)
        dev.upload_envmap(self.envmap())
        dev.upload_alpha(self.alpha_texture())
        dev.upload_albedo_texture(self.albedo_texture())
        dev.upload_bump_texture(self.bump_texture())

    FRAME_DESC_BYTES = 4 * (5 + 16 + 16 + 2 + 256 + 90 + 1 + 2 + 60 + 1 + 2)

    def frame_desc(self, n_cores=1, rank=0, world=1):
        """skc_frame_desc for this scene: camera, film + box filter, LDSampler, render task count (bytes).
        rank / world: this context renders every world-th render task (round-robin tiles)."""
        buf = np.zeros(self.FRAME_DESC_BYTES, np.uint8)
        self.lib.skh_frame_desc(self.h, int(n_cores), buf.ctypes.data_as(C.c_void_p))
        buf[-8:].view(np.int32)[:] = (rank, world)
        return buf

    def film_to_rgb(self, pixels4):
        """ImageFilm::WriteImage's pixel arithmetic: [h][w]{Lxyz, weightSum} -> RGB [h][w][3]."""
        px = np.ascontiguousarray(pixels4, np.float32)
        h, w = px.shape[:2]
        rgb = np.empty((h, w, 3), np.float32)
        self.lib.skh_film_to_rgb(px.ctypes.data_as(C.c_void_p), w, h, rgb.ctypes.data_as(C.c_void_p))
        return rgb

    def write_pfm(self, path, rgb):
        rgb = np.ascontiguousarray(rgb, np.float32)
        if self.lib.skh_write_pfm(path.encode(), rgb.ctypes.data_as(C.c_void_p), rgb.shape[1], rgb.shape[0]):
            raise SkcError("cannot write %s" % path)

    def skin_coefficients(self):
        """LayeredSkin layer coefficients (materials/layeredskin.cpp:49-92): dict of mua/musp [2][30], eta[2], thickness[2]."""
        v = np.zeros(124, np.float32)
        self.lib.skh_skin_coefficients(self.h, v.ctypes.data_as(C.c_void_p))
        c = v[:120].reshape(2, 2, 30)
        return {"flat": v, "mua": c[:, 0], "musp": c[:, 1], "eta": v[120:122], "thickness": v[122:124]}

    def build_profile(self, dev, rho_hd=None, upload=True):
        """ComputeMultipoleProfile for the scene's skin on the device (30 wavelengths, one batched call); uploads the
        Rd tables and rho_hd [1025][30] (the given one, else generated on the device by skc_rho_table) into dev's context when upload is set.  Returns (hdr [30][4], rd [30][len])."""
        ln = C.c_uint32(0)
        if self.lib.skh_build_profile(self.h, dev.ctx, None, C.byref(ln), None, None, 0):
            raise SkcError(dev.lib.skc_last_error(dev.ctx).decode())
        hdr = np.empty((30, 4), np.float32); rd = np.empty((30, int(ln.value)), np.float32)
        rho = None if rho_hd is None else np.ascontiguousarray(rho_hd, np.float32)
        rc = self.lib.skh_build_profile(self.h, dev.ctx, None if rho is None else rho.ctypes.data_as(C.c_void_p), C.byref(ln),
                                        hdr.ctypes.data_as(C.c_void_p), rd.ctypes.data_as(C.c_void_p), int(bool(upload)))
        if rc:
            raise SkcError(dev.lib.skc_last_error(dev.ctx).decode())
        return hdr, rd


def pbrt_render(scene_path, device=0, n_cores=1, quick=False, outfile=None, write_image=False):
    """`pbrt [--ncores n] [--quick] [--outfile f] scene` through the C++ host (host/skin_render.cpp RenderScene): returns
    (rgb [yres][xres][3], stats dict).  The image file is written only when write_image is set."""
    lib = load_host_library()
    err = C.create_string_buffer(1024)
    dims = (C.c_int * 2)()
    of = None if outfile is None else outfile.encode()
    if lib.skh_pbrt_render(scene_path.encode(), device, n_cores, int(bool(quick)), of, 0, dims, None, None, err, 1024):
        raise SkcError(err.value.decode())
    rgb = np.empty((dims[1], dims[0], 3), np.float32)
    st = np.zeros(7, np.float64)
    if lib.skh_pbrt_render(scene_path.encode(), device, n_cores, int(bool(quick)), of, int(bool(write_image)), dims,
                           rgb.ctypes.data_as(C.c_void_p), st.ctypes.data_as(C.c_void_p), err, 1024):
        raise SkcError(err.value.decode())
    keys = ("build_s", "material_s", "preprocess_s", "render_s", "points", "samples", "octree_nodes")
    return rgb, dict(zip(keys, st.tolist()))

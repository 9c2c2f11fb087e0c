"""pbrt-v2-skin_b200 -- B200-native subsurface hot path of Thomi6York/pbrt-v2-skin.

The directory name carries the reference's name and is not a Python identifier; load it with

    import importlib.util, sys
    spec = importlib.util.spec_from_file_location("pbrt_v2_skin_b200", ".../pbrt-v2-skin_b200/__init__.py",
                                                  submodule_search_locations=[".../pbrt-v2-skin_b200"])
    mod = importlib.util.module_from_spec(spec); sys.modules["pbrt_v2_skin_b200"] = mod; spec.loader.exec_module(mod)

(`__graft_entry__.load_package()` and tests/conftest.py do exactly that.)

Python here is test/bench plumbing over the C ABI of include/skin_cuda.h (ctypes): the product
is libskincuda.so (csrc/*.cu, sm_100a) and the C++ host mirror of the reference's plugin
interfaces (host/).  There is no CPU fallback: every entry point raises if the library or a
GPU is missing.
"""
from .skc import SkinDevice, SkinComm, SkcError, load_library, lib_path, DT_HIT, DT_RAY, DT_SURFACE_POINT, DT_IRR_POINT, DT_BVH_NODE, DT_LIGHT  # noqa: F401
from .host_api import HostScene, load_host_library, skin_spectra, pbrt_render, read_points_file, write_points_file  # noqa: F401
from .multigpu import shard_range, exchange_unique_id  # noqa: F401

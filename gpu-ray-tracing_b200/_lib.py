"""ctypes binding of libb2rt.so (include/b2rt.h).  No CPU fallback: if the library is missing or a
call fails, an exception is raised -- nothing in this package computes a result any other way."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libb2rt.so")

LAYOUT_COMPACT2 = 5
TRACE_ANY_HIT = 1
TRACE_WRITE_BARY = 2
TRACE_STATS = 4
TRACE_WATERTIGHT = 8
COUNT_HITS_STATIC_MAX = 16 * 1024 * 1024
TUNE_NO_STREAM, TUNE_GREEDY_WIDE, TUNE_L2_WINDOW, TUNE_DEFAULT_CARVEOUT = 1, 4, 8, 16

RAY_DTYPE = np.dtype([("o", "<f4", 3), ("tmin", "<f4"), ("d", "<f4", 3), ("tmax", "<f4")])
RESULT_DTYPE = np.dtype([("id", "<i4"), ("t", "<f4"), ("padA", "<i4"), ("padB", "<i4")])
KEY_DTYPE = np.dtype([("oldSlot", "<i4"), ("hash", "<u4", 6)])


class B2rtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("b2rt error %d: %s" % (code, msg))
        self.code = code


class TraceStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("node_iters", "node_lanes", "tri_iters", "tri_lanes", "fetches", "rays", "warps", "ns_pool_empty",
                                           "ns_mean_exit", "ns_last_exit", "ns_exit_p50", "ns_exit_p90", "ns_exit_p99", "ns_straggler_fetch", "straggler_node_iters", "straggler_tri_iters", "shares")]


class WorkCounters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("rays", "nodes", "tris", "leaves", "ustage", "vstage", "hits", "max_stack", "max_chain")]


class TraceConfig(C.Structure):
    _fields_ = [("top_nodes", C.c_int), ("fetch_threshold", C.c_int), ("blocks_per_sm", C.c_int), ("variant", C.c_int), ("l2_prefetch", C.c_int), ("tune", C.c_uint)]


class BuildParams(C.Structure):
    _fields_ = [("split_alpha", C.c_float), ("min_leaf", C.c_int), ("max_leaf", C.c_int), ("threads", C.c_int),
                ("spatial_splits", C.c_int)]


_lib = None

vp = C.c_void_p
i32 = C.c_int
i64 = C.c_int64
f32 = C.c_float
u32 = C.c_uint32
sz = C.c_size_t

_SIGNATURES = {
    "b2rt_create": (i32, [i32, C.POINTER(vp)]),
    "b2rt_destroy": (None, [vp]),
    "b2rt_last_error": (C.c_char_p, [vp]),
    "b2rt_launch_count": (i64, [vp]),
    "b2rt_set_trace_config": (i32, [vp, C.POINTER(TraceConfig)]),
    "b2rt_device_info": (i32, [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]),
    "b2rt_set_bvh": (i32, [vp, vp, sz, vp, sz, vp, sz, i32, i32]),
    "b2rt_get_bvh": (i32, [vp, C.POINTER(vp), C.POINTER(sz), C.POINTER(vp), C.POINTER(sz), C.POINTER(vp), C.POINTER(sz)]),
    "b2rt_trace": (i32, [vp, vp, vp, i64, C.c_uint, vp, C.POINTER(f32)]),
    "b2rt_trace_host": (i32, [vp, vp, vp, i64, C.c_uint, C.POINTER(f32), C.POINTER(f32)]),
    "b2rt_trace_host_begin": (i32, [vp, vp, vp, i64, C.c_uint]),
    "b2rt_trace_host_end": (i32, [vp, C.POINTER(f32), C.POINTER(f32)]),
    "b2rt_get_trace_stats": (i32, [vp, C.POINTER(TraceStats)]),
    "b2rt_count_work": (i32, [vp, vp, i64, i32, C.POINTER(WorkCounters), vp]),
    "b2rt_raygen_primary": (i32, [vp, vp, vp, i32, i32, f32, vp, vp, vp, vp, vp]),
    "b2rt_raygen_ao": (i32, [vp, i32, i32, i32, f32, u32, vp, vp, vp, vp, vp, vp, vp]),
    "b2rt_morton_sort": (i32, [vp, vp, vp, vp, i64, vp, vp]),
    "b2rt_set_sort_mode": (i32, [vp, i32]),
    "b2rt_sort_fell_back": (i32, [vp]),
    "b2rt_find_aabb": (i32, [vp, vp, i64, vp, vp]),
    "b2rt_morton_keys": (i32, [vp, vp, i64, vp, vp, vp]),
    "b2rt_reorder_rays": (i32, [vp, vp, vp, vp, vp, vp, vp, i64, vp]),
    "b2rt_count_hits": (i32, [vp, vp, i64, C.POINTER(i64), vp]),
    "b2rt_count_hits_static": (i32, [vp, i64, vp, vp]),
    "b2rt_reconstruct": (i32, [vp, i32, i32, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp]),
    "b2rt_pixel_table": (i32, [i32, i32, vp, vp]),
    "b2rt_camera_matrices": (i32, [vp, vp, vp, f32, f32, f32, i32, i32, vp, vp]),
    "b2rt_bvh_build": (i32, [vp, i32, vp, i32, C.POINTER(BuildParams), C.POINTER(vp)]),
    "b2rt_bvh_free": (None, [vp]),
    "b2rt_bvh_arrays": (i32, [vp, C.POINTER(vp), C.POINTER(sz), C.POINTER(vp), C.POINTER(sz), C.POINTER(vp), C.POINTER(sz)]),
    "b2rt_bvh_stats": (i32, [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32), C.POINTER(f32), C.POINTER(i32)]),
    "b2rt_bvh_save": (i32, [vp, C.c_char_p]),
    "b2rt_bvh_load": (i32, [C.c_char_p, C.POINTER(vp)]),
    "b2rt_bvhcache_key": (i32, [vp, i32, vp, i32, vp, f32, i32, i32, i32, C.POINTER(u32)]),
    "b2rt_bvhcache_path": (i32, [C.c_char_p, u32, C.c_char_p, sz]),
    "b2rt_tri_normals": (i32, [vp, vp, i32, vp]),
    "b2rt_woopify": (i32, [vp, vp, vp, vp, vp, i64, vp, i64, vp]),
    "b2rt_wide_build": (i32, [vp, sz, vp, sz, vp, sz, C.POINTER(vp)]),
    "b2rt_wide_free": (None, [vp]),
    "b2rt_wide_arrays": (i32, [vp, C.POINTER(vp), C.POINTER(sz), C.POINTER(vp), C.POINTER(sz), C.POINTER(vp), C.POINTER(sz), vp]),
}

class MgStats(C.Structure):
    _fields_ = [("rays", C.c_int64), ("hits", C.c_int64), ("ms_trace_max", C.c_float), ("ms_trace_min", C.c_float),
                ("ms_with_gather_max", C.c_float), ("ms_per_device", C.c_float * 16)]


# include/b2rt_mg.h (same library)
_MG_SIGNATURES = {
    "b2rt_mg_create": (i32, [i32, vp, C.POINTER(vp)]),
    "b2rt_mg_destroy": (None, [vp]),
    "b2rt_mg_last_error": (C.c_char_p, [vp]),
    "b2rt_mg_num_devices": (i32, [vp]),
    "b2rt_mg_context": (vp, [vp, i32]),
    "b2rt_mg_stream": (vp, [vp, i32]),
    "b2rt_mg_broadcast_bvh": (i32, [vp, vp, sz, vp, sz, vp, sz, i32, i32]),
    "b2rt_mg_trace": (i32, [vp, vp, vp, vp, C.c_uint, vp, C.POINTER(MgStats)]),
    "b2rt_mg_trace_host": (i32, [vp, vp, vp, i64, C.c_uint, C.POINTER(MgStats)]),
}
MG_EXPORTS = sorted(_MG_SIGNATURES)

_SCENE_SIGNATURES = {
    "b2rt_scene_generate": (i32, [i32, i32, C.c_uint64, C.POINTER(vp)]),
    "b2rt_scene_free": (None, [vp]),
    "b2rt_scene_arrays": (i32, [vp, C.POINTER(vp), C.POINTER(i32), C.POINTER(vp), C.POINTER(i32)]),
    "b2rt_scene_camera": (i32, [vp, vp]),
    "b2rt_scene_write_obj": (i32, [vp, C.c_char_p]),
}

# what include/b2rt.h declares, what include/b2rt_tools.h adds (instrumentation), what libb2rt_scenes.so exports
TOOLS_EXPORTS = sorted(["b2rt_get_trace_stats", "b2rt_count_work", "b2rt_wide_build", "b2rt_wide_free", "b2rt_wide_arrays"])
EXPORTS = sorted(n for n in _SIGNATURES if n not in TOOLS_EXPORTS)
_SIGNATURES.update(_MG_SIGNATURES)
SCENES_EXPORTS = sorted(_SCENE_SIGNATURES)
SCENES_LIB_PATH = os.path.join(HERE, "libb2rt_scenes.so")
_scenes = None


def scenes_lib():
    """libb2rt_scenes.so: the procedural benchmark scenes (include/b2rt_scenes.h).  Host only."""
    global _scenes
    if _scenes is None:
        if not os.path.exists(SCENES_LIB_PATH):
            raise ImportError("libb2rt_scenes.so is not built: run `python gpu-ray-tracing_b200/build.py`")
        l = C.CDLL(SCENES_LIB_PATH)
        for name, (res, args) in _SCENE_SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _scenes = l
    return _scenes


def lib():
    """The loaded library.  Raises if libb2rt.so has not been built (python gpu-ray-tracing_b200/build.py)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("libb2rt.so is not built: run `python gpu-ray-tracing_b200/build.py` "
                              "(there is no CPU fallback)")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(rc, ctx=None):
    if rc != 0:
        msg = lib().b2rt_last_error(ctx)
        raise B2rtError(rc, msg.decode() if msg else "")


def np_ptr(a):
    return a.ctypes.data_as(vp)

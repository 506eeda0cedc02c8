"""gpu-ray-tracing_b200 -- Blackwell-native ray-casting core behind the reference's own interface.

The compute lives in libb2rt.so (hand-written CUDA for sm_100a, C ABI in include/b2rt.h); this
package is the thin Python host side: same names as the reference's C++ classes for the path
(CudaTracer.traceBatch, RayBuffer.mortonSort, RayGen.primary/ao, CudaBVH, Renderer ...).
Import it with importlib (the directory name has a hyphen):

    import importlib; b2rt = importlib.import_module("gpu-ray-tracing_b200")
"""
from ._lib import (B2rtError, EXPORTS, MG_EXPORTS, MgStats, SCENES_EXPORTS, SCENES_LIB_PATH, TOOLS_EXPORTS, scenes_lib, KEY_DTYPE, LAYOUT_COMPACT2, LIB_PATH, RAY_DTYPE, RESULT_DTYPE,
                   TRACE_ANY_HIT, TRACE_STATS, TRACE_WATERTIGHT, TRACE_WRITE_BARY, lib)
from .host import (BVH, Context, CudaBVH, CudaTracer, PixelTable, RayBuffer, RayGen, Renderer, Scene,
                   camera_matrices, countHits, decode_camera_signature, encode_camera_signature, fail, findAABB, leaf_references, mortonKeys, reconstruct, save_image, woopify_device)

__all__ = ["B2rtError", "BVH", "Context", "CudaBVH", "CudaTracer", "EXPORTS", "KEY_DTYPE", "LAYOUT_COMPACT2",
           "LIB_PATH", "PixelTable", "RAY_DTYPE", "RESULT_DTYPE", "RayBuffer", "RayGen", "Renderer", "Scene",
           "TRACE_ANY_HIT", "TRACE_STATS", "TRACE_WATERTIGHT", "TRACE_WRITE_BARY", "camera_matrices", "countHits", "decode_camera_signature", "encode_camera_signature", "fail", "findAABB", "lib", "leaf_references", "mortonKeys", "reconstruct", "save_image", "woopify_device"]

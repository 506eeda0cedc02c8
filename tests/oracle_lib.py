"""ctypes bindings for oracle/libb2rt_oracle.so and readers for the interchange files.

Test infrastructure: imported only by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package never imports this module.
"""
import ctypes as C
import os
import struct
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "libb2rt_oracle.so")
REF_CPU = os.path.join(ORACLE_DIR, "_ref", "ref_cpu")

RAY_DTYPE = np.dtype([("o", "<f4", 3), ("tmin", "<f4"), ("d", "<f4", 3), ("tmax", "<f4")])
RESULT_DTYPE = np.dtype([("id", "<i4"), ("t", "<f4"), ("padA", "<i4"), ("padB", "<i4")])
KEY_DTYPE = np.dtype([("oldSlot", "<i4"), ("hash", "<u4", 6)])
assert RAY_DTYPE.itemsize == 32 and RESULT_DTYPE.itemsize == 16 and KEY_DTYPE.itemsize == 28


class Counters(C.Structure):
    _fields_ = [(n, C.c_int64) for n in
                ("rays", "nodes", "tris", "leaves", "ustage", "vstage", "hits", "maxStack")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


_lib = None


def build_oracle():
    """Compile oracle/libb2rt_oracle.so (and oracle/_ref when /root/reference exists)."""
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "all"], check=True)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(ORACLE_SO):
            build_oracle()
        _lib = C.CDLL(ORACLE_SO)
        _lib.b2o_count_hits.restype = C.c_int64
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


# ---------------------------------------------------------------------------------------------
# wrappers
# ---------------------------------------------------------------------------------------------
def pixel_table(w, h):
    i2p = np.empty(w * h, np.int32)
    p2i = np.empty(w * h, np.int32)
    lib().b2o_pixel_table(C.c_int(w), C.c_int(h), _p(i2p), _p(p2i))
    return i2p, p2i


def camera_matrices(pos, forward, up, fov, znear, zfar, w, h):
    origin = np.empty(3, np.float32)
    m = np.empty(16, np.float32)
    lib().b2o_camera_matrices(_p(_f32(pos)), _p(_f32(forward)), _p(_f32(up)), C.c_float(fov),
                              C.c_float(znear), C.c_float(zfar), C.c_int(w), C.c_int(h), _p(origin), _p(m))
    return origin, m


def raygen_primary(origin, nscreen_to_world, w, h, max_dist, index_to_pixel):
    rays = np.empty(w * h, RAY_DTYPE)
    id2slot = np.empty(w * h, np.int32)
    slot2id = np.empty(w * h, np.int32)
    lib().b2o_raygen_primary(_p(_f32(origin)), _p(_f32(nscreen_to_world)), C.c_int(w), C.c_int(h),
                             C.c_float(max_dist), _p(_i32(index_to_pixel)), _p(rays), _p(id2slot), _p(slot2id))
    return rays, id2slot, slot2id


def raygen_ao(first_slot, num_in, num_samples, max_dist, seed, in_rays, in_results, normals):
    n = num_in * num_samples
    rays = np.empty(n, RAY_DTYPE)
    id2slot = np.empty(n, np.int32)
    slot2id = np.empty(n, np.int32)
    in_rays = np.ascontiguousarray(in_rays)
    in_results = np.ascontiguousarray(in_results)
    normals = _f32(normals)
    lib().b2o_raygen_ao(C.c_int(first_slot), C.c_int(num_in), C.c_int(num_samples), C.c_float(max_dist),
                        C.c_uint32(seed), _p(in_rays), _p(in_results), _p(normals),
                        _p(rays), _p(id2slot), _p(slot2id))
    return rays, id2slot, slot2id


def find_aabb(rays):
    rays = np.ascontiguousarray(rays)
    lo = np.empty(3, np.float32)
    hi = np.empty(3, np.float32)
    lib().b2o_find_aabb(_p(rays), C.c_int(len(rays)), _p(lo), _p(hi))
    return lo, hi


def morton_keys(rays, lo, hi):
    rays = np.ascontiguousarray(rays)
    keys = np.empty(len(rays), KEY_DTYPE)
    lib().b2o_morton_keys(_p(rays), C.c_int(len(rays)), _p(_f32(lo)), _p(_f32(hi)), _p(keys))
    return keys


def sort_keys(keys):
    keys = np.ascontiguousarray(keys).copy()
    lib().b2o_sort_keys(_p(keys), C.c_int(len(keys)))
    return keys


def reorder_rays(keys, in_rays, in_slot2id):
    n = len(keys)
    keys = np.ascontiguousarray(keys)
    in_rays = np.ascontiguousarray(in_rays)
    in_slot2id = _i32(in_slot2id)
    rays = np.empty(n, RAY_DTYPE)
    id2slot = np.empty(n, np.int32)
    slot2id = np.empty(n, np.int32)
    lib().b2o_reorder_rays(C.c_int(n), _p(keys), _p(in_rays), _p(in_slot2id), _p(rays), _p(id2slot), _p(slot2id))
    return rays, id2slot, slot2id


def morton_sort(rays, slot2id=None):
    """RayBuffer::mortonSort (RayBuffer.cc:256-324) end to end, with the oracle's stable sort."""
    if slot2id is None:
        slot2id = np.arange(len(rays), dtype=np.int32)
    lo, hi = find_aabb(rays)
    keys = sort_keys(morton_keys(rays, lo, hi))
    out = reorder_rays(keys, rays, slot2id)
    return out + (keys, lo, hi)


def woopify(v0, v1, v2):
    out = np.empty(12, np.float32)
    lib().b2o_woopify(_p(_f32(v0)), _p(_f32(v1)), _p(_f32(v2)), _p(out))
    return out


def tri_normals(verts, tris):
    verts = _f32(verts)
    tris = _i32(tris)
    out = np.empty((len(tris), 3), np.float32)
    lib().b2o_tri_normals(_p(verts), _p(tris), C.c_int(len(tris)), _p(out))
    return out


def trace_mt(bvh, scene, rays, need_closest_hit=True, results=None):
    rays = np.ascontiguousarray(rays)
    if results is None:
        results = np.zeros(len(rays), RESULT_DTYPE)
    lib().b2o_trace_mt(_p(bvh["nodes"]), _p(bvh["woop"]), _p(bvh["triIndex"]), _p(scene["verts"]),
                       _p(scene["tris"]), _p(rays), _p(results), C.c_int(len(rays)),
                       C.c_int(1 if need_closest_hit else 0))
    return results


def trace_woop(bvh, rays, any_hit=False, counters=False):
    rays = np.ascontiguousarray(rays)
    results = np.zeros(len(rays), RESULT_DTYPE)
    ctr = Counters()
    lib().b2o_trace_woop(_p(bvh["nodes"]), _p(bvh["woop"]), _p(bvh["triIndex"]), _p(rays), _p(results),
                         C.c_int(len(rays)), C.c_int(1 if any_hit else 0), C.byref(ctr) if counters else None)
    return (results, ctr.as_dict()) if counters else results


def count_hits(results):
    results = np.ascontiguousarray(results)
    return int(lib().b2o_count_hits(_p(results), C.c_int64(len(results))))


def reconstruct(num_per_primary, first_primary, num_primary, mode, primary_slot2id, primary_results,
                batch_id2slot, batch_results, tri_material, tri_shaded, pixels):
    """mode in {'primary','ao','diffuse'}; pixels is updated in place (uint32 ABGR)."""
    lib().b2o_reconstruct(C.c_int(num_per_primary), C.c_int(first_primary), C.c_int(num_primary),
                          C.c_int(mode == "primary"), C.c_int(mode == "ao"), C.c_int(mode == "diffuse"),
                          _p(_i32(primary_slot2id)), _p(np.ascontiguousarray(primary_results)),
                          _p(_i32(batch_id2slot)), _p(np.ascontiguousarray(batch_results)),
                          _p(np.ascontiguousarray(tri_material, np.uint32)),
                          _p(np.ascontiguousarray(tri_shaded, np.uint32)), _p(pixels))
    return pixels


# ---------------------------------------------------------------------------------------------
# interchange files (written by oracle/_ref/ref_cpu and by the product's own dump helpers)
# ---------------------------------------------------------------------------------------------
def read_bvh(path):
    """bvhcache format (Renderer.cc:176-211, CudaBVH.cc:113-116, Buffer.cc:327-361):
    S32 layout, then nodes / triWoop / triIndex each as S64 size + raw bytes (little endian)."""
    with open(path, "rb") as f:
        data = f.read()
    layout, = struct.unpack_from("<i", data, 0)
    ofs = 4
    bufs = []
    for _ in range(3):
        size, = struct.unpack_from("<q", data, ofs)
        ofs += 8
        bufs.append(data[ofs:ofs + size])
        ofs += size
    assert ofs == len(data), "trailing bytes in bvh file"
    return {"layout": layout,
            "nodes": np.frombuffer(bufs[0], np.float32).copy(),
            "woop": np.frombuffer(bufs[1], np.float32).copy(),
            "triIndex": np.frombuffer(bufs[2], np.int32).copy()}


def write_bvh(path, bvh):
    with open(path, "wb") as f:
        f.write(struct.pack("<i", int(bvh.get("layout", 5))))
        for k in ("nodes", "woop", "triIndex"):
            b = np.ascontiguousarray(bvh[k]).tobytes()
            f.write(struct.pack("<q", len(b)))
            f.write(b)


def read_scene(path):
    with open(path, "rb") as f:
        magic, nv, nt = struct.unpack("<iii", f.read(12))
        assert magic == 0x43533242, "not a B2SC scene file"
        verts = np.frombuffer(f.read(12 * nv), np.float32).reshape(nv, 3).copy()
        tris = np.frombuffer(f.read(12 * nt), np.int32).reshape(nt, 3).copy()
        normals = np.frombuffer(f.read(12 * nt), np.float32).reshape(nt, 3).copy()
    return {"verts": verts, "tris": tris, "normals": normals}


def _read_counted(path, dtype):
    with open(path, "rb") as f:
        n, = struct.unpack("<i", f.read(4))
        return np.frombuffer(f.read(n * np.dtype(dtype).itemsize), dtype).copy()


def read_rays(path):
    return _read_counted(path, RAY_DTYPE)


def read_results(path):
    return _read_counted(path, RESULT_DTYPE)


def read_ints(path):
    return _read_counted(path, np.int32)


def write_rays(path, rays):
    rays = np.ascontiguousarray(rays, RAY_DTYPE)
    with open(path, "wb") as f:
        f.write(struct.pack("<i", len(rays)))
        f.write(rays.tobytes())


def read_camera(path):
    a = np.fromfile(path, np.float32)
    assert a.size == 20
    return {"origin": a[:3].copy(), "nscreenToWorld": a[3:19].copy(), "far": float(a[19])}


def write_obj(path, verts, tris):
    with open(path, "w") as f:
        for v in np.asarray(verts):
            f.write("v %.9g %.9g %.9g\n" % (v[0], v[1], v[2]))
        for t in np.asarray(tris):
            f.write("f %d %d %d\n" % (t[0] + 1, t[1] + 1, t[2] + 1))


def have_ref_cpu():
    return os.path.exists(REF_CPU) and os.access(REF_CPU, os.X_OK)


def run_ref_cpu(*args):
    """Run oracle/_ref/ref_cpu (the reference's own host code); returns its stdout."""
    r = subprocess.run([REF_CPU] + [str(a) for a in args], check=True, capture_output=True, text=True)
    return r.stdout


# ---------------------------------------------------------------------------------------------
# the reference's own CUDA tracer, compatibility build (oracle/_ref/libref_gpu.so, needs a GPU)
# ---------------------------------------------------------------------------------------------
REF_GPU_SO = os.path.join(ORACLE_DIR, "_ref", "libref_gpu.so")


def have_ref_gpu():
    return os.path.exists(REF_GPU_SO)


class RefGpuTracer:
    """kepler_dynamic_fetch `trace` through its own launcher (kepler_dynamic_fetch.cu:417-474).
    Arguments are torch CUDA tensors.  `warps`: 720 is the reference's literal launch
    (CudaTracer.cc:156); pass SM count x resident warps for the scaled variant."""

    def __init__(self):
        self.lib = C.CDLL(REF_GPU_SO)
        self.lib.launch_tracingKernel.restype = C.c_float
        self.lib.launch_tracingKernel.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_bool] + [C.c_void_p] * 10
        self.lib.bind_CudaBVHTexture.restype = None
        self.lib.bind_CudaBVHTexture.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64]
        self._bvh = None

    def set_bvh(self, nodes, woop, tri_index):
        # the reference reads three rows before it tests the terminator (texture fetches return 0
        # out of range); with plain loads the Woop array needs two rows of padding
        import torch
        woop = torch.cat([woop.reshape(-1), torch.zeros(8, dtype=woop.dtype, device=woop.device)])
        self._bvh = (nodes, woop, tri_index)
        self.lib.bind_CudaBVHTexture(nodes.data_ptr(), nodes.numel() * 4, woop.data_ptr(), woop.numel() * 4,
                                     tri_index.data_ptr(), tri_index.numel() * 4)

    def trace(self, rays, results, any_hit=False, warps=720):
        """Returns the kernel milliseconds the reference's launcher measures (cudaEvents, :449-473)."""
        nodes, woop, tri = self._bvh
        block = (C.c_int * 2)(32, 4)                      # CudaTracer.cc:105-107 asks for 32x4
        threads = int(warps) * 32
        # the launcher prints its grid each call (:447); keep stdout clean for bench.py's JSON line
        fd = os.dup(1)
        devnull = os.open(os.devnull, os.O_WRONLY)
        os.dup2(devnull, 1)
        try:
            ms = self.lib.launch_tracingKernel(threads, block, rays.shape[0], bool(any_hit), rays.data_ptr(),
                                               results.data_ptr(), nodes.data_ptr(), nodes.data_ptr(), nodes.data_ptr(),
                                               nodes.data_ptr(), woop.data_ptr(), woop.data_ptr(), woop.data_ptr(),
                                               tri.data_ptr())
        finally:
            os.dup2(fd, 1)
            os.close(fd)
            os.close(devnull)
        return float(ms)

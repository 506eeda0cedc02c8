"""Host-side mirror of the reference's C++ interface for the ray-casting path.

Same class and method names, argument meaning and error behaviour as the reference (paths
under its src/):  Scene (rt/Scene.hh), BVH (rt/bvh/BVH.hh), CudaBVH (rt/cuda/CudaBVH.hh:86-137),
RayBuffer (rt/ray/RayBuffer.hh:37-97), PixelTable (rt/ray/PixelTable.hh), RayGen
(rt/ray/RayGen.hh:41-69), CudaTracer (rt/cuda/CudaTracer.hh:39-69) and the headless part of
Renderer (rt/cuda/Renderer.hh:49-146).  Everything that computes runs in libb2rt.so through the
C ABI (include/b2rt.h); PyTorch only owns device memory and streams.  There is no CPU path.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import (BuildParams, B2rtError, LAYOUT_COMPACT2, RAY_DTYPE, RESULT_DTYPE, TRACE_ANY_HIT,
                   TRACE_STATS, TRACE_WATERTIGHT, TRACE_WRITE_BARY, TraceConfig, TraceStats, check, lib, np_ptr, scenes_lib)

SCENE_KINDS = {"blob": 0, "interior": 1, "knot": 2, "strands": 3}


def fail(msg):
    """FW::fail (framework/base/Defs.cc:41-54) prints and exits; here it raises."""
    raise RuntimeError(msg)


def _stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _dptr(t):
    return C.c_void_p(t.data_ptr())


# ---------------------------------------------------------------------------------------------
class Context:
    """One b2rt_ctx per device (replaces CudaModule::staticInit + the tracer's process globals)."""
    _cache = {}

    def __init__(self, device=0):
        self.device = int(device)
        h = C.c_void_p()
        rc = lib().b2rt_create(self.device, C.byref(h))
        if rc != 0:
            raise B2rtError(rc, lib().b2rt_last_error(None).decode())
        self.handle = h

    @classmethod
    def get(cls, device=None):
        if device is None:
            device = torch.cuda.current_device()
        device = int(device)
        if device not in cls._cache:
            cls._cache[device] = Context(device)
        return cls._cache[device]

    def launch_count(self):
        return int(lib().b2rt_launch_count(self.handle))

    def sm_count(self):
        sm = C.c_int()
        check(lib().b2rt_device_info(self.handle, C.byref(sm), None, None), self.handle)
        return sm.value

    def set_trace_config(self, top_nodes=-1, fetch_threshold=-1, blocks_per_sm=-1, variant=-1, l2_prefetch=-1, tune=0):
        cfg = TraceConfig(top_nodes, fetch_threshold, blocks_per_sm, variant, l2_prefetch, tune)
        check(lib().b2rt_set_trace_config(self.handle, C.byref(cfg)), self.handle)

    def set_sort_mode(self, mode=0):
        """0 = prefix radix sort + tie settling (default), 1 = always the full 192-bit LSD sort (b2rt_set_sort_mode)."""
        check(lib().b2rt_set_sort_mode(self.handle, int(mode)), self.handle)

    def sort_fell_back(self):
        return bool(lib().b2rt_sort_fell_back(self.handle))

    def torch_device(self):
        return torch.device("cuda", self.device)


# ---------------------------------------------------------------------------------------------
class Scene:
    """rt/Scene.hh: flat triangle list, vertex positions and per-triangle normals (Scene.cc:35-84)."""

    def __init__(self, verts, tris, camera=None, name="mesh"):
        self.verts = np.ascontiguousarray(verts, np.float32).reshape(-1, 3)
        self.tris = np.ascontiguousarray(tris, np.int32).reshape(-1, 3)
        self.normals = np.empty((len(self.tris), 3), np.float32)
        check(lib().b2rt_tri_normals(np_ptr(self.verts), np_ptr(self.tris), len(self.tris), np_ptr(self.normals)))
        self.camera = camera        # dict(pos, forward, up, fov, near, far) or None
        self.name = name
        self._normals_dev = {}

    @staticmethod
    def generate(kind, target_tris, seed):
        """Seeded procedural stand-in for a README scene (libb2rt_scenes.so, csrc/host/scenegen.cpp)."""
        sl = scenes_lib()
        h = C.c_void_p()
        if sl.b2rt_scene_generate(SCENE_KINDS[kind], int(target_tris), C.c_uint64(seed), C.byref(h)) != 0:
            raise ValueError("b2rt_scene_generate(%s, %d) failed" % (kind, target_tris))
        try:
            v, t = C.c_void_p(), C.c_void_p()
            nv, nt = C.c_int(), C.c_int()
            sl.b2rt_scene_arrays(h, C.byref(v), C.byref(nv), C.byref(t), C.byref(nt))
            verts = np.ctypeslib.as_array(C.cast(v, C.POINTER(C.c_float)), (nv.value, 3)).copy()
            tris = np.ctypeslib.as_array(C.cast(t, C.POINTER(C.c_int32)), (nt.value, 3)).copy()
            cam = np.empty(12, np.float32)
            sl.b2rt_scene_camera(h, np_ptr(cam))
        finally:
            sl.b2rt_scene_free(h)
        camera = {"pos": cam[0:3].copy(), "forward": cam[3:6].copy(), "up": cam[6:9].copy(),
                  "fov": float(cam[9]), "near": float(cam[10]), "far": float(cam[11])}
        return Scene(verts, tris, camera, name="%s-%d" % (kind, target_tris))

    def getNumTriangles(self):
        return len(self.tris)

    def getNumVertices(self):
        return len(self.verts)

    def getTriNormalBuffer_dev(self, device):
        key = str(device)
        if key not in self._normals_dev:
            self._normals_dev[key] = torch.from_numpy(self.normals).to(device)
        return self._normals_dev[key]

    def init_camera_for_mesh(self):
        """CameraControls::initForMesh (framework/3d/CameraControls.cc:330-350)."""
        lo, hi = self.verts.min(0), self.verts.max(0)
        size = float((hi - lo).max())
        c = (lo + hi) * np.float32(0.5)
        self.camera = {"pos": np.array([c[0], c[1], c[2] + np.float32(size) * np.float32(0.75)], np.float32),
                       "forward": np.array([0, 0, -1], np.float32), "up": np.array([0, 1, 0], np.float32),
                       "fov": 70.0, "near": float(np.float32(size) * np.float32(0.0005)),
                       "far": float(np.float32(size) * np.float32(1.5))}
        return self.camera

    def write_obj(self, path):
        with open(path, "w") as f:
            for v in self.verts:
                f.write("v %.9g %.9g %.9g\n" % (v[0], v[1], v[2]))
            for t in self.tris:
                f.write("f %d %d %d\n" % (t[0] + 1, t[1] + 1, t[2] + 1))


def camera_matrices(camera, w, h):
    """(origin, nscreenToWorld column-major) as Renderer::beginFrame computes (Renderer.cc:126-129)."""
    origin = np.empty(3, np.float32)
    m = np.empty(16, np.float32)
    check(lib().b2rt_camera_matrices(np_ptr(np.ascontiguousarray(camera["pos"], np.float32)),
                                     np_ptr(np.ascontiguousarray(camera["forward"], np.float32)),
                                     np_ptr(np.ascontiguousarray(camera["up"], np.float32)),
                                     C.c_float(camera["fov"]), C.c_float(camera["near"]), C.c_float(camera["far"]),
                                     int(w), int(h), np_ptr(origin), np_ptr(m)))
    return origin, m


# ---------------------------------------------------------------------------------------------
class BVH:
    """rt/bvh/BVH.hh: SBVH over a Scene.  Built by csrc/host/bvh_builder.cpp; holds the Compact2
    arrays (what CudaBVH(bvh, BVHLayout_Compact2) produces in the reference)."""

    def __init__(self, scene=None, split_alpha=1.0e-5, min_leaf=1, max_leaf=8, threads=0, spatial_splits=True,
                 _handle=None):
        self.scene = scene
        if _handle is None:
            p = BuildParams(split_alpha, min_leaf, max_leaf, threads, 1 if spatial_splits else 0)
            h = C.c_void_p()
            check(lib().b2rt_bvh_build(np_ptr(scene.verts), len(scene.verts), np_ptr(scene.tris), len(scene.tris),
                                       C.byref(p), C.byref(h)))
            _handle = h
        self._h = _handle
        n, w, t = C.c_void_p(), C.c_void_p(), C.c_void_p()
        nb, wb, tb = C.c_size_t(), C.c_size_t(), C.c_size_t()
        check(lib().b2rt_bvh_arrays(self._h, C.byref(n), C.byref(nb), C.byref(w), C.byref(wb), C.byref(t), C.byref(tb)))
        self.nodes = np.ctypeslib.as_array(C.cast(n, C.POINTER(C.c_float)), (nb.value // 4,))
        self.woop = np.ctypeslib.as_array(C.cast(w, C.POINTER(C.c_float)), (wb.value // 4,))
        self.triIndex = np.ctypeslib.as_array(C.cast(t, C.POINTER(C.c_int32)), (tb.value // 4,))
        a, b, c2, d = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        s = C.c_float()
        check(lib().b2rt_bvh_stats(self._h, C.byref(a), C.byref(b), C.byref(c2), C.byref(s), C.byref(d)))
        self.stats = {"numInnerNodes": a.value, "numLeafNodes": b.value, "numTris": c2.value, "SAHCost": s.value,
                      "maxDepth": d.value}

    def __del__(self):
        try:
            if self._h:
                lib().b2rt_bvh_free(self._h)
                self._h = None
        except Exception:
            pass

    def arrays(self):
        return {"layout": LAYOUT_COMPACT2, "nodes": self.nodes, "woop": self.woop, "triIndex": self.triIndex}

    def serialize(self, path):
        """bvhcache file (Renderer.cc:176-211; CudaBVH.cc:113-116)."""
        check(lib().b2rt_bvh_save(self._h, str(path).encode()))

    @staticmethod
    def load(path, scene=None):
        h = C.c_void_p()
        rc = lib().b2rt_bvh_load(str(path).encode(), C.byref(h))
        if rc != 0:
            raise B2rtError(rc, "cannot load BVH cache file %s" % path)
        return BVH(scene, _handle=h)


class CudaBVH:
    """rt/cuda/CudaBVH.hh: the three device arrays of a Compact2 BVH (nodes, triWoop, triIndex)."""

    def __init__(self, bvh, layout=LAYOUT_COMPACT2, device=None):
        if isinstance(bvh, BVH):
            arrays = bvh.arrays()
        else:
            arrays = bvh        # dict with nodes / woop / triIndex (e.g. read from a bvhcache file)
        self.m_layout = int(arrays.get("layout", layout)) if layout == LAYOUT_COMPACT2 else int(layout)
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.m_nodes = self._upload(arrays["nodes"], np.float32)
        self.m_triWoop = self._upload(arrays["woop"], np.float32)
        self.m_triIndex = self._upload(arrays["triIndex"], np.int32)

    def _upload(self, array, dtype):
        """numpy (or anything numpy reads) -> a fresh device tensor; a torch tensor already on the device is taken as it is
        (multigpu.broadcast_bvh leaves the broadcast arrays there)."""
        if isinstance(array, torch.Tensor):
            want = torch.float32 if dtype == np.float32 else torch.int32
            if array.dtype != want:
                raise TypeError("CudaBVH: device array of dtype %s, expected %s" % (array.dtype, want))
            return array.contiguous().reshape(-1).to(self.device)
        return torch.from_numpy(np.ascontiguousarray(array, dtype).copy()).to(self.device)

    @classmethod
    def fromSceneOnDevice(cls, scene, arrays, ctx=None, device=None):
        """Upload nodes and triIndex, compute the Woop rows ON the device from the scene's vertices (b2rt_woopify;
        replaces the woopifyTri loop of CudaBVH::createCompact, CudaBVH.cc:270-380).  `arrays` needs nodes and
        triIndex only; the rows come out bit-identical to the host-built ones."""
        self = cls.__new__(cls)
        self.m_layout = int(arrays.get("layout", LAYOUT_COMPACT2))
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        ctx = ctx or Context.get(self.device.index)
        self.m_nodes = torch.from_numpy(np.ascontiguousarray(arrays["nodes"], np.float32).copy()).to(self.device)
        self.m_triIndex = torch.from_numpy(np.ascontiguousarray(arrays["triIndex"], np.int32).copy()).to(self.device)
        self.m_triWoop = woopify_device(scene, arrays["nodes"], arrays["triIndex"], ctx, self.device)
        return self

    def getLayout(self):
        return self.m_layout

    def getNodeBuffer_dev(self):
        return self.m_nodes

    def getTriWoopBuffer_dev(self):
        return self.m_triWoop

    def getTriIndexBuffer_dev(self):
        return self.m_triIndex


def leaf_references(nodes, triIndex):
    """(first row, triangle) of every triangle reference of a Compact2 BVH, from the node array's leaf pointers: a
    negative child index is ~(first row of the leaf) (CudaBVH.cc:300-330), a leaf is 3 rows per triangle plus one
    terminator row, so the gaps between consecutive leaf starts give the triangle counts."""
    n = np.ascontiguousarray(nodes, np.float32).view(np.int32).reshape(-1, 16)
    tri = np.ascontiguousarray(triIndex, np.int32).reshape(-1)
    child = n[:, 12:14].reshape(-1)
    starts = np.unique(~child[child < 0]).astype(np.int64)
    ends = np.append(starts[1:], len(tri))              # one past the leaf's terminator row
    counts = (ends - starts - 1) // 3
    if ((ends - starts - 1) % 3 != 0).any() or (counts < 0).any():
        raise ValueError("leaf_references: leaf rows are not 3 per triangle + 1 terminator")
    leaf_of_ref = np.repeat(np.arange(len(starts)), counts)
    within = np.arange(int(counts.sum())) - np.repeat(np.cumsum(counts) - counts, counts)
    rows = (starts[leaf_of_ref] + 3 * within).astype(np.int32)
    return rows, tri[rows].astype(np.int32)


def woopify_device(scene, nodes, triIndex, ctx=None, device=None):
    """b2rt_woopify: device tensor of len(triIndex) Woop rows (4 floats each) for the scene's triangles."""
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    ctx = ctx or Context.get(device.index)
    rows, tris = leaf_references(nodes, triIndex)
    num_rows = int(np.asarray(triIndex).size)
    verts_d = torch.from_numpy(np.ascontiguousarray(scene.verts, np.float32).reshape(-1).copy()).to(device)
    index_d = torch.from_numpy(np.ascontiguousarray(scene.tris, np.int32).reshape(-1).copy()).to(device)
    rows_d = torch.from_numpy(rows).to(device)
    tris_d = torch.from_numpy(tris).to(device)
    woop = torch.empty(num_rows * 4, dtype=torch.float32, device=device)
    check(lib().b2rt_woopify(ctx.handle, _dptr(verts_d), _dptr(index_d), _dptr(rows_d), _dptr(tris_d), len(rows), _dptr(woop),
                             num_rows, _stream_ptr()), ctx.handle)
    return woop


# ---------------------------------------------------------------------------------------------
class RayBuffer:
    """rt/ray/RayBuffer.hh:37-97.  Device arrays: rays (32 B), results (16 B), idToSlot, slotToID."""

    def __init__(self, n=0, closestHit=True, device=None):
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.m_size = 0
        self.m_capacity = 0
        self.m_needClosestHit = bool(closestHit)
        self.m_rays = self.m_results = self.m_IDToSlot = self.m_slotToID = None
        self.rayAABB = None
        self.resize(n)

    def getSize(self):
        return self.m_size

    def resize(self, n):
        """RayBuffer.cc:53-81: shrinking is logical, growing reallocates (contents lost)."""
        n = int(n)
        assert n >= 0
        if n <= self.m_capacity and self.m_rays is not None:
            self.m_size = n
            return
        self.m_capacity = n
        self.m_size = n
        self.m_rays = torch.empty((max(n, 1), 8), dtype=torch.float32, device=self.device)
        self.m_results = torch.empty((max(n, 1), 4), dtype=torch.int32, device=self.device)
        self.m_IDToSlot = torch.empty(max(n, 1), dtype=torch.int32, device=self.device)
        self.m_slotToID = torch.empty(max(n, 1), dtype=torch.int32, device=self.device)

    def setNeedClosestHit(self, c):
        self.m_needClosestHit = bool(c)

    def getNeedClosestHit(self):
        return self.m_needClosestHit

    def getRayBuffer_dev(self):
        return self.m_rays[:self.m_size]

    def getResultBuffer_dev(self):
        return self.m_results[:self.m_size]

    def getIDToSlotBuffer_dev(self):
        return self.m_IDToSlot[:self.m_size]

    def getSlotToIDBuffer_dev(self):
        return self.m_slotToID[:self.m_size]

    def setRays(self, rays, ids=None):
        """Upload a whole host array of rays (setRay slot by slot, RayBuffer.cc:227-235)."""
        rays = np.ascontiguousarray(rays)
        n = len(rays)
        self.resize(n)
        if n == 0:
            return
        self.m_rays[:n].copy_(torch.from_numpy(rays.view(np.float32).reshape(n, 8)))
        if ids is None:
            ar = torch.arange(n, dtype=torch.int32, device=self.device)
            self.m_slotToID[:n].copy_(ar)
            self.m_IDToSlot[:n].copy_(ar)
        else:
            ids = torch.from_numpy(np.ascontiguousarray(ids, np.int32)).to(self.device)
            self.m_slotToID[:n].copy_(ids)
            self.m_IDToSlot[ids.long()] = torch.arange(n, dtype=torch.int32, device=self.device)

    def getRays(self):
        return self.m_rays[:self.m_size].cpu().numpy().view(RAY_DTYPE).reshape(-1)

    def getResults(self):
        return self.m_results[:self.m_size].cpu().numpy().view(RESULT_DTYPE).reshape(-1)

    def getResultsByID(self):
        """results looked up through idToSlot (getResultForID, RayBuffer.hh:56-60)."""
        res = self.getResults()
        return res[self.m_IDToSlot[:self.m_size].cpu().numpy()]

    def getRayAABB(self):
        return self.rayAABB

    def mortonSort(self, ctx=None):
        """RayBuffer::mortonSort (RayBuffer.cc:256-324), on the device."""
        ctx = ctx or Context.get(self.device.index)
        if self.m_size == 0:
            return
        check(lib().b2rt_morton_sort(ctx.handle, _dptr(self.m_rays), _dptr(self.m_IDToSlot), _dptr(self.m_slotToID),
                                     self.m_size, None, _stream_ptr()), ctx.handle)


class PixelTable:
    """rt/ray/PixelTable.hh: index <-> pixel LUTs (Morton-ordered 8x8 blocks)."""

    def __init__(self):
        self.m_size = (0, 0)
        self.indexToPixel = None
        self.pixelToIndex = None
        self._dev = None

    def setSize(self, size, device):
        size = (int(size[0]), int(size[1]))
        if size != self.m_size or self._dev is None:
            self.m_size = size
            n = size[0] * size[1]
            self.indexToPixel = np.empty(n, np.int32)
            self.pixelToIndex = np.empty(n, np.int32)
            check(lib().b2rt_pixel_table(size[0], size[1], np_ptr(self.indexToPixel), np_ptr(self.pixelToIndex)))
            self._dev = torch.from_numpy(self.indexToPixel).to(device)

    def getIndexToPixel_dev(self):
        return self._dev


class RayGen:
    """rt/ray/RayGen.hh:41-69."""

    def __init__(self, maxBatchSize=8 * 1024 * 1024, ctx=None):
        self.m_maxBatchSize = int(maxBatchSize)
        self.m_pixelTable = PixelTable()
        self.m_aoStartIdx = 0
        self._ctx = ctx
        self._libc_rand = _LibcRand()

    def _context(self, device):
        return self._ctx or Context.get(device.index)

    def primary(self, orays, origin, nscreenToWorld, w, h, maxDist):
        """RayGen::primary (RayGen.cc:50-73)."""
        ctx = self._context(orays.device)
        self.m_pixelTable.setSize((w, h), orays.device)
        orays.resize(w * h)
        orays.setNeedClosestHit(True)
        origin = np.ascontiguousarray(origin, np.float32)
        m = np.ascontiguousarray(nscreenToWorld, np.float32).reshape(16)
        check(lib().b2rt_raygen_primary(ctx.handle, np_ptr(origin), np_ptr(m), int(w), int(h), C.c_float(maxDist),
                                        _dptr(self.m_pixelTable.getIndexToPixel_dev()), _dptr(orays.m_rays),
                                        _dptr(orays.m_IDToSlot), _dptr(orays.m_slotToID), _stream_ptr()), ctx.handle)

    def batching(self, numInputRays, numSamples, newBatch):
        """RayGen::batching (RayGen.cc:124-142).  Returns (continues, newBatch, lo, hi)."""
        if newBatch:
            newBatch = False
            self.m_aoStartIdx = 0
        if self.m_aoStartIdx == numInputRays:
            return False, newBatch, 0, 0
        lo = self.m_aoStartIdx
        hi = min(numInputRays, lo + self.m_maxBatchSize // numSamples)
        self.m_aoStartIdx = hi
        return True, newBatch, lo, hi

    def ao(self, orays, irays, scene, numSamples, maxDist, newBatch, randomSeed=0):
        """RayGen::ao (RayGen.cc:77-120).  Returns (continues, newBatch).  As in the reference the
        `randomSeed` argument is ignored and the seed is libc's rand() (RayGen.cc:106)."""
        cont, newBatch, lo, hi = self.batching(irays.getSize(), numSamples, newBatch)
        if not cont:
            return False, newBatch
        ctx = self._context(orays.device)
        orays.resize((hi - lo) * numSamples)
        orays.setNeedClosestHit(False)
        seed = self._libc_rand.rand()
        normals = scene.getTriNormalBuffer_dev(orays.device)       # uploaded once, not per batch (RayGen.cc:87-91)
        check(lib().b2rt_raygen_ao(ctx.handle, lo, hi - lo, int(numSamples), C.c_float(maxDist), C.c_uint32(seed),
                                   _dptr(irays.m_rays), _dptr(irays.m_results), _dptr(normals), _dptr(orays.m_rays),
                                   _dptr(orays.m_IDToSlot), _dptr(orays.m_slotToID), _stream_ptr()), ctx.handle)
        self.lastSeed = seed
        return True, newBatch


class _LibcRand:
    """libc rand() with its default seed, which is what RayGen.cc:106 draws from (never srand()'ed)."""

    def __init__(self):
        self._libc = C.CDLL(None)
        self._libc.srand(1)

    def rand(self):
        return int(self._libc.rand())


# ---------------------------------------------------------------------------------------------
class CudaTracer:
    """rt/cuda/CudaTracer.hh:39-69."""

    def __init__(self, ctx=None, device=None):
        self._ctx = ctx or Context.get(device)
        self.m_bvh = None
        self.m_kernelConfig = None
        self.writeBarycentrics = False
        self.collectStats = False
        self.watertight = False         # B2RT_TRACE_WATERTIGHT: edge-tolerant Woop test, order-independent closest hit
        self.initKernel()

    def initKernel(self):
        """CudaTracer.cc:84-115: Compact2, 32x4 blocks, persistent threads."""
        self.m_kernelConfig = {"bvhLayout": LAYOUT_COMPACT2, "blockWidth": 32, "blockHeight": 4, "usePersistentThreads": 1}

    def getDesiredBVHLayout(self):
        return self.m_kernelConfig["bvhLayout"]

    def setBVH(self, bvh):
        """CudaTracer.hh:49 + the texture bind of CudaTracer.cc:142-146."""
        self.m_bvh = bvh
        if bvh is None:
            return
        if bvh.getLayout() != self.getDesiredBVHLayout():
            return      # reported at traceBatch, like the reference
        n, w, t = bvh.getNodeBuffer_dev(), bvh.getTriWoopBuffer_dev(), bvh.getTriIndexBuffer_dev()
        # In the reference this is a pointer assignment and Renderer::beginFrame does it every frame (Renderer.cc:116); here
        # b2rt_set_bvh builds private copies (re-laid nodes, on first use the 8-wide form), so the same CudaBVH object bound
        # again is not uploaded again.  One BVH per context: the last one bound is the one every tracer of the device traces.
        key = (id(bvh), n.data_ptr(), n.numel(), w.data_ptr(), w.numel(), t.data_ptr(), t.numel())
        if getattr(self._ctx, "_bound_bvh", None) == key:
            return
        check(lib().b2rt_set_bvh(self._ctx.handle, _dptr(n), n.numel() * 4, _dptr(w), w.numel() * 4, _dptr(t),
                                 t.numel() * 4, bvh.getLayout(), 1), self._ctx.handle)
        self._ctx._bound_bvh = key
        self._ctx._bound_bvh_ref = bvh          # keeps id(bvh) from being reused while it is the bound one

    def traceBatch(self, rays):
        """CudaTracer::traceBatch (CudaTracer.cc:119-178): returns the kernel time in MILLISECONDS
        (cudaEventElapsedTime, kepler_dynamic_fetch.cu:471-473)."""
        numRays = rays.getSize()
        if not numRays:
            return 0.0
        if self.m_bvh is None:
            fail("CudaTracer: No BVH!")
        if self.m_bvh.getLayout() != self.getDesiredBVHLayout():
            fail("CudaTracer: Incorrect BVH layout!")
        flags = 0 if rays.getNeedClosestHit() else TRACE_ANY_HIT
        if self.writeBarycentrics:
            flags |= TRACE_WRITE_BARY
        if self.collectStats:
            flags |= TRACE_STATS
        if self.watertight:
            flags |= TRACE_WATERTIGHT
        ms = C.c_float()
        check(lib().b2rt_trace(self._ctx.handle, _dptr(rays.m_rays), _dptr(rays.m_results), numRays, flags,
                               _stream_ptr(), C.byref(ms)), self._ctx.handle)
        return ms.value

    def traceBatchAsync(self, rays):
        flags = 0 if rays.getNeedClosestHit() else TRACE_ANY_HIT
        if self.writeBarycentrics:
            flags |= TRACE_WRITE_BARY
        if self.watertight:
            flags |= TRACE_WATERTIGHT
        if rays.getSize():
            check(lib().b2rt_trace(self._ctx.handle, _dptr(rays.m_rays), _dptr(rays.m_results), rays.getSize(), flags,
                                   _stream_ptr(), None), self._ctx.handle)

    def traceHost(self, rays_host, results_host, anyHit=False):
        """End-to-end call for HOST buffers (torch CPU tensors, ideally pinned): H2D, trace, D2H.
        Returns (kernel_ms, total_ms)."""
        n = rays_host.shape[0]
        flags = (TRACE_ANY_HIT if anyHit else 0) | (TRACE_WRITE_BARY if self.writeBarycentrics else 0)
        k, t = C.c_float(), C.c_float()
        check(lib().b2rt_trace_host(self._ctx.handle, C.c_void_p(rays_host.data_ptr()), C.c_void_p(results_host.data_ptr()),
                                    n, flags, C.byref(k), C.byref(t)), self._ctx.handle)
        return k.value, t.value

    def traceHostBegin(self, rays_host, results_host, anyHit=False):
        """First half of traceHost: enqueue upload, trace and download of one host batch and return (up to two batches in
        flight; b2rt_trace_host_begin).  The tensors must stay alive and unread until traceHostEnd."""
        flags = (TRACE_ANY_HIT if anyHit else 0) | (TRACE_WRITE_BARY if self.writeBarycentrics else 0)
        check(lib().b2rt_trace_host_begin(self._ctx.handle, C.c_void_p(rays_host.data_ptr()), C.c_void_p(results_host.data_ptr()),
                                          rays_host.shape[0], flags), self._ctx.handle)

    def traceHostEnd(self):
        """Wait for every batch begun; returns (summed kernel_ms, device ms from the first upload to the last download)."""
        k, t = C.c_float(), C.c_float()
        check(lib().b2rt_trace_host_end(self._ctx.handle, C.byref(k), C.byref(t)), self._ctx.handle)
        return k.value, t.value

    def countWork(self, rays):
        """Canonical per-batch work counters (b2rt_count_work): N_n, N_t, N_l, N_u, N_v, hits."""
        w = _lib.WorkCounters()
        check(lib().b2rt_count_work(self._ctx.handle, _dptr(rays.m_rays), rays.getSize(),
                                    0 if rays.getNeedClosestHit() else 1, C.byref(w), _stream_ptr()), self._ctx.handle)
        return {n: getattr(w, n) for n, _ in w._fields_}

    def getStats(self):
        s = TraceStats()
        check(lib().b2rt_get_trace_stats(self._ctx.handle, C.byref(s)), self._ctx.handle)
        return {n: getattr(s, n) for n, _ in s._fields_}


def countHits(results_dev, ctx=None):
    """launch_countHitsKernel (cuda/RendererKernels.cu:112-162,189-212)."""
    ctx = ctx or Context.get(results_dev.device.index)
    hits = C.c_int64()
    check(lib().b2rt_count_hits(ctx.handle, _dptr(results_dev), results_dev.shape[0], C.byref(hits), _stream_ptr()), ctx.handle)
    return hits.value


# ---------------------------------------------------------------------------------------------
class Renderer:
    """Headless part of rt/cuda/Renderer.hh: the batch state machine around the tracer.
    RayType as Renderer.hh:52-58."""
    RayType_Primary, RayType_AO, RayType_Diffuse = 0, 1, 2

    class Params:
        def __init__(self):                 # Renderer.hh:68-74
            self.rayType = Renderer.RayType_Primary
            self.aoRadius = 1.0
            self.numSamples = 32
            self.sortSecondary = False

    def __init__(self, frameSize, device=None):
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self._ctx = Context.get(self.device.index)
        self.m_raygen = RayGen(1 << 21, self._ctx)          # Renderer.cc:46
        self.m_tracer = CudaTracer(self._ctx)
        self.m_params = Renderer.Params()
        self.m_scene = None
        self.m_bvh = None
        self.m_hostBVH = None
        self.m_ViewSize = (int(frameSize[0]), int(frameSize[1]))
        self.m_primaryRays = RayBuffer(0, True, self.device)
        self.m_secondaryRays = RayBuffer(0, True, self.device)
        self.m_newBatch = True
        self.m_batchRays = None
        self.m_batchStart = 0
        self.m_cameraFar = 0.0
        self.m_buildParams = {"split_alpha": 1.0e-5}
        self.m_bvhCachePath = None

    def setMesh(self, scene):
        if scene is self.m_scene:
            return
        self.m_scene = scene
        self.m_bvh = None

    def setParams(self, params):
        self.m_params = params
        self.m_tracer.initKernel()

    def setBVH(self, host_bvh):
        """Inject an already built BVH (stands in for the bvhcache hit of Renderer.cc:186-196)."""
        self.m_hostBVH = host_bvh
        self.m_bvh = CudaBVH(host_bvh, device=self.device)

    def getCudaBVH(self):
        """Renderer::getCudaBVH (Renderer.cc:157-217): the BVH comes from <m_bvhCachePath>/%08x.dat when that file exists -- same
        name and same bytes as a reference build's cache (b2rt_bvhcache_key) -- and is built and written there otherwise.
        m_bvhCachePath = None (the default here; the reference uses "bvhcache") leaves the file system alone."""
        if self.m_bvh is None:
            path = None
            if self.m_bvhCachePath:
                key = C.c_uint32()
                check(lib().b2rt_bvhcache_key(np_ptr(self.m_scene.verts), len(self.m_scene.verts), np_ptr(self.m_scene.tris),
                                              len(self.m_scene.tris), None, C.c_float(self.m_buildParams["split_alpha"]), 1, 8,
                                              LAYOUT_COMPACT2, C.byref(key)))
                buf = C.create_string_buffer(1024)
                check(lib().b2rt_bvhcache_path(str(self.m_bvhCachePath).encode(), key.value, buf, 1024))
                path = buf.value.decode()
            import os
            if path and os.path.exists(path):
                self.m_hostBVH = BVH.load(path, self.m_scene)
            else:
                self.m_hostBVH = BVH(self.m_scene, split_alpha=self.m_buildParams["split_alpha"], min_leaf=1, max_leaf=8)
                if path:
                    os.makedirs(os.path.dirname(path), exist_ok=True)
                    self.m_hostBVH.serialize(path)
            self.m_bvh = CudaBVH(self.m_hostBVH, device=self.device)
        return self.m_bvh

    def beginFrame(self, camera):
        """Renderer::beginFrame (Renderer.cc:112-152)."""
        self.m_tracer.setBVH(self.getCudaBVH())
        w, h = self.m_ViewSize
        origin, m = camera_matrices(camera, w, h)
        self.m_raygen.primary(self.m_primaryRays, origin, m, w, h, camera["far"])
        if self.m_params.rayType != Renderer.RayType_Primary:
            self.m_tracer.traceBatch(self.m_primaryRays)
        self.m_cameraFar = camera["far"]
        self.m_newBatch = True
        self.m_batchRays = None
        self.m_batchStart = 0

    def getTotalNumRays(self):
        """Renderer::getTotalNumRays (Renderer.cc:221-238): degenerate secondary rays are not counted."""
        if self.m_params.rayType == Renderer.RayType_Primary:
            return self.m_primaryRays.getSize()
        return countHits(self.m_primaryRays.getResultBuffer_dev(), self._ctx) * self.m_params.numSamples

    def nextBatch(self):
        """Renderer::nextBatch (Renderer.cc:242-291)."""
        if self.m_batchRays is not None:
            self.m_batchStart += self.m_batchRays.getSize()
        self.m_batchRays = None
        p = self.m_params
        if p.rayType == Renderer.RayType_Primary:
            if not self.m_newBatch:
                return False
            self.m_newBatch = False
            self.m_batchRays = self.m_primaryRays
        elif p.rayType == Renderer.RayType_AO:
            cont, self.m_newBatch = self.m_raygen.ao(self.m_secondaryRays, self.m_primaryRays, self.m_scene,
                                                     p.numSamples, p.aoRadius, self.m_newBatch, 0)
            if not cont:
                return False
            self.m_batchRays = self.m_secondaryRays
        elif p.rayType == Renderer.RayType_Diffuse:
            cont, self.m_newBatch = self.m_raygen.ao(self.m_secondaryRays, self.m_primaryRays, self.m_scene,
                                                     p.numSamples, self.m_cameraFar, self.m_newBatch, 0)
            if not cont:
                return False
            self.m_secondaryRays.setNeedClosestHit(True)
            self.m_batchRays = self.m_secondaryRays
        else:
            return False
        if p.sortSecondary and p.rayType != Renderer.RayType_Primary:
            self.m_batchRays.mortonSort(self._ctx)
        return True

    def traceBatch(self):
        """Renderer::traceBatch (Renderer.cc:295-300): kernel milliseconds of this batch."""
        assert self.m_batchRays is not None
        return self.m_tracer.traceBatch(self.m_batchRays)


# ---------------------------------------------------------------------------------------------
# the pieces of mortonSort and the reconstruction kernel, exposed for parity tests and the
# headless image path
def findAABB(rays_dev, ctx=None):
    """launch_findAABBKernel (ray/RayBufferKernels.cu:74-140): (lo[3], hi[3]) of origins and end points."""
    ctx = ctx or Context.get(rays_dev.device.index)
    box = np.empty(6, np.float32)
    check(lib().b2rt_find_aabb(ctx.handle, _dptr(rays_dev), rays_dev.shape[0], np_ptr(box), _stream_ptr()), ctx.handle)
    return box[:3].copy(), box[3:].copy()


def mortonKeys(rays_dev, lo, hi, ctx=None):
    """launch_genMortonKeysKernel (ray/RayBufferKernels.cu:144-179): 28-byte MortonKey records."""
    ctx = ctx or Context.get(rays_dev.device.index)
    n = rays_dev.shape[0]
    keys = torch.empty((max(n, 1), 7), dtype=torch.int32, device=rays_dev.device)
    box = np.concatenate([np.asarray(lo, np.float32), np.asarray(hi, np.float32)])
    check(lib().b2rt_morton_keys(ctx.handle, _dptr(rays_dev), n, np_ptr(box), _dptr(keys), _stream_ptr()), ctx.handle)
    return keys[:n].cpu().numpy().view(_lib.KEY_DTYPE).reshape(-1)


def reconstruct(mode, numRaysPerPrimary, firstPrimary, numPrimary, primaryRays, batchRays, triMaterialColor_dev,
                triShadedColor_dev, pixels_dev, ctx=None):
    """launch_reconstructKernel (cuda/RendererKernels.cu:60-108).  mode: 'primary' | 'ao' | 'diffuse'."""
    ctx = ctx or Context.get(pixels_dev.device.index)
    m = {"primary": 0, "ao": 1, "diffuse": 2}[mode]
    check(lib().b2rt_reconstruct(ctx.handle, int(numRaysPerPrimary), int(firstPrimary), int(numPrimary), m,
                                 _dptr(primaryRays.m_slotToID), _dptr(primaryRays.m_results), _dptr(batchRays.m_IDToSlot),
                                 _dptr(batchRays.m_results), _dptr(triMaterialColor_dev), _dptr(triShadedColor_dev),
                                 _dptr(pixels_dev), _stream_ptr()), ctx.handle)


def save_image(pixels, width, height, path, flip=True):
    """Headless stand-in for the reference's display path (App.cc:62-132 draws the pixel buffer with glDrawPixels:
    unavailable on this box): write the U32 ABGR frame that `reconstruct` fills (RendererKernels.cu:60-108, byte 0 =
    red) as a binary PPM.  glDrawPixels puts row 0 at the BOTTOM of the window, so rows are written last-first unless
    flip=False.  `pixels` is a device or host tensor / array of width*height 32-bit words.  Returns the RGB array."""
    if hasattr(pixels, "detach"):
        pixels = pixels.detach().cpu().numpy()
    px = np.ascontiguousarray(pixels).view(np.uint32).reshape(-1)
    if px.size != width * height:
        raise ValueError("save_image: %d pixels for a %dx%d frame" % (px.size, width, height))
    rgb = px.view(np.uint8).reshape(height, width, 4)[:, :, :3]
    if flip:
        rgb = rgb[::-1]
    rgb = np.ascontiguousarray(rgb)
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (width, height))
        f.write(rgb.tobytes())
    return rgb


# ---------------------------------------------------------------------------------------------
# camera signatures: the --camera="..." strings of grtcmdline.txt / test.txt
# (CameraControls::encodeSignature / decodeSignature, framework/3d/CameraControls.cc:354-419,
#  bit packing :472-555: 6 bits per character, floats as 6 characters little-end first)
def _enc_bits(v):
    base = ord('/') if v < 12 else (ord('A') - 12 if v < 38 else ord('a') - 38)
    return chr(v + base)


def _enc_float(f):
    bits = int(np.float32(f).view(np.uint32))
    return "".join(_enc_bits((bits >> i) & 0x3F) for i in range(0, 32, 6))


def _enc_dir(v):
    v = np.asarray(v, np.float32)
    a = np.abs(v)
    axis = 0 if a[0] >= max(a[1], a[2]) else (1 if a[1] >= a[2] else 2)
    tuv = v if axis == 0 else (np.array([v[1], v[2], v[0]], np.float32) if axis == 1 else np.array([v[2], v[0], v[1]], np.float32))
    face = axis | (0 if tuv[0] >= 0 else 4)
    if tuv[1] == 0 and tuv[2] == 0:
        return _enc_bits(face | 8)
    return _enc_bits(face) + _enc_float(tuv[1] / abs(tuv[0])) + _enc_float(tuv[2] / abs(tuv[0]))


def encode_camera_signature(camera, speed=1.0, keep_aligned=False):
    p = camera["pos"]
    return ('"' + _enc_float(p[0]) + _enc_float(p[1]) + _enc_float(p[2]) + _enc_dir(camera["forward"]) + _enc_dir(camera["up"]) +
            _enc_float(speed) + _enc_float(camera["fov"]) + _enc_float(camera["near"]) + _enc_float(camera["far"]) +
            _enc_bits(1 if keep_aligned else 0) + '",')


def decode_camera_signature(sig):
    s = sig.strip()
    if s.startswith('"'):
        s = s[1:]
    pos = [0]

    def bits():
        c = s[pos[0]]
        pos[0] += 1
        if '/' <= c <= ':':
            return ord(c) - ord('/')
        if 'A' <= c <= 'Z':
            return ord(c) - ord('A') + 12
        if 'a' <= c <= 'z':
            return ord(c) - ord('a') + 38
        fail("CameraControls: Invalid signature!")

    def flt():
        b = 0
        for i in range(0, 32, 6):
            b |= bits() << i
        return np.uint32(b & 0xFFFFFFFF).view(np.float32)

    def direction():
        face = bits()
        t = np.array([1.0 if (face & 4) == 0 else -1.0, flt() if (face & 8) == 0 else 0.0, 0.0], np.float32)
        t[2] = flt() if (face & 8) == 0 else 0.0
        l = np.float32(0)
        for k in range(3):
            l = np.float32(l + t[k] * t[k])
        t = t * (np.float32(1) / np.sqrt(l))
        return t if (face & 3) == 0 else (np.array([t[2], t[0], t[1]], np.float32) if (face & 3) == 1 else np.array([t[1], t[2], t[0]], np.float32))

    px, py, pz = flt(), flt(), flt()
    fwd, up = direction(), direction()
    flt()       # speed
    fov, near, far = flt(), flt(), flt()
    return {"pos": np.array([px, py, pz], np.float32), "forward": fwd, "up": up, "fov": float(fov), "near": float(near), "far": float(far)}

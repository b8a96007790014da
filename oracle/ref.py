"""ctypes binding of oracle/_ref/libceres_ref_{strict,fast}.so -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

The .so files are the reference's own headers (render.hpp, lib/bvh, obj_norms.hpp) compiled unchanged from
/root/reference by oracle/Makefile (driver: oracle/ref_driver.cpp).  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import this module.
"""
import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
MISS = 0xFFFFFFFF

NODE_DT = np.dtype([("bounds", "<f4", (6,)), ("primitive_count", "<u4"), ("first_child_or_primitive", "<u4")])
assert NODE_DT.itemsize == 32


def available(kind="strict"):
    return os.path.exists(os.path.join(REF_DIR, f"libceres_ref_{kind}.so"))


_libs = {}


def _lib(kind):
    if kind in _libs:
        return _libs[kind]
    path = os.path.join(REF_DIR, f"libceres_ref_{kind}.so")
    L = C.CDLL(path)
    vp, sz, f32p, u32p, u8p, u64p = C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p
    L.cref_scene_load_obj.restype = vp
    L.cref_scene_load_obj.argtypes = [C.c_char_p, C.c_int, C.c_float, C.c_int]
    L.cref_scene_from_triangles.restype = vp
    L.cref_scene_from_triangles.argtypes = [f32p, f32p, sz, C.c_int]
    L.cref_scene_from_pod.restype = vp
    L.cref_scene_from_pod.argtypes = [vp, sz, vp, vp, vp, sz]
    L.cref_scene_destroy.argtypes = [vp]
    for n in ("cref_scene_node_count", "cref_scene_tri_count"):
        getattr(L, n).restype = sz
        getattr(L, n).argtypes = [vp]
    L.cref_scene_build_seconds.restype = C.c_double
    L.cref_scene_build_seconds.argtypes = [vp]
    for n in ("cref_scene_nodes", "cref_scene_prim_indices", "cref_scene_triangles", "cref_scene_norms"):
        getattr(L, n).restype = vp
        getattr(L, n).argtypes = [vp]
    L.cref_threads.restype = C.c_int
    L.cref_render.restype = C.c_double
    L.cref_render.argtypes = [vp, f32p, f32p, f32p, sz, sz, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.cref_render_ex.restype = C.c_double
    L.cref_render_ex.argtypes = [vp, f32p, f32p, sz, sz, C.c_uint, f32p, u32p, u8p, f32p, u64p]
    L.cref_trace_rays.argtypes = [vp, f32p, sz, u32p, f32p]
    L.cref_slab.argtypes = [f32p, f32p, sz, f32p]
    L.cref_tri_intersect.argtypes = [f32p, f32p, sz, u8p, f32p]
    L.cref_smooth_shading.argtypes = [f32p, sz, f32p]
    L.cref_turntable.argtypes = [f32p, f32p, f32p, C.c_int, f32p]
    _libs[kind] = L
    return L


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def cam_array(eye, dir, up, fov):
    return _f32(list(eye) + list(dir) + list(up) + [fov])


class RefScene:
    """A scene owned by the reference library: triangles, tri_norms and the reference-built Bvh."""

    def __init__(self, handle, kind):
        if not handle:
            raise RuntimeError("reference could not load/build the scene")
        self.h, self.kind, self.L = handle, kind, _lib(kind)

    @classmethod
    def load_obj(cls, path, rot_axis=-1, rot_deg=0.0, builder=0, kind="strict"):
        L = _lib(kind)
        return cls(L.cref_scene_load_obj(path.encode(), rot_axis, rot_deg, builder), kind)

    @classmethod
    def from_triangles(cls, verts, norms=None, builder=0, kind="strict"):
        L = _lib(kind)
        v = _f32(verts).reshape(-1, 9)
        n = _f32(norms).reshape(-1, 9) if norms is not None else None
        return cls(L.cref_scene_from_triangles(_p(v), _p(n), v.shape[0], builder), kind)

    @classmethod
    def from_pod(cls, pod, kind="strict"):
        L = _lib(kind)
        nodes = np.ascontiguousarray(pod["nodes"])
        idx = np.ascontiguousarray(pod["prim_indices"], dtype=np.uint64)
        tris = _f32(pod["triangles"])
        norms = _f32(pod["norms"])
        return cls(L.cref_scene_from_pod(_p(nodes), nodes.shape[0], _p(idx), _p(tris), _p(norms), tris.shape[0]), kind)

    def close(self):
        if self.h:
            self.L.cref_scene_destroy(self.h)
            self.h = None

    @property
    def node_count(self):
        return self.L.cref_scene_node_count(self.h)

    @property
    def tri_count(self):
        return self.L.cref_scene_tri_count(self.h)

    @property
    def build_seconds(self):
        return self.L.cref_scene_build_seconds(self.h)

    def pod(self):
        """Copies of the reference's POD arrays: nodes (32 B), prim_indices (size_t), triangles (n,12), norms (n,9)."""
        nn, nt = self.node_count, self.tri_count

        def view(ptr, nbytes, dt):
            buf = (C.c_char * nbytes).from_address(ptr)
            return np.frombuffer(buf, dtype=dt).copy()

        return {
            "nodes": view(self.L.cref_scene_nodes(self.h), nn * 32, NODE_DT),
            "prim_indices": view(self.L.cref_scene_prim_indices(self.h), nt * 8, np.uint64),
            "triangles": view(self.L.cref_scene_triangles(self.h), nt * 48, np.float32).reshape(nt, 12),
            "norms": view(self.L.cref_scene_norms(self.h), nt * 36, np.float32).reshape(nt, 9),
        }

    def refit(self, tris12):
        """HierarchyRefitter::refit (hierarchy_refitter.hpp) to new triangles (n,12 in primitive order)."""
        t = _f32(tris12).reshape(-1, 12)
        assert t.shape[0] == self.tri_count
        self.L.cref_refit.argtypes = [C.c_void_p, C.c_void_p]
        self.L.cref_refit.restype = None
        self.L.cref_refit(self.h, _p(t))

    def render(self, cam, sun, W, H):
        """The unmodified reference render() (render.hpp:86-156). Returns (pixels[H,W,3], rays, hits, seconds)."""
        cam, sun = _f32(cam), _f32(sun)
        px = np.empty((H, W, 3), np.float32)
        rays, hits = C.c_int(0), C.c_int(0)
        sec = self.L.cref_render(self.h, _p(cam), _p(sun), _p(px), W, H, C.byref(rays), C.byref(hits))
        return px, rays.value, hits.value, sec

    def render_ex(self, cam, sun, W, H, flags=0, want_pixels=True):
        """Instrumented loop. Returns dict(pixels, ids, state, tuv, stats(dict), seconds)."""
        cam, sun = _f32(cam), _f32(sun)
        px = np.empty((H, W, 3), np.float32) if want_pixels else None
        ids = np.empty((H, W), np.uint32)
        st = np.empty((H, W), np.uint8)
        tuv = np.empty((H, W, 3), np.float32)
        stats = np.zeros(8, np.uint64)
        sec = self.L.cref_render_ex(self.h, _p(cam), _p(sun), W, H, flags, _p(px), _p(ids), _p(st), _p(tuv), _p(stats))
        names = ["S_miss", "T_miss", "S_hit", "T_hit", "S_shadow", "T_shadow", "rays", "hits"]
        return {"pixels": px, "ids": ids, "state": st, "tuv": tuv, "seconds": sec,
                "stats": {k: int(v) for k, v in zip(names, stats)}}

    def trace_rays(self, rays):
        rays = _f32(rays).reshape(-1, 8)
        ids = np.empty(rays.shape[0], np.uint32)
        tuv = np.empty((rays.shape[0], 3), np.float32)
        self.L.cref_trace_rays(self.h, _p(rays), rays.shape[0], _p(ids), _p(tuv))
        return ids, tuv


def threads(kind="strict"):
    return _lib(kind).cref_threads()


def slab(boxes, rays, kind="strict"):
    boxes, rays = _f32(boxes).reshape(-1, 6), _f32(rays).reshape(-1, 8)
    out = np.empty((rays.shape[0], 2), np.float32)
    _lib(kind).cref_slab(_p(boxes), _p(rays), rays.shape[0], _p(out))
    return out


def tri_intersect(tris, rays, kind="strict"):
    tris, rays = _f32(tris).reshape(-1, 9), _f32(rays).reshape(-1, 8)
    hit = np.empty(rays.shape[0], np.uint8)
    tuv = np.empty((rays.shape[0], 3), np.float32)
    _lib(kind).cref_tri_intersect(_p(tris), _p(rays), rays.shape[0], _p(hit), _p(tuv))
    return hit, tuv


def smooth_shading(inp, kind="strict"):
    inp = _f32(inp).reshape(-1, 17)
    out = np.empty((inp.shape[0], 3), np.float32)
    _lib(kind).cref_smooth_shading(_p(inp), inp.shape[0], _p(out))
    return out


def turntable(cam, sun, axis, n_frames, kind="strict"):
    out = np.empty((n_frames, 13), np.float32)
    _lib(kind).cref_turntable(_p(_f32(cam)), _p(_f32(sun)), _p(_f32(axis)), n_frames, _p(out))
    return out


def transform_triangles(a, v, tris12, kind="strict"):
    """The reference's transform_triangles (transform.hpp:120-129) with Transform{a, v}."""
    L = _lib(kind)
    out = _f32(tris12).reshape(-1, 12).copy()
    L.cref_transform_triangles.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
    L.cref_transform_triangles.restype = None
    L.cref_transform_triangles(_p(_f32(a).reshape(9)), _p(_f32(v).reshape(3)), _p(out), out.shape[0])
    return out


def rotate_triangles(axis, degrees, tris12, kind="strict"):
    L = _lib(kind)
    out = _f32(tris12).reshape(-1, 12).copy()
    L.cref_rotate_triangles.argtypes = [C.c_int, C.c_float, C.c_void_p, C.c_size_t]
    L.cref_rotate_triangles.restype = None
    L.cref_rotate_triangles(int(axis), float(degrees), _p(out), out.shape[0])
    return out


def fnv1a64(a):
    """FNV-1a-64 over the raw bytes (the id-hash SURVEY.md 8(c) records)."""
    data = np.ascontiguousarray(a).view(np.uint8).ravel()
    # vectorised FNV is awkward; do it in chunks with python ints on a reduced representation
    h = 0xcbf29ce484222325
    prime = 0x100000001b3
    mask = (1 << 64) - 1
    for b in data.tobytes():
        h = ((h ^ b) * prime) & mask
    return f"{h:016x}"

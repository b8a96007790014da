"""ctypes binding of oracle/libceres_oracle.so (oracle/ceres_oracle.c) -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

The C restatement of the reference's render loop.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module; the product package never does.
"""
import ctypes as C
import os
import subprocess
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libceres_oracle.so")
MISS = 0xFFFFFFFF
_L = None


def build(force=False):
    src = os.path.join(HERE, "ceres_oracle.c")
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", HERE, "port"], stdout=subprocess.DEVNULL)


def lib():
    global _L
    if _L is None:
        build()
        L = C.CDLL(LIB)
        vp, sz = C.c_void_p, C.c_size_t
        L.co_frame_setup.argtypes = [vp, sz, sz, vp]
        L.co_render.argtypes = [vp, sz, vp, vp, vp, sz, vp, vp, sz, sz, C.c_uint, vp, vp, vp, vp, vp]
        L.co_trace_rays.argtypes = [vp, sz, vp, vp, sz, vp, sz, vp, vp]
        L.co_slab.argtypes = [vp, vp, sz, vp]
        L.co_tri_intersect.argtypes = [vp, vp, sz, vp, vp]
        L.co_smooth_shading.argtypes = [vp, sz, vp]
        L.co_make_triangles.argtypes = [vp, sz, vp]
        L.co_rotate_triangles.argtypes = [C.c_int, C.c_float, vp, sz]
        L.co_transform_triangles.argtypes = [vp, vp, vp, sz]
        L.co_refit.argtypes = [vp, sz, vp, vp]
        L.co_threads.restype = C.c_int
        _L = L
    return _L


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _scene_args(pod):
    nodes = np.ascontiguousarray(pod["nodes"])
    idx = np.ascontiguousarray(pod["prim_indices"], dtype=np.uint64)
    tris = _f32(pod["triangles"])
    norms = _f32(pod["norms"])
    return nodes, idx, tris, norms


def frame_setup(cam, W, H):
    cam = _f32(cam)
    out = np.empty(9, np.float32)
    lib().co_frame_setup(_p(cam), W, H, _p(out))
    return out


def render(pod, cam, sun, W, H, flags=0, want_pixels=True):
    """pod = dict(nodes, prim_indices, triangles, norms) in the reference's POD layouts."""
    nodes, idx, tris, norms = _scene_args(pod)
    cam, sun = _f32(cam), _f32(sun)
    px = np.empty((H, W, 3), np.float32) if want_pixels else None
    ids = np.empty((H, W), np.uint32)
    st = np.empty((H, W), np.uint8)
    tuv = np.empty((H, W, 3), np.float32)
    stats = np.zeros(8, np.uint64)
    lib().co_render(_p(nodes), nodes.shape[0], _p(idx), _p(tris), _p(norms), tris.shape[0], _p(cam), _p(sun), W, H,
                    flags, _p(px), _p(ids), _p(st), _p(tuv), _p(stats))
    names = ["S_miss", "T_miss", "S_hit", "T_hit", "S_shadow", "T_shadow", "rays", "hits"]
    return {"pixels": px, "ids": ids, "state": st, "tuv": tuv, "stats": {k: int(v) for k, v in zip(names, stats)}}


def trace_rays(pod, rays):
    nodes, idx, tris, _ = _scene_args(pod)
    rays = _f32(rays).reshape(-1, 8)
    ids = np.empty(rays.shape[0], np.uint32)
    tuv = np.empty((rays.shape[0], 3), np.float32)
    lib().co_trace_rays(_p(nodes), nodes.shape[0], _p(idx), _p(tris), tris.shape[0], _p(rays), rays.shape[0], _p(ids), _p(tuv))
    return ids, tuv


def slab(boxes, rays):
    boxes, rays = _f32(boxes).reshape(-1, 6), _f32(rays).reshape(-1, 8)
    out = np.empty((rays.shape[0], 2), np.float32)
    lib().co_slab(_p(boxes), _p(rays), rays.shape[0], _p(out))
    return out


def tri_intersect(tris, rays):
    tris, rays = _f32(tris).reshape(-1, 9), _f32(rays).reshape(-1, 8)
    hit = np.empty(rays.shape[0], np.uint8)
    tuv = np.empty((rays.shape[0], 3), np.float32)
    lib().co_tri_intersect(_p(tris), _p(rays), rays.shape[0], _p(hit), _p(tuv))
    return hit, tuv


def smooth_shading(inp):
    inp = _f32(inp).reshape(-1, 17)
    out = np.empty((inp.shape[0], 3), np.float32)
    lib().co_smooth_shading(_p(inp), inp.shape[0], _p(out))
    return out


def make_triangles(verts):
    verts = _f32(verts).reshape(-1, 9)
    out = np.empty((verts.shape[0], 12), np.float32)
    lib().co_make_triangles(_p(verts), verts.shape[0], _p(out))
    return out


def rotate_triangles(axis, degrees, tris12):
    t = _f32(tris12).reshape(-1, 12).copy()
    lib().co_rotate_triangles(axis, degrees, _p(t), t.shape[0])
    return t


def transform_triangles(a, v, tris12):
    """transform.hpp:120-129 with Transform{a (3x3), v}; returns the new (n,12) triangles."""
    out = _f32(tris12).reshape(-1, 12).copy()
    lib().co_transform_triangles(_p(_f32(a).reshape(9)), _p(_f32(v).reshape(3)), _p(out), out.shape[0])
    return out


def refit(pod, tris12=None):
    """hierarchy_refitter.hpp refit of pod's BVH to tris12 (default: pod's triangles); returns the new nodes."""
    nodes = np.ascontiguousarray(pod["nodes"]).copy()
    idx = np.ascontiguousarray(pod["prim_indices"], dtype=np.uint64)
    tris = _f32(pod["triangles"] if tris12 is None else tris12).reshape(-1, 12)
    lib().co_refit(_p(nodes), nodes.shape[0], _p(idx), _p(tris))
    return nodes


def threads():
    return lib().co_threads()

"""ctypes bindings of the Scalar = double checkers -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

    port_*   oracle/libceres_oracle_f64.so       (oracle/ceres_oracle_f64.c, the plain-C restatement)
    ref_*    oracle/_ref/libceres_ref64_strict.so (oracle/ref_driver_f64.cpp: the reference's own headers, unchanged)

A double-precision scene is a dict of POD arrays in the reference's layouts: nodes NODE64_DT (64 bytes,
lib/bvh/bvh.hpp:20-28), prim_indices uint64, triangles float64[n, 12] ({p0, e1, e2, n}, lib/bvh/triangle.hpp:27),
norms float64[n, 9].  Only tests/ and __graft_entry__.smoke() may import this module.
"""
import ctypes as C
import os
import subprocess
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_LIB = os.path.join(HERE, "libceres_oracle_f64.so")
REF_LIB = os.path.join(HERE, "_ref", "libceres_ref64_strict.so")
MISS = 0xFFFFFFFF
NODE64_DT = np.dtype([("bounds", "<f8", (6,)), ("primitive_count", "<u8"), ("first_child_or_primitive", "<u8")])
assert NODE64_DT.itemsize == 64
_port = _ref = None


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def cam_array(eye, dir, up, fov):
    return _f64(list(eye) + list(dir) + list(up) + [fov])


def widen(pod):
    """The float scene `pod` (reference layouts for Scalar = float) with every value widened to double: a valid
    Scalar = double scene whose numbers happen to be binary32-representable."""
    n32 = np.ascontiguousarray(pod["nodes"])
    nodes = np.zeros(n32.shape[0], NODE64_DT)
    nodes["bounds"] = n32["bounds"]
    nodes["primitive_count"] = n32["primitive_count"]
    nodes["first_child_or_primitive"] = n32["first_child_or_primitive"]
    return {"nodes": nodes, "prim_indices": np.ascontiguousarray(pod["prim_indices"], np.uint64),
            "triangles": _f64(pod["triangles"]).reshape(-1, 12), "norms": _f64(pod["norms"]).reshape(-1, 9)}


def port_lib():
    global _port
    if _port is None:
        src = os.path.join(HERE, "ceres_oracle_f64.c")
        if not os.path.exists(PORT_LIB) or os.path.getmtime(PORT_LIB) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", HERE, "libceres_oracle_f64.so"], stdout=subprocess.DEVNULL)
        L = C.CDLL(PORT_LIB)
        vp, sz = C.c_void_p, C.c_size_t
        L.co64_frame_setup.argtypes = [vp, sz, sz, vp]
        L.co64_render.argtypes = [vp, vp, vp, vp, vp, vp, sz, sz, C.c_uint, vp, vp, vp, vp]
        _port = L
    return _port


def _outputs(W, H):
    return (np.zeros((H, W, 3), np.float64), np.zeros((H, W), np.uint32), np.zeros((H, W), np.uint8), np.zeros(2, np.uint64))


def port_render(pod, cam, sun, W, H, flags=0):
    """(pixels float64[H, W, 3], ids uint32[H, W], state uint8[H, W], (rays, hits)) by the C restatement."""
    px, ids, st, cnt = _outputs(W, H)
    nodes, idx = np.ascontiguousarray(pod["nodes"]), np.ascontiguousarray(pod["prim_indices"], np.uint64)
    tris, norms = _f64(pod["triangles"]), _f64(pod["norms"])
    port_lib().co64_render(_p(nodes), _p(idx), _p(tris), _p(norms), _p(_f64(cam)), _p(_f64(sun)), W, H, flags,
                           _p(px), _p(ids), _p(st), _p(cnt))
    return px, ids, st, (int(cnt[0]), int(cnt[1]))


def port_frame_setup(cam, W, H):
    out = np.zeros(9, np.float64)
    port_lib().co64_frame_setup(_p(_f64(cam)), W, H, _p(out))
    return out


def ref_available():
    return os.path.exists(REF_LIB)


def ref_lib():
    global _ref
    if _ref is None:
        L = C.CDLL(REF_LIB)
        vp, sz = C.c_void_p, C.c_size_t
        L.cref64_scene_load_obj.restype = vp
        L.cref64_scene_load_obj.argtypes = [C.c_char_p, C.c_int, C.c_double]
        L.cref64_scene_from_pod.restype = vp
        L.cref64_scene_from_pod.argtypes = [vp, sz, vp, vp, vp, sz]
        L.cref64_scene_destroy.argtypes = [vp]
        for n in ("cref64_scene_node_count", "cref64_scene_tri_count"):
            getattr(L, n).restype, getattr(L, n).argtypes = sz, [vp]
        for n in ("cref64_scene_nodes", "cref64_scene_prim_indices", "cref64_scene_triangles", "cref64_scene_norms"):
            getattr(L, n).restype, getattr(L, n).argtypes = vp, [vp]
        L.cref64_render.restype = C.c_double
        L.cref64_render.argtypes = [vp, vp, vp, vp, sz, sz, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.cref64_render_ex.argtypes = [vp, vp, vp, sz, sz, C.c_uint, vp, vp, vp, vp]
        L.cref64_turntable.argtypes = [vp, vp, vp, C.c_int, vp]
        _ref = L
    return _ref


def ref_turntable(cam, sun, axis, n_frames):
    """n_frames x 13 doubles (eye, dir, up, fov, sun): the camera path of anim.cpp:77-89 computed by the reference in double."""
    out = np.zeros((n_frames, 13), np.float64)
    ref_lib().cref64_turntable(_p(_f64(cam)), _p(_f64(sun)), _p(_f64(axis)), n_frames, _p(out))
    return out


class RefScene64:
    """A Scalar = double scene owned by the reference library."""

    def __init__(self, handle):
        if not handle:
            raise RuntimeError("reference could not load/build the scene")
        self.h, self.L = handle, ref_lib()

    @classmethod
    def load_obj(cls, path, rot_axis=-1, rot_deg=0.0):
        return cls(ref_lib().cref64_scene_load_obj(os.fsencode(path), rot_axis, rot_deg))

    @classmethod
    def from_pod(cls, pod):
        nodes, idx = np.ascontiguousarray(pod["nodes"]), np.ascontiguousarray(pod["prim_indices"], np.uint64)
        tris, norms = _f64(pod["triangles"]), _f64(pod["norms"])
        return cls(ref_lib().cref64_scene_from_pod(_p(nodes), nodes.shape[0], _p(idx), _p(tris), _p(norms), tris.shape[0]))

    def pod(self):
        L, h = self.L, self.h
        nn, nt = L.cref64_scene_node_count(h), L.cref64_scene_tri_count(h)

        def arr(ptr, dtype, count):
            return np.frombuffer((C.c_ubyte * (count * np.dtype(dtype).itemsize)).from_address(ptr), dtype=dtype).copy()
        return {"nodes": arr(L.cref64_scene_nodes(h), NODE64_DT, nn),
                "prim_indices": arr(L.cref64_scene_prim_indices(h), np.uint64, nt),
                "triangles": arr(L.cref64_scene_triangles(h), np.float64, nt * 12).reshape(nt, 12),
                "norms": arr(L.cref64_scene_norms(h), np.float64, nt * 9).reshape(nt, 9)}

    def render(self, cam, sun, W, H):
        """The unmodified render<double>(): (pixels, (rays, hits), seconds)."""
        px = np.zeros((H, W, 3), np.float64)
        rays, hits = C.c_int(0), C.c_int(0)
        sec = self.L.cref64_render(self.h, _p(_f64(cam)), _p(_f64(sun)), _p(px), W, H, C.byref(rays), C.byref(hits))
        return px, (rays.value, hits.value), sec

    def render_ex(self, cam, sun, W, H, flags=0):
        px, ids, st, cnt = _outputs(W, H)
        self.L.cref64_render_ex(self.h, _p(_f64(cam)), _p(_f64(sun)), W, H, flags, _p(px), _p(ids), _p(st), _p(cnt))
        return px, ids, st, (int(cnt[0]), int(cnt[1]))

    def close(self):
        if self.h:
            self.L.cref64_scene_destroy(self.h)
            self.h = None

    __del__ = close

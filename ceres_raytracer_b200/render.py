"""Host-side mirror of the reference's render interface (include/render.hpp) over the C ABI.

    Camera(eye, dir, up, fov)                                   render.hpp:16-22
    render(camera, sun_position, scene, pixels, width, height)  render.hpp:86-89 -> (rays, hits)

`scene` replaces the (bvh, triangles, tri_norms) triple: it is created once from the reference's POD arrays and
lives on the GPU (the reference also builds them once, outside the frame loop, anim.cpp:46-65).
numpy arrays are host buffers; torch CUDA tensors are used only as device memory / streams for the *_device calls.
"""
import ctypes as C
from dataclasses import dataclass
from typing import Sequence

import numpy as np

from . import _capi
from ._capi import NO_SHADOWS, COMPACT_ROWS, FRAME_RETAINED, PAYLOAD_NO_FLIP, MISS, STATIC_MAX_PIXELS, CeresError  # noqa: F401

NODE_DT = np.dtype([("bounds", "<f4", (6,)), ("primitive_count", "<u4"), ("first_child_or_primitive", "<u4")])
NODE64_DT = np.dtype([("bounds", "<f8", (6,)), ("primitive_count", "<u8"), ("first_child_or_primitive", "<u8")])   # Bvh<double>::Node


@dataclass
class Camera:
    """Camera<float> (render.hpp:16-22); fov in degrees."""
    eye: Sequence[float]
    dir: Sequence[float]
    up: Sequence[float]
    fov: float

    def c(self):
        cam = _capi.Camera()
        for k in range(3):
            cam.eye[k], cam.dir[k], cam.up[k] = float(self.eye[k]), float(self.dir[k]), float(self.up[k])
        cam.fov = float(self.fov)
        return cam

    def as_array(self):
        return np.asarray(list(self.eye) + list(self.dir) + list(self.up) + [self.fov], np.float32)


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _shard(shard):
    if shard is None:
        return None
    block_rows, index, count = shard
    return _capi.Shard(int(block_rows), int(index), int(count))


def _stream_handle(stream, device):
    """cudaStream_t for the C ABI.  Default: torch's current stream.  The C ABI reads NULL as "the scene's own
    stream", so the legacy default stream (handle 0) is passed as cudaStreamLegacy (0x1)."""
    if stream is None:
        import torch
        stream = torch.cuda.current_stream(device).cuda_stream
    return int(stream) or 1


def shard_rows(H, shard):
    """Frame rows owned by a shard (block_rows, index, count), in local order."""
    if shard is None or shard[2] <= 1:
        return np.arange(H)
    br, idx, cnt = shard
    rows = []
    b = idx
    while b * br < H:
        rows.extend(range(b * br, min((b + 1) * br, H)))
        b += cnt
    return np.asarray(rows, dtype=np.int64)


def frame_setup(camera, width, height):
    """dir, image_u, image_v of render.hpp:91-97."""
    out = np.empty(9, np.float32)
    _capi.check(_capi.lib().ceres_frame_setup(C.byref(camera.c()), width, height, _p(out)))
    return out[0:3].copy(), out[3:6].copy(), out[6:9].copy()


def probe_read_bandwidth(working_set_bytes, repeats=20, device=0):
    """GB/s of the library's 128-bit read kernel over a working set of this size (ceres_probe_read_bandwidth)."""
    out = C.c_double(0.0)
    _capi.check(_capi.lib().ceres_probe_read_bandwidth(int(device), int(working_set_bytes), int(repeats), C.byref(out)))
    return float(out.value)


def probe_host_fill_bandwidth(nbytes=256 << 20, repeats=3):
    """GB/s of the library's host-fill pool into page-locked memory (ceres_probe_host_fill_bandwidth)."""
    out = C.c_double(0.0)
    _capi.check(_capi.lib().ceres_probe_host_fill_bandwidth(int(nbytes), int(repeats), C.byref(out)))
    return float(out.value)


def build_options(algorithm=0, radius=0, max_leaf=0, triangle_cost=0.0):
    """Which tree build_lbvh / Scene.from_lbvh make: algorithm 0 = 2 = top-down binned SAH (default), 3 = locally-ordered
    clustering, 1 = radix tree of the Morton codes; radius = clustering neighbours per side (0 = 8); max_leaf = SAH leaf collapsing
    (0 = 8, 1 = off); triangle_cost = cost of a triangle test in node steps (0 = 1.5).  Process-wide (ceres_build_options)."""
    _capi.check(_capi.lib().ceres_build_options(int(algorithm), int(radius), int(max_leaf), float(triangle_cost)))


def build_lbvh(triangles, device=0):
    """GPU linear-BVH build (ceres_build_lbvh): returns (nodes NODE_DT[2n-1], prim_indices uint64[n], build_ms) in the
    reference's layout for `triangles` float32[n, 12]."""
    tris = np.ascontiguousarray(triangles, np.float32).reshape(-1, 12)
    n = tris.shape[0]
    nodes = np.zeros(max(2 * n - 1, 1), NODE_DT)
    idx = np.zeros(max(n, 1), np.uint64)[:n]
    ms, count = C.c_double(0.0), C.c_size_t(0)
    _capi.check(_capi.lib().ceres_build_bvh(_p(tris), n, int(device), _p(nodes), C.byref(count), _p(idx), C.byref(ms)))
    return nodes[:count.value].copy(), idx, float(ms.value)


class Scene:
    """Flattened BVH + triangles + normals resident on one GPU (ceres_scene)."""

    def __init__(self, nodes, prim_indices, triangles, tri_norms=None, device=0):
        L = _capi.lib()
        nodes = np.ascontiguousarray(nodes)
        if nodes.dtype.itemsize != 32:
            nodes = np.ascontiguousarray(nodes).view(np.uint8).reshape(-1, 32)
        idx = np.ascontiguousarray(prim_indices, dtype=np.uint64)
        tris = np.ascontiguousarray(triangles, dtype=np.float32).reshape(-1, 12)
        norms = np.ascontiguousarray(tri_norms, dtype=np.float32).reshape(-1, 9) if tri_norms is not None else None
        if idx.shape[0] != tris.shape[0] or (norms is not None and norms.shape[0] != tris.shape[0]):
            raise ValueError("prim_indices / triangles / tri_norms length mismatch")
        h = C.c_void_p()
        _capi.check(L.ceres_scene_create(_p(nodes), nodes.shape[0], _p(idx), _p(tris), _p(norms), tris.shape[0],
                                         int(device), C.byref(h)))
        self._h, self._L = h, L
        self.device = int(device)
        self.node_count, self.tri_count = int(nodes.shape[0]), int(tris.shape[0])

    @classmethod
    def from_lbvh(cls, triangles, tri_norms=None, device=0):
        """Scene over a linear BVH built AND kept on the GPU (ceres_scene_create_lbvh); `build_ms` = device build time."""
        L = _capi.lib()
        tris = np.ascontiguousarray(triangles, dtype=np.float32).reshape(-1, 12)
        norms = np.ascontiguousarray(tri_norms, dtype=np.float32).reshape(-1, 9) if tri_norms is not None else None
        if norms is not None and norms.shape[0] != tris.shape[0]:
            raise ValueError("triangles / tri_norms length mismatch")
        h, ms = C.c_void_p(), C.c_double(0.0)
        _capi.check(L.ceres_scene_create_lbvh(_p(tris), _p(norms), tris.shape[0], int(device), C.byref(h), C.byref(ms)))
        self = cls.__new__(cls)
        self._h, self._L, self.device = h, L, int(device)
        self.tri_count = int(tris.shape[0])
        self.node_count = int(L.ceres_scene_node_count(h))
        self.build_ms = float(ms.value)
        return self

    @classmethod
    def from_pod(cls, pod, device=0):
        return cls(pod["nodes"], pod["prim_indices"], pod["triangles"], pod.get("norms"), device)

    def close(self):
        if getattr(self, "_h", None):
            self._L.ceres_scene_destroy(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def device_bytes(self):
        return int(self._L.ceres_scene_device_bytes(self._h))

    @property
    def last_d2h_bytes(self):
        """Bytes the last host-frame render moved over PCIe (the scene's screen rectangle + counters)."""
        return int(self._L.ceres_scene_last_d2h_bytes(self._h))

    @property
    def stack_depth(self):
        return int(self._L.ceres_scene_stack_depth(self._h))

    @property
    def last_rectangle(self):
        """(i0, i1, l0, l1): columns x local rows of the conservative screen rectangle traced for the last frame."""
        out = np.zeros(4, np.uint32)
        _capi.check(self._L.ceres_scene_last_rectangle(self._h, _p(out)))
        return tuple(int(x) for x in out)

    def set_tuning(self, variant=0, primary_block=0, shadow_block=0, ctas_per_sm=0, shadow_order=0, refill_min=0, subframes=0, chunk=0,
                   share=0, share_period=0, share_min_idle=0):
        """variant 0 = the frame kernel; 1-6 = the wave pipelines (experiments build only).  shadow_order: 0 default,
        1 nearer child first, 2 left child first.  subframes = row bands per frame.  share (frame kernel): shadow-subtree
        sharing between the lanes of a warp, 0 default / 1 off / 2 on, with its period (steps) and idle-lane threshold."""
        if variant == 0:
            refill_min, chunk = share, (share_period & 0xFF) | ((share_min_idle & 0xFF) << 8)
        t = _capi.Tuning(variant, primary_block, shadow_block, ctas_per_sm, shadow_order, (C.c_uint32 * 3)(refill_min, subframes, chunk))
        _capi.check(self._L.ceres_scene_set_tuning(self._h, C.byref(t)))
        self._variant = int(variant)

    # ---- scene preparation on the device ---------------------------------------------------
    def rotate(self, axis, degrees):
        """rotate_triangles<axis>(degrees) of render.hpp:24-44 on the resident triangles (call refit() afterwards)."""
        _capi.check(self._L.ceres_scene_rotate(self._h, int(axis), float(degrees)))

    def transform(self, a, v):
        """transform_triangles(Transform{a, v}) of transform.hpp:106-129 on the resident triangles."""
        a = np.ascontiguousarray(a, np.float32).reshape(9)
        v = np.ascontiguousarray(v, np.float32).reshape(3)
        _capi.check(self._L.ceres_scene_transform(self._h, _p(a), _p(v)))

    def refit(self):
        """HierarchyRefitter::refit (hierarchy_refitter.hpp) of the resident BVH to the resident triangles."""
        _capi.check(self._L.ceres_scene_refit(self._h))

    def download(self):
        """(nodes, triangles) as the reference would hold them now: NODE_DT[node_count], float32[tri_count, 12]."""
        nodes = np.zeros(self.node_count, NODE_DT)
        tris = np.zeros((self.tri_count, 12), np.float32)
        _capi.check(self._L.ceres_scene_download(self._h, _p(nodes), self.node_count, _p(tris), self.tri_count))
        return nodes, tris

    def set_stats(self, enable):
        _capi.check(self._L.ceres_scene_set_stats(self._h, int(bool(enable))))

    def get_stats(self):
        out = np.zeros(8, np.uint64)
        _capi.check(self._L.ceres_scene_get_stats(self._h, _p(out)))
        names = ["S_miss", "T_miss", "S_hit", "T_hit", "S_shadow", "T_shadow", "rays", "hits"]
        return {k: int(v) for k, v in zip(names, out)}

    def step_histogram(self):
        """(primary[64], shadow[64]) counts of rays by node-pair steps // 4 since the last call (statistics mode)."""
        out = np.zeros(128, np.uint64)
        _capi.check(self._L.ceres_scene_get_step_histogram(self._h, _p(out)))
        return out[:64].copy(), out[64:].copy()

    def frame_error(self):
        """Device-side error bits of the frames enqueued so far (synchronises; 0 = none; include/ceres_b200.h)."""
        w = C.c_uint64(0)
        _capi.check(self._L.ceres_scene_frame_error(self._h, C.byref(w)))
        return int(w.value)

    def set_timeline(self, enable):
        _capi.check(self._L.ceres_scene_set_timeline(self._h, int(bool(enable))))

    def timeline(self):
        """uint64 [warps, 12] per warp of the last frame kernel (set_timeline(True)); columns: include/ceres_b200.h."""
        out = np.zeros((16 * 4 * 256, 12), np.uint64)
        n = C.c_size_t(0)
        _capi.check(self._L.ceres_scene_get_timeline(self._h, _p(out), out.shape[0], C.byref(n)))
        out = out[:n.value]
        return out[out[:, 2] != 0]

    def set_profiling(self, enable):
        """Per-kernel CUDA-event timing of every render_device call (see kernel_ms)."""
        _capi.check(self._L.ceres_scene_set_profiling(self._h, int(bool(enable))))

    def kernel_ms(self):
        """Sums over the frames recorded since set_profiling(True): ({clear, frame, total} ms, frames) for the frame
        kernel; ({primary, shadow, shade, total} ms, frames) for the wave pipelines of the experiments build."""
        ms = np.zeros(4, np.float64)
        n = C.c_uint64(0)
        _capi.check(self._L.ceres_scene_get_kernel_ms(self._h, _p(ms), C.byref(n)))
        if getattr(self, "_variant", 0) == 0:
            return {"clear": float(ms[0]), "frame": float(ms[1]), "total": float(ms[3])}, int(n.value)
        return dict(zip(["primary", "shadow", "shade", "total"], ms.tolist())), int(n.value)

    # ---- host buffers ---------------------------------------------------------------------------------------
    def render(self, camera, sun_position, width, height, flags=0, pixels=None, want_ids=False, shard=None, ids=None):
        """Returns (pixels[H,W,3] float32, ids[H,W] uint32 or None, (rays, hits)).  flags may carry FRAME_RETAINED (the same
        pixels / ids buffers as in the previous call, untouched since: include/ceres_b200.h)."""
        rows = len(shard_rows(height, shard)) if (shard is not None and flags & COMPACT_ROWS) else height
        if pixels is None:
            pixels = np.zeros((rows, width, 3), np.float32) if shard is not None else np.empty((rows, width, 3), np.float32)
        if pixels.dtype != np.float32 or not pixels.flags.c_contiguous or pixels.size != rows * width * 3:
            raise ValueError("pixels must be a C-contiguous float32 buffer of 3*width*height")
        if ids is None:
            ids = (np.full((rows, width), MISS, np.uint32) if shard is not None else np.empty((rows, width), np.uint32)) if want_ids else None
        elif ids.dtype != np.uint32 or not ids.flags.c_contiguous or ids.size != rows * width:
            raise ValueError("ids must be a C-contiguous uint32 buffer of width*height")
        sun = np.asarray(sun_position, np.float32)
        cnt = np.zeros(2, np.uint64)
        sh = _shard(shard)
        _capi.check(self._L.ceres_render_shard(self._h, C.byref(camera.c()), _p(sun), width, height, flags,
                                               C.byref(sh) if sh is not None else None, _p(pixels), _p(ids), _p(cnt)))
        return pixels, ids, (int(cnt[0]), int(cnt[1]))

    def render_ppm8(self, camera, sun_position, width, height, flags=0, out=None, shard=None):
        """The frame as P6 payload bytes (static.cpp:135-147), quantised and flipped on the device: (uint8[H,W,3], (rays, hits)).
        shard = (block_rows, index, count): only that shard's rows of the full-size payload are written."""
        if out is None:
            out = np.empty((height, width, 3), np.uint8) if shard is None else np.zeros((height, width, 3), np.uint8)
        sun = np.asarray(sun_position, np.float32)
        cnt = np.zeros(2, np.uint64)
        sh = _shard(shard)
        _capi.check(self._L.ceres_render_ppm8_shard(self._h, C.byref(camera.c()), _p(sun), width, height, flags,
                                                    C.byref(sh) if sh is not None else None, _p(out), _p(cnt)))
        return out, (int(cnt[0]), int(cnt[1]))

    def render_ppm16(self, camera, sun_position, width, height, flags=0, out=None, shard=None):
        """The frame as 16-bit P6 payload (maxval 65535, big-endian; the x 65535 scaling of anim.cpp:117-119), quantised and
        flipped on the device: (uint16 big-endian [H,W,3], (rays, hits))."""
        if out is None:
            out = np.empty((height, width, 3), ">u2") if shard is None else np.zeros((height, width, 3), ">u2")
        sun = np.asarray(sun_position, np.float32)
        cnt = np.zeros(2, np.uint64)
        sh = _shard(shard)
        _capi.check(self._L.ceres_render_ppm16_shard(self._h, C.byref(camera.c()), _p(sun), width, height, flags,
                                                     C.byref(sh) if sh is not None else None, _p(out), _p(cnt)))
        return out, (int(cnt[0]), int(cnt[1]))

    def render_sequence(self, cams, width, height, flags=0, pixels_all=None):
        """cams: (n,13) float32 rows of eye,dir,up,fov,sun.  Returns (pixels[n,H,W,3], counters[n,2])."""
        cams = np.ascontiguousarray(cams, np.float32).reshape(-1, 13)
        n = cams.shape[0]
        if pixels_all is None:
            pixels_all = np.empty((n, height, width, 3), np.float32)
        cnt = np.zeros((n, 2), np.uint64)
        _capi.check(self._L.ceres_render_sequence(self._h, n, _p(cams), width, height, flags, _p(pixels_all), _p(cnt)))
        return pixels_all, cnt

    def render_sequence_payload(self, cams, width, height, flags=0, bytes_per_channel=1, out=None):
        """The sequence as quantised P6 payloads: (uint8[n,H,W,3] or big-endian uint16[n,H,W,3], counters[n,2])."""
        cams = np.ascontiguousarray(cams, np.float32).reshape(-1, 13)
        n = cams.shape[0]
        if out is None:
            out = np.empty((n, height, width, 3), np.uint8 if bytes_per_channel == 1 else ">u2")
        cnt = np.zeros((n, 2), np.uint64)
        _capi.check(self._L.ceres_render_sequence_payload(self._h, n, _p(cams), width, height, flags, int(bytes_per_channel), _p(out), _p(cnt)))
        return out, cnt

    def trace_rays(self, rays):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        ids = np.empty(rays.shape[0], np.uint32)
        tuv = np.empty((rays.shape[0], 3), np.float32)
        _capi.check(self._L.ceres_trace_rays(self._h, _p(rays), rays.shape[0], _p(ids), _p(tuv)))
        return ids, tuv

    # ---- device buffers (torch tensors as plain device memory) -----------------------------------------------
    def render_device(self, camera, sun_position, width, height, flags=0, shard=None, pixels=None, ids=None,
                      counters=None, stream=None):
        """Enqueue one frame on `stream` (int cudaStream_t; default torch's current stream).  pixels: float32 CUDA
        tensor with 3*rows*width elements; ids: int32/uint32 CUDA tensor or None; counters: int64 CUDA tensor[2]."""
        stream = _stream_handle(stream, self.device)
        sun = np.asarray(sun_position, np.float32)
        sh = _shard(shard)
        _capi.check(self._L.ceres_render_device(
            self._h, C.byref(camera.c()), _p(sun), width, height, flags, C.byref(sh) if sh is not None else None,
            C.c_void_p(stream), C.c_void_p(pixels.data_ptr()), C.c_void_p(ids.data_ptr()) if ids is not None else None,
            C.c_void_p(counters.data_ptr()) if counters is not None else None))

    def render_device_static(self, camera, sun_position, width, height, flags=0, pixels=None, ids=None, counters=None,
                             stream=None):
        """render_device over the library's static device scratch: the call allocates nothing (ceres_render_device_static,
        include/ceres_b200.h; no counterpart in the reference).  width * height <= STATIC_MAX_PIXELS, whole frames only;
        all calls on a device share the scratch, so calls that overlap on the GPU race."""
        stream = _stream_handle(stream, self.device)
        sun = np.asarray(sun_position, np.float32)
        _capi.check(self._L.ceres_render_device_static(
            self._h, C.byref(camera.c()), _p(sun), width, height, flags,
            C.c_void_p(stream), C.c_void_p(pixels.data_ptr()), C.c_void_p(ids.data_ptr()) if ids is not None else None,
            C.c_void_p(counters.data_ptr()) if counters is not None else None))

    def quantize_device(self, pixels, width, height, out, stream=None):
        stream = _stream_handle(stream, self.device)
        _capi.check(self._L.ceres_quantize_device(self._h, C.c_void_p(pixels.data_ptr()), width, height,
                                                  C.c_void_p(stream), C.c_void_p(out.data_ptr())))


def render(camera, sun_position, scene, pixels, width, height, flags=0):
    """Drop-in for render.hpp:86-156: fills `pixels` (float32, 3*width*height, index 3*(width*j+i)) and returns
    the pair (rays, hits) of render.hpp:155."""
    _, _, counters = scene.render(camera, sun_position, width, height, flags=flags, pixels=pixels)
    return counters


class PinnedBuffer:
    """Page-locked host memory from ceres_host_alloc, exposed as a numpy array (`.array`).  A framebuffer living
    here is filled by DMA directly from the device."""

    def __init__(self, shape, dtype=np.float32):
        L = _capi.lib()
        self._L = L
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self._ptr = L.ceres_host_alloc(n)
        if not self._ptr:
            raise CeresError(_capi.ERR_CUDA, L.ceres_last_error().decode(errors="replace"))
        buf = (C.c_char * n).from_address(self._ptr)
        self.array = np.frombuffer(buf, dtype=dtype).reshape(shape)

    def close(self):
        if getattr(self, "_ptr", None):
            self.array = None
            self._L.ceres_host_free(self._ptr)
            self._ptr = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def kernel_launch_count():
    return int(_capi.lib().ceres_kernel_launch_count())


def device_count():
    return int(_capi.lib().ceres_device_count())


def write_ppm(path, pixels):
    """static.cpp:135-147: P6, rows from j=height-1 down to 0, uint8(max(min(p*255,255),0))."""
    H, W, _ = pixels.shape
    x = pixels.astype(np.float32) * np.float32(255)
    x = np.where(np.float32(255) < x, np.float32(255), x)
    x = np.where(x < 0, np.float32(0), x)
    with np.errstate(invalid="ignore"):
        b = np.nan_to_num(x, nan=0.0).astype(np.uint8)[::-1]
    with open(path, "wb") as f:
        f.write(f"P6 {W} {H} 255\n".encode())
        f.write(b.tobytes())
    return b


class SceneF64:
    """A Scalar = double scene on one GPU (ceres_scene_f64; `anim -d`, anim.cpp:154-155).  nodes: NODE64_DT or raw 64-byte
    records, triangles float64[n, 12], tri_norms float64[n, 9] or None."""

    def __init__(self, nodes, prim_indices, triangles, tri_norms=None, device=0):
        L = _capi.lib()
        nodes = np.ascontiguousarray(nodes)
        if nodes.dtype.itemsize != 64:
            raise ValueError("double-precision nodes are 64-byte records (bvh.hpp:20-28)")
        idx = np.ascontiguousarray(prim_indices, dtype=np.uint64)
        tris = np.ascontiguousarray(triangles, dtype=np.float64).reshape(-1, 12)
        norms = np.ascontiguousarray(tri_norms, dtype=np.float64).reshape(-1, 9) if tri_norms is not None else None
        if idx.shape[0] != tris.shape[0] or (norms is not None and norms.shape[0] != tris.shape[0]):
            raise ValueError("prim_indices / triangles / tri_norms length mismatch")
        h = C.c_void_p()
        _capi.check(L.ceres_scene_f64_create(_p(nodes), nodes.shape[0], _p(idx), _p(tris), _p(norms), tris.shape[0], int(device), C.byref(h)))
        self._h, self._L, self.device = h, L, int(device)
        self.kernel_ms = 0.0

    @classmethod
    def from_pod(cls, pod, device=0):
        return cls(pod["nodes"], pod["prim_indices"], pod["triangles"], pod.get("norms"), device=device)

    def render(self, camera, sun_position, width, height, flags=0, want_ids=False, pixels=None, shard=None, ids=None):
        """Returns (pixels float64[H, W, 3], ids uint32[H, W] or None, (rays, hits)); camera: Camera or 10 numbers.  pixels: an
        existing C-contiguous float64 frame to fill (the frame loop of anim.cpp:76-105 reuses one).  shard = (block_rows,
        index, count): only that shard's rows of the full-size pixels / ids are written and the counters are the shard's
        (ceres_render_f64_shard); pass the same pixels / ids to every shard's call to assemble the frame."""
        cam = _capi.CameraF64()
        vals = (list(camera.eye) + list(camera.dir) + list(camera.up) + [camera.fov]) if hasattr(camera, "eye") else [float(x) for x in camera]
        for k in range(3):
            cam.eye[k], cam.dir[k], cam.up[k] = vals[k], vals[3 + k], vals[6 + k]
        cam.fov = vals[9]
        if pixels is None:
            pixels = np.empty((height, width, 3), np.float64)
        elif pixels.dtype != np.float64 or not pixels.flags.c_contiguous or pixels.size != 3 * width * height:
            raise ValueError("pixels must be a C-contiguous float64 buffer of 3*width*height")
        if ids is not None and (ids.dtype != np.uint32 or not ids.flags.c_contiguous or ids.size != width * height):
            raise ValueError("ids must be a C-contiguous uint32 buffer of width*height")
        if ids is None and want_ids:
            ids = np.empty((height, width), np.uint32)
        sun = np.ascontiguousarray(sun_position, np.float64)
        cnt, ms = np.zeros(2, np.uint64), C.c_double(0.0)
        sh = _shard(shard)
        if sh is None:
            _capi.check(self._L.ceres_render_f64(self._h, C.byref(cam), _p(sun), width, height, flags, _p(pixels), _p(ids), _p(cnt), C.byref(ms)))
        else:
            _capi.check(self._L.ceres_render_f64_shard(self._h, C.byref(cam), _p(sun), width, height, flags, C.byref(sh), _p(pixels), _p(ids), _p(cnt), C.byref(ms)))
        self.kernel_ms = float(ms.value)
        return pixels, ids, (int(cnt[0]), int(cnt[1]))

    def close(self):
        if getattr(self, "_h", None):
            self._L.ceres_scene_f64_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    __del__ = close

"""ctypes binding of libceres_b200.so (include/ceres_b200.h).  No fallback: if the CUDA library is missing and
cannot be built, importing the entry points raises."""
import ctypes as C
import os

from . import build as _build

HERE = os.path.dirname(os.path.abspath(__file__))

OK, ERR_INVALID, ERR_CUDA, ERR_NO_DEVICE, ERR_UNSUPPORTED = 0, 1, 2, 3, 4
NO_SHADOWS, COMPACT_ROWS, FRAME_RETAINED, PAYLOAD_NO_FLIP = 1, 2, 4, 8
MISS = 0xFFFFFFFF
STATIC_MAX_PIXELS = 1 << 20     # CERES_STATIC_MAX_PIXELS (ceres_render_device_static)


class CeresError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"ceres_b200 error {code}: {msg}")
        self.code = code


class Camera(C.Structure):
    """ceres_camera = Camera<float> of include/render.hpp:16-22."""
    _fields_ = [("eye", C.c_float * 3), ("dir", C.c_float * 3), ("up", C.c_float * 3), ("fov", C.c_float)]


class CameraF64(C.Structure):
    """ceres_camera_f64 = Camera<double> of include/render.hpp:16-22."""
    _fields_ = [("eye", C.c_double * 3), ("dir", C.c_double * 3), ("up", C.c_double * 3), ("fov", C.c_double)]


class Shard(C.Structure):
    _fields_ = [("block_rows", C.c_uint32), ("index", C.c_uint32), ("count", C.c_uint32)]


class Tuning(C.Structure):
    _fields_ = [("variant", C.c_uint32), ("primary_block", C.c_uint32), ("shadow_block", C.c_uint32),
                ("ctas_per_sm", C.c_uint32), ("shadow_order", C.c_uint32), ("reserved", C.c_uint32 * 3)]


# every symbol include/ceres_b200.h declares (tests/test_capi_symbols.py checks the header against this list)
SYMBOLS = [
    "ceres_scene_create", "ceres_scene_destroy", "ceres_scene_set_tuning", "ceres_scene_device_bytes",
    "ceres_render", "ceres_render_shard", "ceres_render_multi", "ceres_render_sequence", "ceres_render_device",
    "ceres_scene_set_stats", "ceres_scene_get_stats", "ceres_trace_rays", "ceres_frame_setup",
    "ceres_scene_set_profiling", "ceres_scene_get_kernel_ms", "ceres_host_register", "ceres_host_unregister",
    "ceres_host_is_pinned", "ceres_quantize_device", "ceres_host_alloc", "ceres_host_free", "ceres_abi_version", "ceres_device_count",
    "ceres_kernel_launch_count", "ceres_last_error",
    "ceres_peer_frame_create", "ceres_peer_frame_open", "ceres_peer_frame_alias", "ceres_peer_frame_destroy",
    "ceres_peer_frame_pixels", "ceres_peer_frame_ids", "ceres_render_peer", "ceres_peer_frame_error",
    "ceres_peer_frame_invalidate",
]
SYMBOLS += ["ceres_render_ppm8", "ceres_render_ppm16", "ceres_probe_read_bandwidth", "ceres_build_lbvh", "ceres_scene_create_lbvh", "ceres_scene_f64_create", "ceres_scene_f64_destroy", "ceres_render_f64", "ceres_scene_get_step_histogram", "ceres_scene_last_d2h_bytes", "ceres_scene_rotate", "ceres_scene_transform", "ceres_scene_refit", "ceres_scene_download",
            "ceres_scene_last_rectangle", "ceres_probe_host_fill_bandwidth", "ceres_scene_get_timeline", "ceres_scene_set_timeline", "ceres_scene_frame_error", "ceres_build_is_checked", "ceres_render_ppm8_shard", "ceres_render_ppm16_shard", "ceres_build_options", "ceres_scene_stack_depth", "ceres_build_bvh", "ceres_scene_node_count", "ceres_render_sequence_payload", "ceres_render_device_static", "ceres_render_f64_shard"]
PEER_HANDLE_BYTES = 64

_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    # CERES_B200_EXPERIMENTS=1 (profiles/kbench.py, tests/test_gpu_experiments.py): the -DCERES_EXPERIMENTS build
    experiments = os.environ.get("CERES_B200_EXPERIMENTS", "") not in ("", "0")
    path = _build.LIB_EXP if experiments else _build.LIB
    # development aid (profiles/kbench.py A/B of compile-time options): a library built by `build.py --variant NAME -D...`
    variant = os.environ.get("CERES_B200_VARIANT", "")
    if variant:
        path = _build.variant_path(variant)
        if not os.path.exists(path):
            raise RuntimeError(f"{path} is missing: build it with python -m ceres_raytracer_b200.build --variant {variant} [nvcc flags]")
    elif _build.needs_build(path):
        try:
            _build.build(experiments=experiments)
        except Exception as e:  # no nvcc on the box and no prebuilt library: fail loudly, never fall back
            if not os.path.exists(path):
                raise RuntimeError(f"{os.path.basename(path)} is missing and could not be built ({e}); "
                                   "ceres_raytracer_b200 has no CPU fallback") from e
    L = C.CDLL(path)
    vp, sz, u32, u64 = C.c_void_p, C.c_size_t, C.c_uint32, C.c_uint64
    camp, shp = C.POINTER(Camera), C.POINTER(Shard)
    L.ceres_scene_create.argtypes = [vp, sz, vp, vp, vp, sz, C.c_int, C.POINTER(vp)]
    L.ceres_scene_destroy.argtypes = [vp]
    L.ceres_scene_destroy.restype = None
    L.ceres_scene_set_tuning.argtypes = [vp, C.POINTER(Tuning)]
    L.ceres_scene_device_bytes.argtypes = [vp]
    L.ceres_scene_device_bytes.restype = sz
    L.ceres_render.argtypes = [vp, camp, vp, sz, sz, C.c_uint, vp, vp, vp]
    L.ceres_render_shard.argtypes = [vp, camp, vp, sz, sz, C.c_uint, shp, vp, vp, vp]
    L.ceres_render_multi.argtypes = [C.POINTER(vp), C.c_int, camp, vp, sz, sz, C.c_uint, vp, vp, vp]
    L.ceres_render_sequence.argtypes = [vp, sz, vp, sz, sz, C.c_uint, vp, vp]
    L.ceres_render_device.argtypes = [vp, camp, vp, sz, sz, C.c_uint, shp, vp, vp, vp, vp]
    L.ceres_render_device_static.argtypes = [vp, camp, vp, sz, sz, C.c_uint, vp, vp, vp, vp]
    L.ceres_scene_set_stats.argtypes = [vp, C.c_int]
    L.ceres_scene_get_stats.argtypes = [vp, vp]
    L.ceres_trace_rays.argtypes = [vp, vp, sz, vp, vp]
    L.ceres_frame_setup.argtypes = [camp, sz, sz, vp]
    L.ceres_quantize_device.argtypes = [vp, vp, sz, sz, vp, vp]
    L.ceres_host_alloc.argtypes = [sz]
    L.ceres_host_alloc.restype = vp
    L.ceres_host_free.argtypes = [vp]
    L.ceres_host_free.restype = None
    L.ceres_host_register.argtypes = [vp, sz]
    L.ceres_host_unregister.argtypes = [vp]
    L.ceres_host_is_pinned.argtypes = [vp]
    L.ceres_scene_set_profiling.argtypes = [vp, C.c_int]
    L.ceres_scene_get_kernel_ms.argtypes = [vp, vp, vp]
    L.ceres_peer_frame_create.argtypes = [vp, sz, sz, C.c_int, C.c_int, C.POINTER(vp), vp]
    L.ceres_peer_frame_open.argtypes = [vp, vp, sz, sz, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]
    L.ceres_peer_frame_alias.argtypes = [vp, vp, C.c_int, C.POINTER(vp)]
    L.ceres_peer_frame_destroy.argtypes = [vp]
    L.ceres_peer_frame_destroy.restype = None
    L.ceres_peer_frame_pixels.argtypes = [vp]
    L.ceres_peer_frame_pixels.restype = vp
    L.ceres_peer_frame_ids.argtypes = [vp]
    L.ceres_peer_frame_ids.restype = vp
    L.ceres_render_peer.argtypes = [vp, camp, vp, C.c_uint, vp, vp, vp]
    L.ceres_peer_frame_error.argtypes = [vp]
    L.ceres_peer_frame_invalidate.argtypes = [vp]
    L.ceres_render_ppm8.argtypes = [vp, camp, vp, sz, sz, C.c_uint, vp, vp]
    L.ceres_render_ppm16.argtypes = [vp, camp, vp, sz, sz, C.c_uint, vp, vp]
    L.ceres_render_ppm8_shard.argtypes = [vp, camp, vp, sz, sz, C.c_uint, shp, vp, vp]
    L.ceres_render_ppm16_shard.argtypes = [vp, camp, vp, sz, sz, C.c_uint, shp, vp, vp]
    L.ceres_probe_read_bandwidth.argtypes = [C.c_int, sz, C.c_int, vp]
    L.ceres_build_lbvh.argtypes = [vp, sz, C.c_int, vp, vp, vp]
    L.ceres_scene_create_lbvh.argtypes = [vp, vp, sz, C.c_int, vp, vp]
    L.ceres_scene_f64_create.argtypes = [vp, sz, vp, vp, vp, sz, C.c_int, vp]
    L.ceres_scene_f64_destroy.argtypes = [vp]
    L.ceres_scene_f64_destroy.restype = None
    L.ceres_render_f64.argtypes = [vp, vp, vp, sz, sz, C.c_uint, vp, vp, vp, vp]
    L.ceres_render_f64_shard.argtypes = [vp, vp, vp, sz, sz, C.c_uint, shp, vp, vp, vp, vp]
    L.ceres_scene_get_step_histogram.argtypes = [vp, vp]
    L.ceres_scene_last_d2h_bytes.argtypes = [vp]
    L.ceres_scene_last_d2h_bytes.restype = u64
    L.ceres_scene_rotate.argtypes = [vp, C.c_int, C.c_float]
    L.ceres_scene_transform.argtypes = [vp, vp, vp]
    L.ceres_scene_refit.argtypes = [vp]
    L.ceres_scene_download.argtypes = [vp, vp, sz, vp, sz]
    L.ceres_scene_last_rectangle.argtypes = [vp, vp]
    L.ceres_probe_host_fill_bandwidth.argtypes = [sz, C.c_int, vp]
    L.ceres_scene_get_timeline.argtypes = [vp, vp, sz, vp]
    L.ceres_scene_set_timeline.argtypes = [vp, C.c_int]
    L.ceres_scene_frame_error.argtypes = [vp, vp]
    L.ceres_build_options.argtypes = [C.c_int, C.c_int, C.c_int, C.c_float]
    L.ceres_build_bvh.argtypes = [vp, sz, C.c_int, vp, vp, vp, vp]
    L.ceres_scene_stack_depth.argtypes = [vp]
    L.ceres_scene_stack_depth.restype = C.c_uint32
    L.ceres_scene_node_count.argtypes = [vp]
    L.ceres_render_sequence_payload.argtypes = [vp, sz, vp, sz, sz, C.c_uint, C.c_int, vp, vp]
    L.ceres_scene_node_count.restype = sz
    L.ceres_build_is_checked.restype = C.c_int
    L.ceres_abi_version.restype = C.c_int
    L.ceres_device_count.restype = C.c_int
    L.ceres_kernel_launch_count.restype = u64
    L.ceres_last_error.restype = C.c_char_p
    _lib = L
    return L


def check(rc):
    if rc != OK:
        raise CeresError(rc, lib().ceres_last_error().decode(errors="replace"))

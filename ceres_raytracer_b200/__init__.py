"""ceres_raytracer_b200 -- the per-pixel render loop of iracigt/ceres-raytracer (include/render.hpp:86-156) as
hand-written sm_100a CUDA behind the C ABI of include/ceres_b200.h.  No CPU fallback: the compute entry points
raise when the CUDA library or a B200-class GPU is missing."""
from .render import (Camera, Scene, SceneF64, NODE64_DT, build_lbvh, build_options, probe_read_bandwidth, probe_host_fill_bandwidth, render, frame_setup, shard_rows, write_ppm, kernel_launch_count, PinnedBuffer, device_count,
                     NO_SHADOWS, COMPACT_ROWS, FRAME_RETAINED, PAYLOAD_NO_FLIP, MISS, STATIC_MAX_PIXELS, CeresError, NODE_DT)
from . import configs
from . import scenes

__all__ = ["Camera", "Scene", "SceneF64", "NODE64_DT", "build_lbvh", "build_options", "probe_read_bandwidth", "probe_host_fill_bandwidth", "render", "frame_setup", "shard_rows", "write_ppm", "kernel_launch_count", "PinnedBuffer",
           "device_count", "NO_SHADOWS", "COMPACT_ROWS", "FRAME_RETAINED", "PAYLOAD_NO_FLIP", "MISS", "STATIC_MAX_PIXELS", "CeresError", "NODE_DT", "configs", "scenes"]

"""Multi-GPU sharding of the render path: one process per GPU, scene replicated, no exchange step inside the path
(render.hpp:104-153 writes only pixels[index..index+2] and two '+'-reduced counters).

    single frame   rows are grouped in blocks of `block_rows`; block b belongs to rank b mod N (interleaved, because
                   the dragon covers only the middle of the screen and contiguous bands would be unbalanced).  Every
                   rank renders its blocks into a packed band (COMPACT_ROWS); FrameGather collects the bands on rank 0
                   with ONE gather and scatters the rows to their place in the frame; the two counters are summed.
    sequence       frame f of an `anim` sequence belongs to rank f mod N (anim.cpp:82-105: frames share only the
                   read-only scene); no collective at all, every rank keeps / writes its own frames.

The classes only use torch.distributed collectives on plain tensors, so the same code runs over NCCL on CUDA tensors
(bench.py, N GPUs) and over gloo on CPU tensors (tests/test_shard_gloo.py, world_size 2).
"""
import os

import numpy as np
import torch
import torch.distributed as dist

from .render import shard_rows


def sequence_frames(n_frames, rank, world):
    """Frames of a sequence owned by `rank`: f = rank, rank + world, ... (round-robin)."""
    return list(range(rank, n_frames, max(world, 1)))


class FrameGather:
    """Buffers + collective that assemble one frame from the packed bands of all ranks.

    band      this rank's rows, packed in local order (what ceres_render_device writes with COMPACT_ROWS);
              padded to the largest band so that the gather is regular
    frame     rank 0 only: the assembled row-major frame, row j at 3*W*j (render.hpp:107)
    counters  int64[2] (rays, hits) of this rank's shard; summed over ranks by assemble()
    """

    def __init__(self, width, height, block_rows, rank, world, device, channels=3, dtype=torch.float32):
        self.W, self.H, self.block_rows, self.rank, self.world = int(width), int(height), int(block_rows), rank, world
        self.shard = (self.block_rows, rank, world) if world > 1 else None
        self.row_floats = self.W * channels
        self.rows = shard_rows(self.H, self.shard)
        self.n_rows = len(self.rows)
        self.rows_max = len(shard_rows(self.H, (self.block_rows, 0, world))) if world > 1 else self.H
        self.band = torch.zeros(self.rows_max * self.row_floats, dtype=dtype, device=device)
        self.counters = torch.zeros(2, dtype=torch.int64, device=device)
        self.frame = self.band
        self._bands = None
        if world > 1 and rank == 0:
            self._bands = [torch.empty_like(self.band) for _ in range(world)]
            self.frame = torch.empty(self.H * self.row_floats, dtype=dtype, device=device)
            self._row_index = [torch.as_tensor(np.asarray(shard_rows(self.H, (self.block_rows, r, world)), np.int64),
                                               device=device) for r in range(world)]

    def assemble(self):
        """Collective: gather the bands to rank 0, put every row in its place, sum the counters on all ranks."""
        if self.world <= 1:
            return self.frame
        dist.gather(self.band, self._bands, dst=0)
        if self.rank == 0:
            fv = self.frame.view(self.H, self.row_floats)
            for r in range(self.world):
                idx = self._row_index[r]
                fv.index_copy_(0, idx, self._bands[r].view(self.rows_max, self.row_floats)[: idx.numel()])
        dist.all_reduce(self.counters)
        return self.frame if self.rank == 0 else None


class _DevMem:
    """A raw device pointer presented through __cuda_array_interface__ so that torch can view it without a copy."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 3}


class PeerFrame:
    """One frame on N GPUs, fused with its gather over NVLink peer memory (ceres_peer_frame, include/ceres_b200.h).

    Rank 0 owns the device frame; every other rank maps it and its shade kernel stores its row blocks straight into
    rank 0's memory, so render() is the whole multi-GPU step: no NCCL call, no gather pass.  The 64-byte mapping
    handle travels once, at construction, through a torch.distributed broadcast (any backend)."""

    def __init__(self, scene, width, height, rank=0, world=1, want_ids=False, owner=None):
        import ctypes as C
        from . import _capi
        self._L, self._scene = _capi.lib(), scene
        self.W, self.H, self.rank, self.world, self.want_ids = int(width), int(height), rank, world, bool(want_ids)
        h = C.c_void_p()
        if owner is not None:          # same process (tests, single-process multi-GPU hosts)
            _capi.check(self._L.ceres_peer_frame_alias(owner._h, scene._h, rank, C.byref(h)))
        else:
            blob = (C.c_ubyte * _capi.PEER_HANDLE_BYTES)()
            if rank == 0:
                _capi.check(self._L.ceres_peer_frame_create(scene._h, self.W, self.H, int(want_ids), world, C.byref(h), blob))
            if world > 1 and dist.is_initialized():      # not initialised: the other ranks are aliases in this process
                box = [bytes(blob) if rank == 0 else None]
                dist.broadcast_object_list(box, src=0)
                if rank != 0:
                    blob = (C.c_ubyte * _capi.PEER_HANDLE_BYTES).from_buffer_copy(box[0])
                    _capi.check(self._L.ceres_peer_frame_open(scene._h, blob, self.W, self.H, int(want_ids), world, rank, C.byref(h)))
        self._h = h
        self.counters = torch.zeros(2, dtype=torch.int64, device=torch.device("cuda", scene.device)) if rank == 0 else None

    def render(self, camera, sun_position, flags=0, stream=None):
        """Enqueue this rank's share of the next frame (all ranks call it once per frame)."""
        import ctypes as C
        from . import _capi
        from .render import _stream_handle, _p
        sun = np.asarray(sun_position, np.float32)
        cnt = C.c_void_p(self.counters.data_ptr()) if self.counters is not None else None
        _capi.check(self._L.ceres_render_peer(self._scene._h, C.byref(camera.c()), _p(sun), flags,
                                              C.c_void_p(_stream_handle(stream, self._scene.device)), self._h, cnt))

    @property
    def pixels(self):
        """float32 [H, W, 3] view of the assembled frame in rank 0's memory."""
        ptr = self._L.ceres_peer_frame_pixels(self._h)
        return torch.as_tensor(_DevMem(ptr, (self.H, self.W, 3), "<f4"), device=torch.device("cuda", self._scene.device))

    @property
    def ids(self):
        ptr = self._L.ceres_peer_frame_ids(self._h)
        if not ptr:
            return None
        return torch.as_tensor(_DevMem(ptr, (self.H, self.W), "<i4"), device=torch.device("cuda", self._scene.device))

    def error(self):
        return int(self._L.ceres_peer_frame_error(self._h))

    def invalidate(self):
        """Somebody wrote into the frame (through .pixels / .ids): this rank's next render() re-clears all of its rows."""
        from . import _capi
        _capi.check(self._L.ceres_peer_frame_invalidate(self._h))

    def close(self):
        if getattr(self, "_h", None):
            self._L.ceres_peer_frame_destroy(self._h)
            self._h = None

    __del__ = close


class SharedHostFrame:
    """A HOST framebuffer shared by the ranks of one box (a file in /dev/shm mapped by every process and page-locked
    in each with ceres_host_register).  Every rank renders its row blocks and DMA-writes them in place over ITS OWN
    PCIe link, so the end-to-end frame needs no gather at all and the drain bandwidth grows with the GPU count
    (the one-process equivalent is ceres_render_multi).  `array` is the frame, valid on every rank after render()."""

    CTL_BYTES = 4096      # control page behind the frame: epoch[r] at word r, (rays, hits) of rank r at words 64 + 2r

    def __init__(self, name, shape, rank, world, dtype=np.float32, block_rows=None):
        from . import _capi, configs
        self._L = _capi.lib()
        self.rank, self.world = rank, world
        self.block_rows = block_rows or configs.TILE_BLOCK_ROWS
        self.path = os.path.join("/dev/shm", name)
        nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self._frame_bytes = (nbytes + 4095) // 4096 * 4096
        if rank == 0:
            with open(self.path, "wb") as f:
                f.truncate(self._frame_bytes + self.CTL_BYTES)
        if world > 1:
            dist.barrier()
        self._map = np.memmap(self.path, dtype=np.uint8, mode="r+", shape=(self._frame_bytes + self.CTL_BYTES,))
        self.array = self._map[:nbytes].view(dtype).reshape(shape)
        self._ctl = self._map[self._frame_bytes:].view(np.uint64)
        self._ptr, self._nbytes = self.array.ctypes.data, nbytes
        _capi.check(self._L.ceres_host_register(self._ptr, nbytes))
        self._epoch = 0

    def render(self, scene, camera, sun_position, width, height, flags=0, timeout_s=30.0):
        """Collective: every rank renders + copies its rows; returns the frame's (rays, hits) on every rank.
        Completion is a barrier in the shared control page (no NCCL/gloo call on the data path): a rank publishes its
        counters, then its epoch, and waits until every rank's epoch has reached the frame's.  A uint8 / big-endian uint16 frame
        receives the quantised, flipped P6 payload (ceres_render_ppm8_shard / ppm16_shard) instead of floats."""
        import time
        shard = (self.block_rows, self.rank, self.world) if self.world > 1 else None
        if self.array.dtype == np.uint8:
            _, cnt = scene.render_ppm8(camera, sun_position, width, height, flags=flags, out=self.array, shard=shard)
        elif self.array.dtype.itemsize == 2:
            _, cnt = scene.render_ppm16(camera, sun_position, width, height, flags=flags, out=self.array, shard=shard)
        else:
            _, _, cnt = scene.render(camera, sun_position, width, height, flags=flags, pixels=self.array, shard=shard)
        if self.world <= 1:
            return cnt
        self._epoch += 1
        c = self._ctl
        base = 64 + (self._epoch & 1) * 2 * self.world      # counter slots alternate: a rank may already be one frame ahead
        c[base + 2 * self.rank], c[base + 2 * self.rank + 1] = cnt
        c[self.rank] = self._epoch          # x86 stores are not reordered: the counters are visible before the epoch
        t0 = time.perf_counter()
        while int(c[: self.world].min()) < self._epoch:
            if time.perf_counter() - t0 > timeout_s:
                raise RuntimeError("SharedHostFrame: a rank did not deliver its rows in time")
        both = c[base: base + 2 * self.world].reshape(self.world, 2).sum(axis=0)
        return int(both[0]), int(both[1])

    def close(self):
        if getattr(self, "array", None) is not None:
            self._L.ceres_host_unregister(self._ptr)
            self.array = self._ctl = self._map = None
            if self.world > 1:
                dist.barrier()
            if self.rank == 0 and os.path.exists(self.path):
                os.unlink(self.path)

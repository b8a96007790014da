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

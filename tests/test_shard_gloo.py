"""N > 1 host logic on CPU: world_size-2 `gloo` processes run ceres_raytracer_b200.shard over CPU tensors.

The band every rank contributes is cut from the ORACLE's frame (the CUDA kernels cannot run here), so the test pins
exactly what the multi-GPU path adds to the single-GPU path: row ownership, packing order, the gather, the row scatter
on rank 0 and the counter reduction.  (tests/test_gpu_parity.py::test_row_block_shards_reassemble_exactly pins that
the kernels write those packed bands; bench.py --gpus N uses the same FrameGather over NCCL.)"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, W, H, block_rows, frame_path, out_path):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from ceres_raytracer_b200.shard import FrameGather, sequence_frames
        from ceres_raytracer_b200.render import shard_rows
        z = np.load(frame_path)
        full, ids = z["pixels"], z["ids"]
        g = FrameGather(W, H, block_rows, rank, world, torch.device("cpu"))
        rows = shard_rows(H, (block_rows, rank, world))
        assert g.n_rows == len(rows) and g.rows_max >= g.n_rows
        # what ceres_render_device(COMPACT_ROWS) leaves in the band: this rank's rows, packed in local order
        g.band[: len(rows) * W * 3] = torch.from_numpy(full[rows].reshape(-1))
        hit = ids[rows] != 0xFFFFFFFF
        g.counters[0] = len(rows) * W + int(hit.sum())        # rays = pixels + primary hits (render.hpp:115,138)
        g.counters[1] = int(hit.sum())
        frame = g.assemble()
        if rank == 0:
            np.savez(out_path, frame=frame.numpy().reshape(H, W, 3), counters=g.counters.numpy())
        else:
            assert frame is None
        # sequences: round-robin ownership covers every frame exactly once
        mine = torch.zeros(7, dtype=torch.int64)
        mine[sequence_frames(7, rank, world)] = 1
        dist.all_reduce(mine)
        assert mine.tolist() == [1] * 7
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("W,H,block_rows", [(96, 54, 16), (64, 37, 4), (40, 8, 16)])
def test_two_rank_gather_reassembles_the_oracle_frame(tmp_path, bunny_pod, W, H, block_rows):
    import ceres_raytracer_b200 as cr
    from oracle import port
    cam, sun = cr.configs.BUNNY_CAMERA.as_array(), np.array(cr.configs.BUNNY_SUN, np.float32)
    o = port.render(bunny_pod, cam, sun, W, H, 1)
    frame_path, out_path = str(tmp_path / "frame.npz"), str(tmp_path / "out.npz")
    np.savez(frame_path, pixels=o["pixels"].reshape(H, W, 3), ids=o["ids"].reshape(H, W))
    mp.spawn(_worker, args=(2, _free_port(), W, H, block_rows, frame_path, out_path), nprocs=2, join=True)
    z = np.load(out_path)
    assert np.array_equal(z["frame"].view(np.uint32), o["pixels"].reshape(H, W, 3).view(np.uint32))
    hits = int((o["ids"] != 0xFFFFFFFF).sum())
    assert z["counters"].tolist() == [W * H + hits, hits]


def test_shard_rows_partition():
    from ceres_raytracer_b200.render import shard_rows
    for H, br, n in [(2160, 16, 8), (1080, 16, 3), (37, 4, 2), (5, 16, 4)]:
        got = np.concatenate([shard_rows(H, (br, r, n)) for r in range(n)])
        assert sorted(got.tolist()) == list(range(H))

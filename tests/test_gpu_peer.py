"""GPU tests of the multi-GPU frame paths: the peer-memory frame (ceres_render_peer: every rank's shade kernel stores
its row blocks straight into rank 0's frame) and the shared host frame (every rank DMA-writes its row blocks into one
page-locked buffer).  Both must reproduce the single-GPU frame bit for bit, which test_gpu_parity.py pins to the
oracle.  The ranks run on distinct GPUs when the box has them and share GPU 0 otherwise (CUDA IPC works between
processes on one device), so the protocol is exercised on a 1-GPU box too."""
import os
import socket
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SUN = np.array([-50, -20, 0], np.float32)


@pytest.fixture(scope="module")
def cr():
    import ceres_raytracer_b200 as cr
    return cr


@pytest.mark.parametrize("world", [1, 2, 3])
@pytest.mark.parametrize("flags", [0, 1])
def test_peer_frame_same_process(cr, dragon_pod, world, flags):
    """world ranks in ONE process (aliases of the owner's frame), each with its own scene and stream."""
    import torch
    from ceres_raytracer_b200.shard import PeerFrame
    cam, W, H = cr.configs.DRAGON_CAMERA, 1280, 720
    ngpu = cr.device_count()
    scenes = [cr.Scene.from_pod(dragon_pod, device=r % ngpu) for r in range(world)]
    ref_px, ref_ids, ref_cnt = scenes[0].render(cam, SUN, W, H, flags=flags, want_ids=True)
    streams = [torch.cuda.Stream(device=s.device) for s in scenes]
    owner = PeerFrame(scenes[0], W, H, 0, world, want_ids=True)
    frames = [owner] + [PeerFrame(scenes[r], W, H, r, world, want_ids=True, owner=owner) for r in range(1, world)]
    # a second view whose screen rectangle neither contains nor is contained in the first one's: the dirty-rectangle clear
    # has to wipe what the previous epoch left outside the new rectangle
    cam2 = cr.Camera((3.0, -15.0, 2.5), (-0.1, 1.0, 0.0), (0.0, 0.0, 1.0), 45.0)
    ref2 = scenes[0].render(cam2, SUN, W, H, flags=flags, want_ids=True)
    try:
        # epochs 0-2: the frame is poisoned by the caller, who says so (invalidate): every pixel must be rewritten;
        # epochs 3-4: untouched frame, same view (no clear at all); 5: other view; 6: back; 7: poisoned again
        for it, (view, poison) in enumerate([(cam, True)] * 3 + [(cam, False)] * 2 + [(cam2, False), (cam, False), (cam2, True)]):
            want = (ref_px, ref_ids, ref_cnt) if view is cam else ref2
            if poison:
                owner.pixels.fill_(float("nan"))
                owner.ids.fill_(12345)
                for f in frames:
                    f.invalidate()
            torch.cuda.synchronize(scenes[0].device)
            for r in range(world):        # rank 0 first: the ranks may share one GPU (include/ceres_b200.h)
                frames[r].render(view, SUN, flags=flags, stream=streams[r].cuda_stream)
            for s in streams:
                s.synchronize()
            assert owner.error() == 0
            px = owner.pixels.cpu().numpy()
            ids = owner.ids.cpu().numpy().view(np.uint32)
            assert np.array_equal(px.view(np.uint32), want[0].view(np.uint32)), f"epoch {it}"
            assert np.array_equal(ids, want[1]), f"epoch {it}"
            assert tuple(owner.counters.cpu().tolist()) == want[2]
    finally:
        for f in frames[1:] + [owner]:
            f.close()
        for s in scenes:
            s.close()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_main(rank, world, port, W, H, out_path):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch
    import torch.distributed as dist
    import ceres_raytracer_b200 as cr
    from ceres_raytracer_b200.shard import PeerFrame, SharedHostFrame
    from conftest import scene_pod
    dev = rank % cr.device_count()
    torch.cuda.set_device(dev)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cam = cr.configs.DRAGON_CAMERA
        scene = cr.Scene.from_pod(scene_pod("dragon"), device=dev)
        # ---- device frame in rank 0's memory, written by every rank's kernels
        pf = PeerFrame(scene, W, H, rank, world, want_ids=True)
        out = {}
        for it in range(3):
            dist.barrier()
            pf.render(cam, SUN)
            torch.cuda.synchronize()
            dist.barrier()
        if rank == 0:
            assert pf.error() == 0
            out.update(px=pf.pixels.cpu().numpy(), ids=pf.ids.cpu().numpy().view(np.uint32), cnt=pf.counters.cpu().numpy())
        dist.barrier()
        pf.close()
        # ---- host frame shared by the ranks, every rank DMA-writes its own row blocks
        hf = SharedHostFrame(f"ceres_test_{port}", (H, W, 3), rank, world)
        cnt = hf.render(scene, cam, SUN, W, H)
        if rank == 0:
            out.update(host_px=hf.array.copy(), host_cnt=np.asarray(cnt))
        dist.barrier()
        hf.close()
        if rank == 0:
            np.savez(out_path, **out)
        scene.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_peer_and_shared_host_frames_across_processes(cr, dragon_pod, tmp_path, world):
    import torch.multiprocessing as mp
    cam, W, H = cr.configs.DRAGON_CAMERA, 1280, 720
    with cr.Scene.from_pod(dragon_pod) as s:
        ref_px, ref_ids, ref_cnt = s.render(cam, SUN, W, H, want_ids=True)
    out_path = str(tmp_path / "out.npz")
    mp.spawn(_rank_main, args=(world, _free_port(), W, H, out_path), nprocs=world, join=True)
    z = np.load(out_path)
    assert np.array_equal(z["px"].view(np.uint32), ref_px.view(np.uint32))
    assert np.array_equal(z["ids"], ref_ids)
    assert tuple(z["cnt"].tolist()) == ref_cnt
    assert np.array_equal(z["host_px"].view(np.uint32), ref_px.view(np.uint32))
    assert tuple(z["host_cnt"].tolist()) == ref_cnt

import os, sys
sys.path.insert(0, os.getcwd())
import numpy as np, torch
import ceres_raytracer_b200 as cr
with np.load("tests/golden/dragon_scene.npz") as z:
    pod = {k: z[k] for k in z.files}
cam = cr.configs.DRAGON_CAMERA
sun = np.array([-50, -20, 0], np.float32)
W, H = 1280, 720
s = cr.Scene.from_pod(pod)
px = torch.full((H * W * 3,), float("nan"), dtype=torch.float32, device="cuda")
ids = torch.zeros(H * W, dtype=torch.int32, device="cuda")
cnt = torch.zeros(2, dtype=torch.int64, device="cuda")
ref_px = torch.empty_like(px); ref_ids = torch.empty_like(ids); ref_cnt = torch.zeros_like(cnt)
s.render_device(cam, sun, W, H, pixels=ref_px, ids=ref_ids, counters=ref_cnt)
s.render_device_static(cam, sun, W, H, pixels=px, ids=ids, counters=cnt)      # warm-up: symbol address, occupancy
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    s.render_device_static(cam, sun, W, H, pixels=px, ids=ids, counters=cnt)
for rep in range(3):
    px.fill_(float("nan")); ids.zero_(); cnt.zero_()
    g.replay()
    torch.cuda.synchronize()
    print(rep, bool(torch.equal(px.view(torch.int32), ref_px.view(torch.int32))), bool(torch.equal(ids, ref_ids)), cnt.tolist(), ref_cnt.tolist())
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(200): g.replay()
e1.record(); torch.cuda.synchronize()
print("graph replay us/frame", e0.elapsed_time(e1) * 5)
e0.record()
for _ in range(200): s.render_device_static(cam, sun, W, H, pixels=px, ids=ids, counters=cnt)
e1.record(); torch.cuda.synchronize()
print("direct calls us/frame", e0.elapsed_time(e1) * 5)

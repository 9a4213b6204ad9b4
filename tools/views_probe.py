#!/usr/bin/env python
"""Development probe: nsdf_render_views (n views in one launch) against n calls of nsdf_render: bytes equal? ms per frame?
usage: views_probe.py [W H] [batch sizes...]"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import host
W, H = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1920, 1080)
batches = [int(a) for a in sys.argv[3:] if "=" not in a] or [2, 8, 16, 64]
tuning = {a.split("=")[0]: int(a.split("=")[1]) for a in sys.argv[3:] if "=" in a}     # nsdf_tuning knobs for the batched calls
nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", "plane_1.h5"))
matcap = host.Image(); assert matcap.loadPNG(os.path.join(ROOT, "data", "matcaps", "blue.png"))
host.copyStaticSettings(host.COLOR_MATCAP, 3)
host.setSceneOptions(host.SCENE_MANYSPHERE, 6000, 0, 0, host.MATH_REFERENCE)
N = max(batches)
views = [(0.0, k * 5.625) for k in range(N)]
single = torch.zeros(N * W * H, dtype=torch.int32, device="cuda")
def one_by_one(n):
    for v in range(n):
        host.updateViewMatrices((views[v][0], views[v][1], 0.0), (0.0, 0.0, -2.0), 0)
        host.render_kernel(single[v * W * H:(v + 1) * W * H], W, H, nn, matcap)
one_by_one(N); torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record(); one_by_one(N); b.record(); torch.cuda.synchronize()
print(f"one launch per view: {a.elapsed_time(b) / N:.2f} ms per frame ({N} views)", flush=True)
ref = single.cpu().numpy()
out = torch.zeros(N * W * H, dtype=torch.int32, device="cuda")
for B in batches:
    def batched():
        for s in range(0, N, B):
            cams = host.view_cameras(views[s:s + B])
            host.render_views(out[s * W * H:(s + B) * W * H], W, H, nn, cams, matcap)
    host.setTuning(**tuning)
    out.zero_(); batched(); torch.cuda.synchronize()
    eq = bool(np.array_equal(out.cpu().numpy(), ref))
    a.record(); batched(); b.record(); torch.cuda.synchronize()
    host.setTuning()
    print(f"batches of {B:3d} {tuning or ''}: {a.elapsed_time(b) / N:.2f} ms per frame, bytes equal to one-by-one: {eq}", flush=True)

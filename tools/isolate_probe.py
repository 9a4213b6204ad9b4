#!/usr/bin/env python
"""Development sweep: long-ray isolation (NSDF_COOP_ISOLATE) in the schedules of the tracer."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import _capi as capi, host
def run(nn, mc, W, H, flags, evict=0, iters=4):
    host.setSceneOptions(host.SCENE_TANH, 6000, flags, evict)
    fr = torch.zeros(W * H, dtype=torch.int32, device="cuda")
    for _ in range(2): host.render_kernel(fr, W, H, nn, mc)
    torch.cuda.synchronize(); ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); host.render_kernel(fr, W, H, nn, mc); b.record(); b.synchronize(); ts.append(a.elapsed_time(b))
    return min(ts), fr.cpu().numpy()
for model, W, H in (("plane_1", 1920, 1080), ("3a3d4a90a2db90b4203936772104a82d.obj", 3840, 2160), ("car_1", 512, 512)):
    nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", model + ".h5"))
    mc = host.Image(); mc.loadPNG(os.path.join(ROOT, "data", "matcaps", "blue.png"))
    host.copyStaticSettings(host.COLOR_MATCAP, 3); host.updateViewMatrices((0, 0, 0), (0, 0, -2), 0)
    os.environ["NSDF_COOP_ISOLATE"] = "0"; os.environ["NSDF_COOP_QUEUE"] = "1"
    t0, ref = run(nn, mc, W, H, 0)
    print(model[:8], f"default (queue, no isolation) {t0:.2f}", flush=True)
    os.environ["NSDF_TWO_PASS2_COOP"] = "1"; os.environ["NSDF_TRACE_P"] = "1"
    for iso in (0, 64, 128, 256, 512, 1024):
        os.environ["NSDF_COOP_ISOLATE"] = str(iso)
        os.environ["NSDF_COOP_QUEUE"] = "0"
        t1, a = run(nn, mc, W, H, 0)
        os.environ["NSDF_COOP_QUEUE"] = "1"
        t2, b = run(nn, mc, W, H, capi.RENDER_TWO_KERNELS, 128)
        t3, _ = run(nn, mc, W, H, capi.RENDER_TC_FARFIELD)
        print(f"  isolate {iso:4d}: single launch, no queue {t1:.2f} | two-launch(128)+coop {t2:.2f} | mixed {t3:.2f} | exact frames identical {np.array_equal(a, ref)} {np.array_equal(b, ref)}", flush=True)

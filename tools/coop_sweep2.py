#!/usr/bin/env python
"""Development sweep: hand-over threshold and quiet-SM period of the default single-launch tracer."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import host
def run(nn, mc, W, H, iters=4):
    fr = torch.zeros(W * H, dtype=torch.int32, device="cuda")
    for _ in range(2): host.render_kernel(fr, W, H, nn, mc)
    torch.cuda.synchronize(); ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); host.render_kernel(fr, W, H, nn, mc); b.record(); b.synchronize(); ts.append(a.elapsed_time(b))
    return min(ts)
for model, W, H in (("plane_1", 1920, 1080), ("3a3d4a90a2db90b4203936772104a82d.obj", 3840, 2160)):
    nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", model + ".h5"))
    mc = host.Image(); mc.loadPNG(os.path.join(ROOT, "data", "matcaps", "blue.png"))
    host.copyStaticSettings(host.COLOR_MATCAP, 3); host.updateViewMatrices((0, 0, 0), (0, 0, -2), 0)
    host.setSceneOptions(host.SCENE_TANH, 6000, 0, 0)
    for period in (8, 16, 32, 74):
        row = []
        for ev in (128, 256, 512, 1024, 2048):
            os.environ["NSDF_COOP_QUIET_PERIOD"] = str(period); os.environ["NSDF_COOP_EVICT"] = str(ev)
            row.append(f"evict {ev}: {run(nn, mc, W, H):.2f}")
        print(model[:8], "period", period, " | ".join(row), flush=True)
    os.environ["NSDF_COOP_QUEUE"] = "0"
    print(model[:8], "no queue:", round(run(nn, mc, W, H), 2), flush=True)
    os.environ["NSDF_COOP_QUEUE"] = "1"

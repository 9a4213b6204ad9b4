#!/usr/bin/env python
"""Development sweep: tile priority (tile_priority.cuh) x hand-over / isolation settings of the exact and mixed tracers."""
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
def setenv(**kw):
    for k, v in kw.items(): os.environ[k] = str(v)
for model, W, H in (("plane_1", 1920, 1080), ("3a3d4a90a2db90b4203936772104a82d.obj", 3840, 2160), ("car_1", 512, 512)):
    nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", model + ".h5"))
    mc = host.Image(); mc.loadPNG(os.path.join(ROOT, "data", "matcaps", "blue.png"))
    host.copyStaticSettings(host.COLOR_MATCAP, 3); host.updateViewMatrices((0, 0, 0), (0, 0, -2), 0)
    setenv(NSDF_TILE_PRIORITY_STEPS=0, NSDF_COOP_ISOLATE=0, NSDF_COOP_QUEUE=1, NSDF_TWO_PASS_MIN_PIXELS=1 << 30)
    t0, ref = run(nn, mc, W, H, 0)
    print(model[:8], f"baseline single launch (queue 256/16, no priority) {t0:.2f}", flush=True)
    for steps in (0, 32, 48, 96):
        setenv(NSDF_TILE_PRIORITY_STEPS=steps)
        row = []
        for q, iso in ((1, 0), (1, 256), (0, 0), (0, 128), (0, 256), (0, 512)):
            setenv(NSDF_COOP_QUEUE=q, NSDF_COOP_ISOLATE=iso)
            t, a = run(nn, mc, W, H, 0)
            row.append(f"q{q}/iso{iso}: {t:.2f}{'' if np.array_equal(a, ref) else ' DIFF'}")
        setenv(NSDF_COOP_QUEUE=1, NSDF_COOP_ISOLATE=0)
        t2, b = run(nn, mc, W, H, capi.RENDER_TWO_PASS)
        setenv(NSDF_COOP_ISOLATE=256)
        t2i, b2 = run(nn, mc, W, H, capi.RENDER_TWO_PASS)
        t3i, _ = run(nn, mc, W, H, capi.RENDER_TC_FARFIELD)
        setenv(NSDF_COOP_ISOLATE=0)
        t3, _ = run(nn, mc, W, H, capi.RENDER_TC_FARFIELD)
        print(f"  probe_steps {steps:3d}: " + " | ".join(row) + f" || two-pass {t2:.2f} iso256 {t2i:.2f}{'' if np.array_equal(b, ref) and np.array_equal(b2, ref) else ' DIFF'} || mixed {t3:.2f} iso256 {t3i:.2f}", flush=True)

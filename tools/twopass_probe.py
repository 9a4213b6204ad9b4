#!/usr/bin/env python
"""Development probe: frame time of the exact tracer's schedules (default cooperative launch, two-launch with the latency
kernel, two-launch with the cooperative tracer refilling from the static queue) for several hand-over step counts."""
import os, sys, json
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import _capi as capi, host
def run(nn, mc, W, H, flags, evict=0, iters=5):
    host.setSceneOptions(host.SCENE_TANH, 6000, flags, evict)
    fr = torch.zeros(W * H, dtype=torch.int32, device="cuda")
    for _ in range(2): host.render_kernel(fr, W, H, nn, mc)
    torch.cuda.synchronize(); ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); host.render_kernel(fr, W, H, nn, mc); b.record(); b.synchronize(); ts.append(a.elapsed_time(b))
    return min(ts), fr.cpu().numpy()
for model, W, H in (("plane_1", 1920, 1080), ("3a3d4a90a2db90b4203936772104a82d.obj", 3840, 2160)):
    nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", model + ".h5"))
    mc = host.Image(); mc.loadPNG(os.path.join(ROOT, "data", "matcaps", "blue.png"))
    host.copyStaticSettings(host.COLOR_MATCAP, 3); host.updateViewMatrices((0, 0, 0), (0, 0, -2), 0)
    t0, ref = run(nn, mc, W, H, 0)
    print(model, "coop default", round(t0, 2), flush=True)
    for P in ("1",):
        os.environ["NSDF_TRACE_P"] = P
        for ev in (128,):
            os.environ["NSDF_TWO_PASS2_COOP"] = "0"
            t1, a = run(nn, mc, W, H, capi.RENDER_TWO_KERNELS, ev)
            os.environ["NSDF_TWO_PASS2_COOP"] = "1"
            t2, b = run(nn, mc, W, H, capi.RENDER_TWO_KERNELS, ev)
            t3, _ = run(nn, mc, W, H, capi.RENDER_PHASE_THROUGHPUT, ev)
            print(f"  P={P} evict={ev}: two-launch latency-kernel {t1:.2f} ms | two-launch coop-from-queue {t2:.2f} ms | pass1 alone {t3:.2f} ms | identical {np.array_equal(a, ref)} {np.array_equal(b, ref)}", flush=True)
    os.environ["NSDF_FARFIELD_SKIP_PASS2"] = "1"
    t4, _ = run(nn, mc, W, H, capi.RENDER_TC_FARFIELD)
    os.environ["NSDF_FARFIELD_SKIP_PASS2"] = "0"
    t5, _ = run(nn, mc, W, H, capi.RENDER_TC_FARFIELD)
    print(f"  mixed: pass1 (tensor cores) alone {t4:.2f} ms, both passes {t5:.2f} ms", flush=True)

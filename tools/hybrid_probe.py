#!/usr/bin/env python
"""Development probe: the cooperative tracer with ray-per-lane producers (NSDF_COOP_HYBRID) against the default
schedule, over hand-over step counts and quiet-SM periods; frames must stay bytewise identical."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import _capi as capi, host
def run(nn, mc, W, H, flags=0, evict=0, iters=5):
    host.setSceneOptions(host.SCENE_TANH, 6000, flags, evict)
    fr = torch.zeros(W * H, dtype=torch.int32, device="cuda")
    for _ in range(2): host.render_kernel(fr, W, H, nn, mc)
    torch.cuda.synchronize(); ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); host.render_kernel(fr, W, H, nn, mc); b.record(); b.synchronize(); ts.append(a.elapsed_time(b))
    return min(ts), fr.cpu().numpy()
def main():
    cases = [("plane_1", 1920, 1080), ("3a3d4a90a2db90b4203936772104a82d.obj", 1920, 1080), ("3a3d4a90a2db90b4203936772104a82d.obj", 3840, 2160)]
    if len(sys.argv) > 1: cases = cases[:int(sys.argv[1])]
    for model, W, H in cases:
        nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", model + ".h5"))
        mc = host.Image(); mc.loadPNG(os.path.join(ROOT, "data", "matcaps", "blue.png"))
        host.copyStaticSettings(host.COLOR_MATCAP, 3); host.updateViewMatrices((0, 0, 0), (0, 0, -2), 0)
        os.environ["NSDF_COOP_HYBRID"] = "0"
        os.environ["NSDF_TWO_PASS_MIN_PIXELS"] = str(1 << 30)
        t0, ref = run(nn, mc, W, H)
        os.environ["NSDF_TWO_PASS_MIN_PIXELS"] = "3000000"
        t1, _ = run(nn, mc, W, H)
        os.environ["NSDF_TWO_PASS_MIN_PIXELS"] = str(1 << 30)
        print(f"{model[:8]} {W}x{H}: coop {t0:.2f} ms, default schedule {t1:.2f} ms", flush=True)
        for period in (8, 16, 32):
            for ev in (96, 192, 256):
                for drain in (8, 20, 32):
                    os.environ["NSDF_COOP_HYBRID"] = str(drain)
                    os.environ["NSDF_COOP_QUIET_PERIOD"] = str(period)
                    os.environ["NSDF_COOP_EVICT"] = str(ev)
                    t, a = run(nn, mc, W, H, iters=3)
                    print(f"  hybrid period={period} evict={ev} drain={drain}: {t:.2f} ms identical={np.array_equal(a, ref)}", flush=True)
        for k in ("NSDF_COOP_HYBRID", "NSDF_COOP_QUIET_PERIOD", "NSDF_COOP_EVICT"): os.environ.pop(k, None)


if __name__ == '__main__':
    main()

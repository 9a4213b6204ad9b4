#!/usr/bin/env python
"""Development probe: the three schedules a large call can take (NSDF_LARGE_SCHEDULE 0 cooperative, 1 hybrid, 2 two-pass)
on large frames of the shipped shapes, the hybrid one over quiet-SM periods and backlog limits; bytes must agree."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import host
from tools.hybrid_probe import run
sizes = ((3840, 2160), (1920, 1080)) if len(sys.argv) < 2 else ((3840, 2160),)
for model in ("3a3d4a90a2db90b4203936772104a82d.obj", "plane_1", "car_1"):
    for (W, H) in sizes:
        nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", model + ".h5"))
        mc = host.Image(); mc.loadPNG(os.path.join(ROOT, "data", "matcaps", "blue.png"))
        host.copyStaticSettings(host.COLOR_MATCAP, 3); host.updateViewMatrices((0, 0, 0), (0, 0, -2), 0)
        os.environ["NSDF_TWO_PASS_MIN_PIXELS"] = "1000000"
        res = {}
        for sched in ("0", "2"):
            os.environ["NSDF_LARGE_SCHEDULE"] = sched
            res[sched] = run(nn, mc, W, H, iters=3)
        os.environ["NSDF_LARGE_SCHEDULE"] = "1"
        line = f"{model[:8]} {W}x{H}: coop {res['0'][0]:.2f} two-pass {res['2'][0]:.2f} | hybrid"
        same = np.array_equal(res["0"][1], res["2"][1])
        for period in ("16", "32"):
            for room in ("0", "128", "512", "2048"):
                os.environ["NSDF_COOP_QUIET_PERIOD"] = period; os.environ["NSDF_COOP_ROOM"] = room
                t, a = run(nn, mc, W, H, iters=3)
                same = same and np.array_equal(a, res["0"][1])
                line += f" p{period}/r{room} {t:.2f}"
        for k in ("NSDF_LARGE_SCHEDULE", "NSDF_COOP_QUIET_PERIOD", "NSDF_COOP_ROOM"): os.environ.pop(k)
        print(line, f"identical={same}", flush=True)

#!/usr/bin/env python
"""Development probe: the 64-view 1080p turntable of bench.py through nsdf_render_views, default tuning against variants
(bytes equal? ms per frame?).  usage: bulk4_probe.py [WxH] [N_VIEWS] "k=v,k=v" "k=v" ..."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import host
W, H = 1920, 1080
args = sys.argv[1:]
if args and "x" in args[0] and "=" not in args[0]:
    W, H = (int(v) for v in args.pop(0).split("x"))
N = int(args.pop(0)) if args and args[0].isdigit() else 64
variants = [dict((kv.split("=")[0], int(kv.split("=")[1])) for kv in a.split(",") if kv) for a in args]
nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", "plane_1.h5"))
matcap = host.Image(); assert matcap.loadPNG(os.path.join(ROOT, "data", "matcaps", "blue.png"))
host.copyStaticSettings(host.COLOR_MATCAP, 3)
host.setSceneOptions(host.SCENE_MANYSPHERE, 6000, 0, 0, host.MATH_REFERENCE)
cams = host.view_cameras([(0.0, k * 5.625) for k in range(N)])
out = torch.zeros(N * W * H, dtype=torch.int32, device="cuda")
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
def run(tuning, reps=3):
    host.setTuning(**tuning)
    out.zero_()
    stats = torch.zeros(6, dtype=torch.int64, device="cuda")
    host.render_views(out, W, H, nn, cams, matcap, stats=stats); torch.cuda.synchronize()
    st = stats.cpu().tolist()
    frame = out.cpu().numpy().copy()
    a.record()
    for _ in range(reps):
        host.render_views(out, W, H, nn, cams, matcap)
    b.record(); torch.cuda.synchronize()
    host.setTuning()
    return frame, a.elapsed_time(b) / reps / N, st
ref, ms, st = run({})
print(f"default: {ms:.3f} ms per frame", flush=True)
for t in variants:
    f, ms, st = run(t)
    print(f"{t}: {ms:.3f} ms per frame, bytes equal: {bool(np.array_equal(f, ref))}, queries {st[0]}, evicted {st[3]}", flush=True)

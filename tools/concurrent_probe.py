#!/usr/bin/env python
"""Development probe: V views of each of the five shapes at 512x512 (a rank's share of BASELINE config 5), the five calls one after the
other on one stream against the five calls side by side on five streams with nsdf_tuning.max_ctas = SMs / 5.
usage: concurrent_probe.py [V=8] [W H]"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import host
V = int(sys.argv[1]) if len(sys.argv) > 1 else 8
W, H = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (512, 512)
SHAPES = ["car_1", "plane_1", "plane_2", "plane_3", "3a3d4a90a2db90b4203936772104a82d.obj"]
nets = []
for s in SHAPES:
    nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", s + ".h5")); nets.append(nn)
matcap = host.Image(); assert matcap.loadPNG(os.path.join(ROOT, "data", "matcaps", "blue.png"))
host.copyStaticSettings(host.COLOR_MATCAP, 3)
host.setSceneOptions(host.SCENE_MANYSPHERE, 6000, 0, 0, host.MATH_REFERENCE)
cams = host.view_cameras([(0.0, k * (360.0 / 64) * (64 // V)) for k in range(V)])
sms = torch.cuda.get_device_properties(0).multi_processor_count
outs = [torch.zeros(V * W * H, dtype=torch.int32, device="cuda") for _ in SHAPES]
refs = [torch.zeros(V * W * H, dtype=torch.int32, device="cuda") for _ in SHAPES]
streams = [torch.cuda.Stream() for _ in SHAPES]
def sequential(bufs):
    for nn, o in zip(nets, bufs): host.render_views(o, W, H, nn, cams, matcap)
def concurrent(bufs, share):
    host.setTuning(max_ctas=share)
    start = torch.cuda.Event(); start.record()
    for nn, o, st in zip(nets, bufs, streams):
        st.wait_event(start)
        with torch.cuda.stream(st): host.render_views(o, W, H, nn, cams, matcap)
    for st in streams: torch.cuda.current_stream().wait_stream(st)
    host.setTuning()
def timed(fn, reps=5):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps
t_seq = timed(lambda: sequential(refs))
print(f"{V} views x 5 shapes at {W}x{H}: one call after the other {t_seq:.2f} ms = {5 * V / t_seq * 1e3:.0f} frames/s", flush=True)
for share in (sms // 5, sms // 4, sms // 3, sms // 2):
    for o in outs: o.fill_(0x55)
    t = timed(lambda: concurrent(outs, share))
    eq = all(bool(torch.equal(o, r)) for o, r in zip(outs, refs))
    print(f"  side by side on 5 streams, max_ctas {share:3d}: {t:.2f} ms = {5 * V / t * 1e3:.0f} frames/s, bytes equal: {eq}", flush=True)

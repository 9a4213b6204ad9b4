#!/usr/bin/env python
"""Development probe: the tracer's timeline of long rays (NSDF_COOP_TIMELINE=1): when the pixels ran out, when the launch
ended, and for the longest rays when they were pushed, popped and how fast they stepped (ns per step between events)."""
import ctypes as C, os, sys
os.environ["NSDF_LIB"] = "dev"      # the timeline hooks exist in the development build only (make -C cudaneuralrender_b200/csrc dev)
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import host, _capi as capi
model, W, H = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
scene, ry, tile_world = host.SCENE_TANH, 0.0, 1
for kv in sys.argv[4:]:
    k, v = kv.split("=")
    if k == "SCENE": scene = int(v)          # 0 tanh, 1 manySphere (the reference's HEAD)
    elif k == "RY": ry = float(v)
    elif k == "TILE_WORLD": tile_world = int(v)   # time rank 0's share of a tile partition
    else: os.environ[k] = v
nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", model + ".h5"))
matcap = host.Image(); assert matcap.loadPNG(os.path.join(ROOT, "data", "matcaps", "blue.png"))
host.copyStaticSettings(host.COLOR_MATCAP, 3)
host.updateViewMatrices((0.0, ry, 0.0), (0.0, 0.0, -2.0), 0)
host.setSceneOptions(scene, 6000, 0, 0, host.MATH_REFERENCE)
frame = torch.zeros(W * H, dtype=torch.int32, device="cuda")
for _ in range(3): host.render_kernel(frame, W, H, nn, matcap, tile_rank=0, tile_world=tile_world)
torch.cuda.synchronize()
os.environ["NSDF_COOP_TIMELINE"] = "1"
L = capi.lib()
n = 4 + 4 * 32768
buf = np.zeros(n, dtype=np.uint64)
L.nsdf_dev_timeline_read.argtypes = [C.c_void_p, C.c_int]
L.nsdf_dev_timeline_read(buf.ctypes.data, n)     # clear
host.render_kernel(frame, W, H, nn, matcap, tile_rank=0, tile_world=tile_world)
assert L.nsdf_dev_timeline_read(buf.ctypes.data, n) == 0
t0, t_dry, t_end, nev = (int(v) for v in buf[:4])
print(f"launch {(t_end - t0) / 1e6:.2f} ms; last warp out of pixels at {(t_dry - t0) / 1e6:.2f} ms; events {nev}")
ev = buf[4:4 + 4 * min(nev, 32768)].reshape(-1, 4).astype(np.int64)
ev = ev[np.argsort(ev[:, 0])]
pushes = ev[ev[:, 3] == 100]; pops = ev[ev[:, 3] == 101]
print('express pushes', int((ev[:, 3] == 102).sum()), 'express pops', int((ev[:, 3] == 103).sum()))
if len(pushes):
    pt = (pushes[:, 0] - t0) / 1e6
    print(f"pushes {len(pushes)}: first {pt.min():.2f} ms, median {np.median(pt):.2f}, last {pt.max():.2f}")
if len(pops):
    pt = (pops[:, 0] - t0) / 1e6
    print(f"pops {len(pops)}: first {pt.min():.2f} ms, median {np.median(pt):.2f}, last {pt.max():.2f}")
    push_t = {int(p[1]): int(p[0]) for p in pushes}
    waits = np.array([(int(p[0]) - push_t[int(p[1])]) / 1e6 for p in pops if int(p[1]) in push_t])
    if len(waits): print(f"queue wait ms: median {np.median(waits):.3f}, p90 {np.percentile(waits, 90):.3f}, max {waits.max():.3f}")
last = {}
for t, pix, steps, kind in ev:
    last[pix] = max(last.get(pix, 0), steps)
longest = sorted(last, key=lambda p: -last[p])[:6]
for pix in longest:
    rows = ev[ev[:, 1] == pix]
    print(f"ray pix {pix} (x {pix % W}, y {pix // W}), last logged step {last[pix]}:")
    prev = None
    for t, _, steps, kind in rows:
        rate = "" if prev is None or steps == prev[1] else f"  {(t - prev[0]) / (steps - prev[1]):.0f} ns/step"
        if kind in (100, 101, 102, 103, 104) or steps % 1024 < 64 or steps < 1100:
            print(f"   t {(t - t0) / 1e6:7.3f} ms  steps {steps:5d}  { {100: 'pushed', 101: 'popped', 102: 'pushed to express', 103: 'popped by express lane', 104: 'handed over (tail)'}.get(int(kind), 'rays in warp ' + str(kind)) }{rate}")
        prev = (t, steps)

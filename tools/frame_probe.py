#!/usr/bin/env python
"""Development probe: device time of one frame through render_kernel for a list of step caps and NSDF_* knob settings
(where the bulk of a frame ends and its serial tail begins).  usage: frame_probe.py model W H [KEY=VAL,KEY=VAL ...]"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import host, _capi as capi
model, W, H = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
settings = sys.argv[4:] or [""]
nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", model + ".h5"))
matcap = host.Image(); assert matcap.loadPNG(os.path.join(ROOT, "data", "matcaps", "blue.png"))
host.copyStaticSettings(host.COLOR_MATCAP, 3)
host.updateViewMatrices((0.0, 0.0, 0.0), (0.0, 0.0, -2.0), 0)
frame = torch.zeros(W * H, dtype=torch.int32, device="cuda")
stats = torch.zeros(6, dtype=torch.int64, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ref = {}
FLAGS = getattr(capi, os.environ["PROBE_FLAGS"]) if os.environ.get("PROBE_FLAGS") else 0   # e.g. RENDER_TC_FARFIELD
for setting in settings:
    keys = []
    for kv in filter(None, setting.split(",")):
        k, v = kv.split("="); os.environ[k] = v; keys.append(k)
    for ms_cap in (int(s) for s in os.environ.get("PROBE_STEPS", "256,1024,6000").split(",")):
        host.setSceneOptions(host.SCENE_TANH, ms_cap, FLAGS, 0)
        for _ in range(2): host.render_kernel(frame, W, H, nn, matcap)
        ts = []
        for _ in range(5):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); host.render_kernel(frame, W, H, nn, matcap); b.record(); b.synchronize()
            ts.append(a.elapsed_time(b))
        stats.zero_(); host.render_kernel(frame, W, H, nn, matcap, stats=stats); torch.cuda.synchronize()
        s = stats.cpu().tolist()
        img = frame.clone()
        same = torch.equal(img, ref[ms_cap]) if ms_cap in ref else None
        ref.setdefault(ms_cap, img)
        print(f"[{setting or 'default'}] max_steps {ms_cap}: min {min(ts):.2f} ms  median {sorted(ts)[2]:.2f} ms  queries {s[0]} evicted {s[3]} "
              f"same frame as first setting: {same}", flush=True)
    for k in keys: del os.environ[k]

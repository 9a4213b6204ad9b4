#!/usr/bin/env python
"""Development sweep: device time of a few frames under lists of tracer knobs.  lower-case names are nsdf_tuning fields (any
build); NSDF_* names are environment variables, read per call by the DEVELOPMENT build only (lib=dev as first argument, after
`make -C cudaneuralrender_b200/csrc dev`).  usage: sweep.py [lib=dev] [reps=N] [cases=a,b,..] 'k1=v1 k2=v2' 'k1=v3' ...
cases: tanh0 (plane_1 1080p tanh view 0), head0, head8 (ry=45), head24 (ry=135), car512 (car_1 512^2 HEAD), k4 (3a3d 4K HEAD)"""
import os, sys
if "lib=dev" in sys.argv[1:]:
    os.environ["NSDF_LIB"] = "dev"
    sys.argv.remove("lib=dev")
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import host

CASES = {"tanh0": ("plane_1", 1920, 1080, host.SCENE_TANH, 0.0), "head0": ("plane_1", 1920, 1080, host.SCENE_MANYSPHERE, 0.0),
         "head8": ("plane_1", 1920, 1080, host.SCENE_MANYSPHERE, 45.0), "head24": ("plane_1", 1920, 1080, host.SCENE_MANYSPHERE, 135.0),
         "car512": ("car_1", 512, 512, host.SCENE_MANYSPHERE, 0.0),
         "k4": ("3a3d4a90a2db90b4203936772104a82d.obj", 3840, 2160, host.SCENE_MANYSPHERE, 0.0),
         "k4r8": ("3a3d4a90a2db90b4203936772104a82d.obj", 3840, 2160, host.SCENE_MANYSPHERE, 0.0),     # rank 0's share of an 8-way tile partition
         "k4r4": ("3a3d4a90a2db90b4203936772104a82d.obj", 3840, 2160, host.SCENE_MANYSPHERE, 0.0),
         "k4r8t": ("3a3d4a90a2db90b4203936772104a82d.obj", 3840, 2160, host.SCENE_TANH, 0.0)}
WORLD = {"k4r8": 8, "k4r4": 4, "k4r8t": 8}
reps, cases, configs = 5, ["tanh0", "head0", "head8"], []
for a in sys.argv[1:]:
    if a.startswith("reps="): reps = int(a[5:])
    elif a.startswith("cases="): cases = a[6:].split(",")
    else: configs.append(a)
if not configs: configs = [""]
matcap = host.Image(); assert matcap.loadPNG(os.path.join(ROOT, "data", "matcaps", "blue.png"))
host.copyStaticSettings(host.COLOR_MATCAP, 3)
nets = {}
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
base_env = dict(os.environ)
ref = {}
for cfg in configs:
    for k in list(os.environ):
        if k.startswith("NSDF_") and k not in base_env: del os.environ[k]
    tuning = {}
    for kv in cfg.split():
        k, v = kv.split("=")
        if k.islower(): tuning[k] = int(v)
        else: os.environ[k] = v
    host.setTuning(**tuning)
    row = []
    for c in cases:
        model, W, H, scene, ry = CASES[c]
        if model not in nets:
            nets[model] = host.NeuralNetwork(); assert nets[model].load(os.path.join(ROOT, "data", "neuralGeometries", model + ".h5"))
        nn = nets[model]
        host.updateViewMatrices((0.0, ry, 0.0), (0.0, 0.0, -2.0), 0)
        host.setSceneOptions(scene, 6000, 0, 0, host.MATH_REFERENCE)
        frame = torch.zeros(W * H, dtype=torch.int32, device="cuda")
        tw = WORLD.get(c, 1)
        for _ in range(2): host.render_kernel(frame, W, H, nn, matcap, tile_world=tw)
        ts = []
        for _ in range(reps):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); host.render_kernel(frame, W, H, nn, matcap, tile_world=tw); b.record(); torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        img = frame.cpu().numpy()
        if c not in ref: ref[c] = img
        same = bool(np.array_equal(img, ref[c]))
        row.append(f"{c} {np.mean(ts):6.2f} (min {min(ts):6.2f}){'' if same else ' DIFFERENT-BYTES'}")
    print(f"[{cfg or 'defaults'}]  " + "  |  ".join(row), flush=True)

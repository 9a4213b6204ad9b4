#!/usr/bin/env python
"""Workload for NVLink-traffic captures: ONE process drives all GPUs of the box (nsdf_multi) and renders one 3840x2160 frame of
3a3d4a90...obj tile-partitioned, every device's tracer storing its finished tiles into GPU 0's frame over NVLink."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import host, multigpu
from oracle import oracle
nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", "3a3d4a90a2db90b4203936772104a82d.obj.h5"))
host.copyStaticSettings(host.COLOR_MATCAP, 3)
host.setSceneOptions(host.SCENE_MANYSPHERE, 6000, 0, 0, host.MATH_REFERENCE)
mc = oracle.load_matcap_png(os.path.join(ROOT, "data", "matcaps", "blue.png"))
n = torch.cuda.device_count()
W, H = 3840, 2160
mr = multigpu.MultiRenderer(nn, list(range(n)))
out = torch.zeros(W * H, dtype=torch.int32, device="cuda:0")
cams = host.view_cameras([(0.0, 0.0)])
for _ in range(2):
    st = mr.render(out, W, H, cams, mc)
print("ok", n, "devices", st["pixels"], "pixels", int((out != 0).sum()), "hit pixels")

#!/usr/bin/env python
"""Small fixed workload for ncu captures: one batch of V 1080p views of plane_1 (reference HEAD scene) through nsdf_render_views,
then (optional) one dense 256^3 sweep per listed precision.  usage: ncu_target.py [views=4] [dense=0,1,3] [nsdf_tuning knob=value ...]"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import host
V, dense = 4, []
for a in sys.argv[1:]:
    if a.startswith("views="): V = int(a[6:])
    if a.startswith("dense="): dense = [int(x) for x in a[6:].split(",") if x]
tuning = {a.split("=")[0]: int(a.split("=")[1]) for a in sys.argv[1:] if "=" in a and a.split("=")[0] not in ("views", "dense")}
W, H = 1920, 1080
nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", "plane_1.h5"))
matcap = host.Image(); assert matcap.loadPNG(os.path.join(ROOT, "data", "matcaps", "blue.png"))
host.copyStaticSettings(host.COLOR_MATCAP, 3)
host.setSceneOptions(host.SCENE_MANYSPHERE, 6000, 0, 0, host.MATH_REFERENCE)
if tuning:
    host.setTuning(**tuning)
if V > 0:
    out = torch.zeros(V * W * H, dtype=torch.int32, device="cuda")
    cams = host.view_cameras([(0.0, 5.625 * k) for k in range(V)])
    for _ in range(2):
        host.render_views(out, W, H, nn, cams, matcap)
    torch.cuda.synchronize()
car = host.NeuralNetwork(); assert car.load(os.path.join(ROOT, "data", "neuralGeometries", "car_1.h5"))
for prec in dense:
    for _ in range(2):
        car.forward_lattice(256, precision=prec)
torch.cuda.synchronize()
print("ok")

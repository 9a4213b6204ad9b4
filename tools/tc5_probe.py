#!/usr/bin/env python
"""Development probe (dev library): the dense 256^3 sweep of car_1 on the fp16 / bf16 tensor-core rungs, schedule v2 against the
v5 configurations (NSDF_TC_KERNEL / NSDF_TC5_CFG), with bit-equality of the outputs against v2.  usage: tc5_probe.py"""
import os, sys
os.environ.setdefault("NSDF_LIB", "dev")
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import host
nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", "car_1.h5"))
res = 256
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def run(prec, env):
    for k, v in env.items(): os.environ[k] = str(v)
    y = torch.empty(res ** 3, dtype=torch.float32, device="cuda")
    for _ in range(3): nn.forward_lattice(res, out=y, precision=prec)
    ts = []
    for _ in range(7):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); nn.forward_lattice(res, out=y, precision=prec); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return y, res ** 3 / (min(ts) * 1e-3) / 1e9, res ** 3 / (np.mean(ts) * 1e-3) / 1e9
for name, prec in (("fp16", host.PREC_TC_FP16),):
    ref, best, mean = run(prec, {"NSDF_TC_KERNEL": 2})
    print(f"{name} v2            : {best:.2f} Gq/s best, {mean:.2f} mean", flush=True)
    for cfg, label in ((0, "5 groups x 2 CTAs"), (1, "4 groups x 2 CTAs"), (2, "2 groups x 4 CTAs"), (3, "1 group  x 8 CTAs")):
        try:
            y, best, mean = run(prec, {"NSDF_TC_KERNEL": 5, "NSDF_TC5_CFG": cfg})
            print(f"{name} v5 {label}: {best:.2f} Gq/s best, {mean:.2f} mean, bits equal v2: {bool(torch.equal(y.view(torch.int32), ref.view(torch.int32)))}", flush=True)
        except Exception as e:
            print(f"{name} v5 {label}: failed: {e}", flush=True)
    for cfg, label in ((0, "5 groups x 1 CTA, 10 tiles"), (1, "2 x 2, 8 tiles"), (2, "1 x 4, 8 tiles"), (3, "4 x 1, 8 tiles")):
        try:
            y, best, mean = run(prec, {"NSDF_TC_KERNEL": 6, "NSDF_TC6_CFG": cfg})
            print(f"{name} v6 {label}: {best:.2f} Gq/s best, {mean:.2f} mean, bits equal v2: {bool(torch.equal(y.view(torch.int32), ref.view(torch.int32)))}", flush=True)
        except Exception as e:
            print(f"{name} v6 {label}: failed: {e}", flush=True)

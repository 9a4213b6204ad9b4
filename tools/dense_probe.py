#!/usr/bin/env python
"""Development probe: dense 256^3 car_1 sweep through the fp32 evaluators (bit-equality against the default rung + Gq/s).
usage: dense_probe.py 'NSDF_FP32_WARP_R=32 NSDF_FP32_WARP_THREADS=512' ...   (each argument = one configuration of the warp rung)"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import host
nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", "car_1.h5"))
res = 256
ref = nn.forward_lattice(res, precision=host.PREC_FP32).clone()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def run(prec, label):
    y = torch.empty(res ** 3, dtype=torch.float32, device="cuda")
    for _ in range(2): nn.forward_lattice(res, out=y, precision=prec)
    ts = []
    for _ in range(5):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); nn.forward_lattice(res, out=y, precision=prec); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    eq = bool(torch.equal(y.view(torch.int32), ref.view(torch.int32)))
    print(f"[{label}] {res**3 / (min(ts) * 1e-3) / 1e9:.3f} Gq/s (min {min(ts):.3f} ms, mean {np.mean(ts):.3f})  bit-equal {eq}", flush=True)
run(host.PREC_FP32, "default fp32 rung")
for cfg in sys.argv[1:]:
    for kv in cfg.split():
        k, v = kv.split("="); os.environ[k] = v
    try:
        run(host.PREC_FP32_WARP, cfg)
    except Exception as e:
        print(f"[{cfg}] failed: {e}")

#!/usr/bin/env python
"""Development probe (dev library): the fp16x3 rung on the v4 configurations (NSDF_TC4_CFG), rate and bits against cfg 1."""
import os, sys
os.environ.setdefault("NSDF_LIB", "dev")
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import host
nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", "car_1.h5"))
res = 256
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def run(cfg):
    os.environ["NSDF_TC_KERNEL"] = "4"; os.environ["NSDF_TC4_CFG"] = str(cfg)
    y = torch.empty(res ** 3, dtype=torch.float32, device="cuda")
    for _ in range(3): nn.forward_lattice(res, out=y, precision=host.PREC_TC_FP16X3)
    ts = []
    for _ in range(7):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); nn.forward_lattice(res, out=y, precision=host.PREC_TC_FP16X3); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return y, res ** 3 / (min(ts) * 1e-3) / 1e9
ref, r = run(1)
print(f"cfg 1 (7 groups x 1 CTA, all warps poll): {r:.2f} Gq/s", flush=True)
for cfg in (0, 2, 5, 6, 7):
    try:
        y, r = run(cfg)
        print(f"cfg {cfg}: {r:.2f} Gq/s, bits equal cfg 1: {bool(torch.equal(y.view(torch.int32), ref.view(torch.int32)))}", flush=True)
    except Exception as e:
        print(f"cfg {cfg}: failed {e}", flush=True)

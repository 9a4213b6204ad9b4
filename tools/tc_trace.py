#!/usr/bin/env python
"""Development probe: per-step clock64 timeline of CTA 0 of the dense TC kernel (stamps: tile start,
then per step: before sync, after sync, after issue, after mbarrier wait)."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import _capi, host  # noqa: E402

nn = host.NeuralNetwork()
assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", "car_1.h5"))
buf = torch.zeros(1024, dtype=torch.int64, device="cuda")
lib = _capi.lib()
lib.nsdf_debug_tc_trace.argtypes = [C.c_void_p]
lib.nsdf_debug_tc_trace.restype = None
out = torch.empty(256 ** 3, dtype=torch.float32, device="cuda")
prec = {"fp16": host.PREC_TC_FP16, "fp16x3": host.PREC_TC_FP16X3}[sys.argv[1] if len(sys.argv) > 1 else "fp16"]
nn.forward_lattice(256, precision=prec, out=out)
lib.nsdf_debug_tc_trace(buf.data_ptr())
nn.forward_lattice(256, precision=prec, out=out)
torch.cuda.synchronize()
lib.nsdf_debug_tc_trace(None)
t = buf.cpu().numpy()
for who, off in (("thread0", 0), ("thread96", 512)):
    s = t[off:off + 500]
    s = s[s > 0]
    per_tile = 1 + 4 * 9
    # skip first 3 tiles (cold), average over next tiles
    tiles = (len(s) // per_tile)
    rows = []
    for k in range(3, tiles - 1):
        a = s[k * per_tile:(k + 1) * per_tile + 1]
        rows.append(np.diff(a))
    d = np.mean(np.array(rows), axis=0)
    print(who, "tiles", tiles, "cycles/tile", d.sum())
    print("  store_query", d[0])
    for st in range(9):
        q = d[1 + 4 * st: 5 + 4 * st]
        print(f"  step {st}: sync {q[0]:.0f} issue {q[1]:.0f} wait {q[2]:.0f} epilogue {q[3]:.0f}  total {q.sum():.0f}")

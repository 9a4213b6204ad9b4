#!/usr/bin/env python
"""Development probe: rate of schedule v5 with parts of the step compiled out (-DNSDF_TC5_DIAG=bits: 1 no conversion, 2 no
tcgen05.st, 4 no tcgen05.ld, 8 no bias UMMA; results are wrong under it, only the rate is read).  One library per variant:
usage: tc5_diag.py LIB.so [cfg ...]   (run once per library: the library is chosen at import)"""
import os, sys
os.environ["NSDF_LIB_PATH"] = os.path.abspath(sys.argv[1])
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import host
nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", "car_1.h5"))
res = 256
y = torch.empty(res ** 3, dtype=torch.float32, device="cuda")
def run(env):
    for k, v in env.items(): os.environ[k] = str(v)
    for _ in range(3): nn.forward_lattice(res, out=y, precision=host.PREC_TC_FP16)
    ts = []
    for _ in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); nn.forward_lattice(res, out=y, precision=host.PREC_TC_FP16); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return res ** 3 / (min(ts) * 1e-3) / 1e9
out = [f"v2 {run({'NSDF_TC_KERNEL': 2}):.2f}"]
for cfg in [int(a) for a in sys.argv[2:]] or [2, 1, 0]:
    out.append(f"v5 cfg{cfg} {run({'NSDF_TC_KERNEL': 5, 'NSDF_TC5_CFG': cfg}):.2f}")
print(os.path.basename(sys.argv[1]), " | ".join(out), "Gq/s", flush=True)

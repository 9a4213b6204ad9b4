#!/usr/bin/env python
"""Development probe: throughput of the two fp32 evaluators on a dense batch."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import host
nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", "car_1.h5"))
n = 1 << 22
x = torch.from_numpy(np.random.default_rng(0).uniform(-1, 1, (n, 3)).astype(np.float32)).cuda()
out = torch.empty(n, dtype=torch.float32, device="cuda")
for name, prec in (("thread", host.PREC_FP32), ("warp", host.PREC_FP32_WARP)):
    for _ in range(3): nn.forward_tensor(x, prec, out=out)
    torch.cuda.synchronize()
    ts = []
    for _ in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); nn.forward_tensor(x, prec, out=out); b.record(); b.synchronize()
        ts.append(a.elapsed_time(b))
    print(name, "ms", min(ts), "Gq/s", n / min(ts) / 1e6, flush=True)

#!/usr/bin/env python
"""Development probe: the thread-per-point fp32 evaluators (P = 1, 2 and the 4-point shared-memory-activation form)
on a dense batch: throughput and bit-equality of their outputs."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import host
nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", "car_1.h5"))
n = (1 << 24) + 12345
x = torch.from_numpy(np.random.default_rng(0).uniform(-1, 1, (n, 3)).astype(np.float32)).cuda()
ref = None
for name, env in (("P2", {"NSDF_FP32_P": "2"}), ("P1", {"NSDF_FP32_P": "1"}), ("P2s/256x2", {"NSDF_FP32_P": "22"}), ("P4/128", {"NSDF_FP32_P": "4", "NSDF_FP32_P4_THREADS": "128"}),
                  ("P4/256", {"NSDF_FP32_P": "4", "NSDF_FP32_P4_THREADS": "256"}),
                  ("P4/384", {"NSDF_FP32_P": "4", "NSDF_FP32_P4_THREADS": "384"}),
                  ("const P2", {"NSDF_FP32_P": "40"}), ("const P1", {"NSDF_FP32_P": "41"})):
    os.environ.update(env)
    out = torch.zeros(n, dtype=torch.float32, device="cuda")
    for _ in range(3): nn.forward_tensor(x, host.PREC_FP32, out=out)
    torch.cuda.synchronize()
    ts = []
    for _ in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); nn.forward_tensor(x, host.PREC_FP32, out=out); b.record(); b.synchronize()
        ts.append(a.elapsed_time(b))
    if ref is None: ref = out.clone()
    same = bool(torch.equal(out.view(torch.int32), ref.view(torch.int32)))
    print(name, "ms", round(min(ts), 4), "Gq/s", round(n / min(ts) / 1e6, 3), "bit-equal to P2:", same, flush=True)

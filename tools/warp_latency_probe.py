#!/usr/bin/env python
"""Development probe: latency of ONE dependent fp32 network evaluation in the warp-cooperative evaluator (lane = neuron),
as a function of rays per warp (R) and warps per SM.  Each warp evaluates `iters` batches back to back."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import host
nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", "car_1.h5"))
iters = 2000
for R in (1, 2):
    for threads in (32, 128, 512):
        os.environ["NSDF_FP32_WARP_R"] = str(R); os.environ["NSDF_FP32_WARP_THREADS"] = str(threads); os.environ["NSDF_FP32_WARP_CTAS"] = "1"
        n = 148 * (threads // 32) * R * iters
        x = torch.from_numpy(np.random.default_rng(0).uniform(-1, 1, (n, 3)).astype(np.float32)).cuda()
        out = torch.empty(n, dtype=torch.float32, device="cuda")
        for _ in range(2): nn.forward_tensor(x, host.PREC_FP32_WARP, out=out)
        torch.cuda.synchronize(); ts = []
        for _ in range(3):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); nn.forward_tensor(x, host.PREC_FP32_WARP, out=out); b.record(); b.synchronize(); ts.append(a.elapsed_time(b))
        print(f"R={R} warps/SM={threads // 32:2d}: {min(ts) * 1e3 / iters:.3f} us per batch of {R} ({n / min(ts) / 1e6:.2f} Gq/s)", flush=True)

#!/usr/bin/env python
"""Development probe (GPU box): error of each MLP rung against the CPU oracle and dense throughput.

  python tools/tc_probe.py [--n 4194304] [--res 256] [--out gpurun_out/tc_probe.json]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import host  # noqa: E402
from oracle import oracle  # noqa: E402

FLOP = 14592


def time_ms(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[0], ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1 << 20)
    ap.add_argument("--res", type=int, default=256)
    ap.add_argument("--model", default="car_1")
    ap.add_argument("--out", default="")
    ap.add_argument("--precs", default="fp32,fp16,bf16,fp16x3,fp16acc")
    args = ap.parse_args()
    path = os.path.join(ROOT, "data", "neuralGeometries", args.model + ".h5")
    nn = host.NeuralNetwork()
    assert nn.load(path)
    om = oracle.Model.from_h5(path)
    x = np.random.default_rng(0).uniform(-1, 1, (args.n, 3)).astype(np.float32)
    ref = oracle.mlp_forward(om, x).ravel().astype(np.float64)
    xd = torch.from_numpy(x).cuda().contiguous()
    precs = {"fp32": host.PREC_FP32, "fp16": host.PREC_TC_FP16, "bf16": host.PREC_TC_BF16, "fp16x3": host.PREC_TC_FP16X3,
             "fp16acc": host.PREC_TC_FP16_ACC16}
    res = {}
    for name in args.precs.split(","):
        prec = precs[name]
        try:
            y = nn.forward_tensor(xd, prec).cpu().numpy().astype(np.float64)
        except Exception as e:  # noqa: BLE001
            res[name] = {"error": str(e)}
            print(name, "FAILED", e, flush=True)
            continue
        err = np.abs(y - ref)
        out = torch.empty(args.res ** 3, dtype=torch.float32, device="cuda")
        best, med = time_ms(lambda: nn.forward_lattice(args.res, precision=prec, out=out))
        q = args.res ** 3
        best2, med2 = time_ms(lambda: nn.forward_tensor(xd, prec))
        res[name] = {"err_max": float(err.max()), "err_p99": float(np.quantile(err, 0.99)), "err_median": float(np.median(err)),
                     "nan": int(np.isnan(y).sum()),
                     "lattice_ms_best": best, "lattice_ms_median": med, "lattice_gqps": q / best / 1e6,
                     "lattice_tflops": q * FLOP / best / 1e9,
                     "points_ms_best": best2, "points_gqps": args.n / best2 / 1e6}
        print(name, json.dumps(res[name]), flush=True)
    if args.out:
        os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
        json.dump(res, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()

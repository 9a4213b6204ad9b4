#!/usr/bin/env python
"""Development probe: the warp-cooperative fp32 evaluators on a dense batch (8 points per warp with lane = neuron
against the register-tiled 16-point form), and whole frames through the default tracer."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import host
from oracle import oracle
nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", "car_1.h5"))
n = (1 << 23) + 12345
x = torch.from_numpy(np.random.default_rng(0).uniform(-1, 1, (n, 3)).astype(np.float32)).cuda()
os.environ["NSDF_FP32_P"] = "2"
ref = nn.forward_tensor(x, host.PREC_FP32).clone()
for r in ("8", "16", "77", "78", "79"):
    os.environ["NSDF_FP32_WARP_R"] = r
    out = torch.zeros(n, dtype=torch.float32, device="cuda")
    for _ in range(3): nn.forward_tensor(x, host.PREC_FP32_WARP, out=out)
    torch.cuda.synchronize()
    ts = []
    for _ in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); nn.forward_tensor(x, host.PREC_FP32_WARP, out=out); b.record(); b.synchronize()
        ts.append(a.elapsed_time(b))
    same = bool(torch.equal(out.view(torch.int32), ref.view(torch.int32)))
    print("warp R", r, "ms", round(min(ts), 4), "Gq/s", round(n / min(ts) / 1e6, 3), "bit-equal to thread P2:", same, flush=True)
del os.environ["NSDF_FP32_P"]
if os.environ.get("PROBE_DENSE_ONLY"): sys.exit(0)

matcap = oracle.load_matcap_png(os.path.join(ROOT, "data", "matcaps", "blue.png"))
iv, nm = oracle.view_matrices()
host.copyStaticSettings(host.COLOR_MATCAP, 3)
host.copyViewMatrices(iv, nm, 0, 3)
host.setSceneOptions(host.SCENE_TANH, 6000)
for model, (w, h) in (("plane_1", (1920, 1080)), ("car_1", (1920, 1080)), ("plane_1", (3840, 2160))):
    m = host.NeuralNetwork(); assert m.load(os.path.join(ROOT, "data", "neuralGeometries", model + ".h5"))
    ts = []
    for _ in range(6):
        t0 = time.perf_counter(); img, stats = host.render_host(m, w, h, matcap); ts.append(time.perf_counter() - t0)
    print(model, w, h, "render_host ms (min of 6, includes copies)", round(min(ts) * 1e3, 2), "kernel ms", stats.get("ms"), stats, flush=True)

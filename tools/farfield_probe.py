#!/usr/bin/env python
"""Development probe (GPU box): mixed-precision tracer (NSDF_RENDER_TC_FARFIELD) against the bit-exact default:
hit-mask mismatches, fraction of pixels within 1/255 per channel, frame time.  Env: NSDF_FARFIELD_FP16/THR/MARGIN."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import _capi as capi, host  # noqa: E402


def frame(nn, mc, W, H, flags, iters=5):
    host.setSceneOptions(host.SCENE_TANH, 6000, flags, 0)
    fr = torch.zeros(W * H, dtype=torch.int32, device="cuda")
    st = torch.zeros(6, dtype=torch.int64, device="cuda")
    host.render_kernel(fr, W, H, nn, mc, stats=st)
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        host.render_kernel(fr, W, H, nn, mc)
        b.record()
        b.synchronize()
        ts.append(a.elapsed_time(b))
    return fr.cpu().numpy().view(np.uint32), st.cpu().numpy(), min(ts)


def main():
    cases = [("plane_1", 1920, 1080), ("car_1", 512, 512), ("3a3d4a90a2db90b4203936772104a82d.obj", 3840, 2160)]
    for model, W, H in cases:
        nn = host.NeuralNetwork()
        assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", model + ".h5"))
        mc = host.Image()
        mc.loadPNG(os.path.join(ROOT, "data", "matcaps", "blue.png"))
        host.copyStaticSettings(host.COLOR_MATCAP, 3)
        host.updateViewMatrices((0, 0, 0), (0, 0, -2), 0)
        ref, s0, t0 = frame(nn, mc, W, H, 0)
        img, s1, t1 = frame(nn, mc, W, H, capi.RENDER_TC_FARFIELD)
        hit0, hit1 = ref != 0, img != 0
        both = hit0 & hit1
        ch = lambda a, k: ((a >> (8 * k)) & 255).astype(np.int32)  # noqa: E731
        d = np.max(np.stack([np.abs(ch(ref, k) - ch(img, k)) for k in range(3)]), axis=0)
        out = {"model": model, "size": [W, H], "ms_exact": t0, "ms_farfield": t1, "speedup": t0 / t1,
               "queries_exact": int(s0[0]), "queries_farfield": int(s1[0]), "handed_over": int(s1[3]),
               "hit_pixels": int(hit0.sum()), "hit_mask_mismatches": int((hit0 != hit1).sum()),
               "identical_px_frac": float((ref == img).mean()),
               "hit_px_within_1_255": float((d[both] <= 1).mean()) if both.any() else 1.0,
               "hit_px_identical": float((d[both] == 0).mean()) if both.any() else 1.0, "max_channel_diff": int(d[both].max()) if both.any() else 0}
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()

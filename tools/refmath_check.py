#!/usr/bin/env python
"""Compares the product in NSDF_MATH_REFERENCE mode with outputs of the reference's own CUDA code (golden vectors made by
tools/make_ref_golden.py): frames of render_kernel at HEAD and of the tanh variant, and the analytic-SDF device functions.
Prints one JSON object.  GPU box:  python tools/refmath_check.py [--golden DIR]   (default tests/golden)"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

# scene programs equal to the functions oracle/ref_sdf_probe.cu evaluates, in its order
BOX = [-0.5, -0.2, 0.4, 0.1, 0.1, 0.1]


def sdf_programs(c):
    return [
        [(c.PRIM_SPHERE, c.COMB_SET, [0, 0, 0, 0.9])],
        [(c.PRIM_BOX, c.COMB_SET, BOX + [0.0])],
        [(c.PRIM_BOX, c.COMB_SET, BOX + [0.04])],
        [(c.PRIM_PLANE, c.COMB_SET, [0.5])],
        [(c.PRIM_CYLINDER, c.COMB_SET, BOX)],
        [(c.PRIM_TANH, c.COMB_SET, []), (c.MOD_DISPLACE, 0, [5.0, 0.05])],
        [(c.PRIM_TANH, c.COMB_SET, []), (c.MOD_ROUND, 0, [0.04])],
        [(c.PRIM_NSDF, c.COMB_SET, []), (c.MOD_ONION, 0, [0.02])],
        [(c.PRIM_TANH, c.COMB_SET, []), (c.PRIM_CYLINDER, c.COMB_SMOOTH_SUBTRACT_FROM, BOX + [0.0, 0.05])],
        [(c.PRIM_NSDF, c.COMB_SET, []), (c.MOD_MANY_CYLINDER_CUT, 0, [])],
        [(c.PRIM_NSDF, c.COMB_SET, []), (c.MOD_MANY_SPHERE, 0, [0.0])],
        [(c.PRIM_NSDF, c.COMB_SET, []), (c.MOD_MANY_SPHERE, 0, [1.0])],
        [(c.PRIM_NSDF, c.COMB_SET, []), (c.PRIM_SPHERE, c.COMB_SMOOTH_UNION, [0, 0, 0, 0.5, 0, 0, 0, 0.05])],
        [(c.PRIM_TANH, c.COMB_SET, [])],
        [(c.PRIM_NSDF, c.COMB_SET, []), (c.PRIM_SPHERE, c.COMB_SUBTRACT, [0, 0, 0, 0.5])],
        [(c.PRIM_NSDF, c.COMB_SET, []), (c.PRIM_PLANE, c.COMB_UNION, [0.5]), (c.PRIM_SPHERE, c.COMB_INTERSECT, [0, 0, 0, 0.9])],
    ]


def compare_frames(mine, ref):
    hm, hr = mine != 0, ref != 0
    mism = hm != hr
    mism[-1, -1] = False           # the reference reads a stale SDF slot for the last pixel
    both = hm & hr
    both[-1, -1] = False
    ch = lambda v: np.stack([(v >> s) & 255 for s in (0, 8, 16)], -1).astype(np.int32)  # noqa: E731
    diff = np.abs(ch(mine) - ch(ref))[both]
    neq = mine != ref
    neq[-1, -1] = False
    return {"pixels": int(mine.size), "hit_pixels": int(both.sum()), "mask_mismatches": int(mism.sum()),
            "differing_pixels": int(neq.sum()), "frac_within_1_255": float((diff.max(1) <= 1).mean()) if diff.size else 1.0,
            "frac_identical": float(1.0 - neq.sum() / max(1, both.sum())), "max_channel_diff": int(diff.max()) if diff.size else 0}


def render_case(host, torch, m, cam, scene, math, matcaps, geom):
    nn = host.NeuralNetwork()
    assert nn.load(os.path.join(geom, m["model"] + ".h5"))
    matcap = None
    if m["matcap"]:
        matcap = host.Image()
        assert matcap.loadPNG(os.path.join(matcaps, m["matcap"]))
    host.copyStaticSettings(host.COLOR_MATCAP if matcap is not None else host.COLOR_FACING, 3)
    host.copyViewMatrices(cam[:12], cam[12:], m["frame"], 3)
    host.setSceneOptions(scene, 6000, 0, 0, math)
    out = torch.zeros(m["W"] * m["H"], dtype=torch.int32, device="cuda")
    host.render_kernel(out, m["W"], m["H"], nn, matcap)
    return out.cpu().numpy().view(np.uint32).reshape(m["H"], m["W"])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--golden", default=os.path.join(ROOT, "tests", "golden"))
    args = ap.parse_args()
    import torch
    from cudaneuralrender_b200 import host, _capi as capi
    matcaps, geom = os.path.join(ROOT, "data", "matcaps"), os.path.join(ROOT, "data", "neuralGeometries")
    meta = json.load(open(os.path.join(args.golden, "ref_cuda_meta.json")))
    res = {"frames": {}, "frames_tanh": {}, "sdf": {}}
    for key, fname, scene in (("frames", "ref_cuda_frames.npz", host.SCENE_MANYSPHERE), ("frames_tanh", "ref_cuda_frames_tanh.npz", host.SCENE_TANH)):
        path = os.path.join(args.golden, fname)
        if not os.path.exists(path) or key not in meta:
            continue
        f = np.load(path)
        for k in sorted(meta[key]):
            m = meta[key][k]
            ref, cam = f[k + "/image"], f[k + "/camera"]
            r = {"case": f"{m['model']} {m['W']}x{m['H']} matcap={m['matcap']} frame={m['frame']}"}
            for label, math in (("reference_math", host.MATH_REFERENCE), ("deterministic_math", host.MATH_DETERMINISTIC)):
                r[label] = compare_frames(render_case(host, torch, m, cam, scene, math, matcaps, geom), ref)
            res[key][k] = r
    path = os.path.join(args.golden, "ref_cuda_sdf.npz")
    if os.path.exists(path):
        d = np.load(path)
        xyz = torch.from_numpy(d["xyz"]).cuda().contiguous()
        nsdf = torch.from_numpy(d["nsdf"]).cuda().contiguous()
        names = meta.get("sdf_probe", {}).get("functions", [])
        for frame in (0, 40):
            ref = d[f"frame{frame}"]
            for i, prog in enumerate(sdf_programs(capi)):
                r = {}
                for label, math in (("reference_math", host.MATH_REFERENCE), ("deterministic_math", host.MATH_DETERMINISTIC)):
                    y = host.scene_eval(prog, xyz, nsdf, frame, math).cpu().numpy()
                    neq = y.view(np.uint32) != ref[i].view(np.uint32)
                    neq &= ~(np.isnan(y) & np.isnan(ref[i]))
                    r[label] = {"differing": int(neq.sum()), "max_abs": float(np.nanmax(np.abs(y.astype(np.float64) - ref[i]))) if neq.any() else 0.0}
                res["sdf"][f"frame{frame}/{i}:{names[i] if i < len(names) else ''}"] = r
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Runs the REFERENCE's own CUDA code (oracle/_ref/ref_driver, built by oracle/build_ref.sh from
/root/reference) on this GPU and records its outputs as golden vectors:

  tests/golden/ref_cuda_mlp.npz     per model: points (seeded) and nn.forward outputs (pre-tanh)
  tests/golden/ref_cuda_frames.npz  frames of render_kernel at HEAD (sceneSDF = manySphere union), facing
                                    and matcap colouring, plus the camera matrices used
  tests/golden/ref_cuda_frames_tanh.npz  the same through oracle/_ref/ref_driver_tanh: the reference with sceneSDF's two early
                                    returns commented out, so that its own `return tanh(nSDF)` (:229) runs (oracle/build_ref.sh)
  tests/golden/ref_cuda_sdf.npz     the reference's analytic-SDF device functions (:67-196) on seeded points, through
                                    oracle/_ref/ref_sdf_probe (oracle/ref_sdf_probe.cu #includes the reference's file unmodified)
  tests/golden/ref_cuda_meta.json   timings of the reference on this box, sizes, command lines

Usage (GPU box):  python tools/make_ref_golden.py --out gpurun_out/ref_golden
The committed copies under tests/golden/ were produced by exactly this script.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle  # noqa: E402  (camera matrices only: plain trigonometry, src/main.cpp:207-222)

DRIVER = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
DRIVER_TANH = os.path.join(ROOT, "oracle", "_ref", "ref_driver_tanh")
SDF_PROBE = os.path.join(ROOT, "oracle", "_ref", "ref_sdf_probe")
MODELS = ["car_1", "plane_1", "plane_2", "plane_3", "3a3d4a90a2db90b4203936772104a82d.obj"]
KAT_POINTS = [[0, 0, 0], [0.1, 0.2, 0.3], [0.5, -0.25, 0.125], [-0.3, 0.05, 0.4], [0.9, 0.9, -0.9]]


def model_path(name):
    return os.path.join(ROOT, "data", "neuralGeometries", name + ".h5")


def run(cmd, timeout=600):
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
    if r.returncode != 0:
        raise RuntimeError(f"{' '.join(cmd)} failed ({r.returncode}):\n{r.stdout[-2000:]}\n{r.stderr[-2000:]}")
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    return json.loads(line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "ref_golden"))
    ap.add_argument("--n", type=int, default=20000)
    ap.add_argument("--big", action="store_true", help="also time 1 000 000 points and a 512x512 frame")
    args = ap.parse_args()
    os.makedirs(args.out, exist_ok=True)
    tmp = tempfile.mkdtemp()
    meta = {"driver": "oracle/_ref/ref_driver (reference TUs compiled for sm_100, oracle/build_ref.sh)", "mlp": {}, "frames": {}}
    mlp = {}
    for name in MODELS:
        rng = np.random.default_rng(abs(hash(name)) % 1000 if False else MODELS.index(name) + 11)
        x = rng.uniform(-1, 1, (args.n, 3)).astype(np.float32)
        x[:len(KAT_POINTS)] = np.asarray(KAT_POINTS, dtype=np.float32)
        px, py = os.path.join(tmp, "x.f32"), os.path.join(tmp, "y.f32")
        x.tofile(px)
        info = run([DRIVER, "mlp", model_path(name), px, py, "2"])
        y = np.fromfile(py, dtype=np.float32)
        assert y.size == args.n
        mlp[name + "/x"] = x
        mlp[name + "/y"] = y
        meta["mlp"][name] = info
        print(name, "mlp", info, flush=True)
    np.savez_compressed(os.path.join(args.out, "ref_cuda_mlp.npz"), **mlp)

    def record_frames(driver, cases, fname, meta_key):
        frames = {}
        meta[meta_key] = {}
        for i, (name, W, H, matcap, rx, ry, frame) in enumerate(cases):
            iv, nm = oracle.view_matrices(rx, ry)
            cam = np.concatenate([np.asarray(iv, np.float32).ravel(), np.asarray(nm, np.float32).ravel()])
            pc, po = os.path.join(tmp, "cam.f32"), os.path.join(tmp, "out.u32")
            cam.tofile(pc)
            mc = os.path.join(ROOT, "data", "matcaps", matcap) if matcap else "-"
            info = run([driver, "render", model_path(name), str(W), str(H), mc, pc, str(frame), po, "1"], timeout=900)
            img = np.fromfile(po, dtype=np.uint32).reshape(H, W)
            key = f"case{i}"
            frames[key + "/image"] = img
            frames[key + "/camera"] = cam
            meta[meta_key][key] = {"model": name, "W": W, "H": H, "matcap": matcap, "rx": rx, "ry": ry, "frame": frame, **info}
            print(meta_key, key, meta[meta_key][key], "nonzero", int((img != 0).sum()), flush=True)
        np.savez_compressed(os.path.join(args.out, fname), **frames)

    # HEAD (sceneSDF = manySphere union): four small frames, the 512x512 car_1 frame of BASELINE config 1 and the 1920x1080
    # plane_1 frame of config 2 (the frame bench.py times)
    cases = [("car_1", 128, 128, "blue.png", 0.0, 0.0, 0), ("car_1", 128, 128, None, 20.0, 135.0, 0),
             ("plane_1", 160, 90, "red.png", 0.0, 30.0, 40), ("plane_3", 96, 96, "blue.png", -15.0, 200.0, 7),
             ("car_1", 512, 512, "blue.png", 0.0, 0.0, 0), ("plane_1", 1920, 1080, "blue.png", 0.0, 0.0, 0)]
    record_frames(DRIVER, cases, "ref_cuda_frames.npz", "frames")
    # the tanh(nSDF) variant of sceneSDF (the reference's line :229 made reachable, oracle/build_ref.sh)
    if os.path.exists(DRIVER_TANH):
        tcases = [("car_1", 128, 128, "blue.png", 0.0, 0.0, 0), ("plane_1", 160, 90, "red.png", 0.0, 30.0, 0),
                  ("3a3d4a90a2db90b4203936772104a82d.obj", 96, 96, None, 10.0, 300.0, 0),
                  ("car_1", 512, 512, "blue.png", 0.0, 0.0, 0), ("plane_1", 1920, 1080, "blue.png", 0.0, 0.0, 0)]
        record_frames(DRIVER_TANH, tcases, "ref_cuda_frames_tanh.npz", "frames_tanh")

    # the analytic-SDF device functions of the reference (:67-196) on seeded points: half of the "network outputs" are the
    # reference's own nn.forward of car_1 at those points, half are uniform in [-0.3, 1.2] (so that every blend band is hit)
    if os.path.exists(SDF_PROBE):
        n = 20000
        rng = np.random.default_rng(77)
        xyz = rng.uniform(-1.2, 1.2, (n, 3)).astype(np.float32)
        px, py = os.path.join(tmp, "sx.f32"), os.path.join(tmp, "sy.f32")
        xyz.tofile(px)
        run([DRIVER, "mlp", model_path("car_1"), px, py, "1"])
        nsdf = np.fromfile(py, dtype=np.float32)
        nsdf[n // 2:] = rng.uniform(-0.3, 1.2, n - n // 2).astype(np.float32)
        ps, po = os.path.join(tmp, "ss.f32"), os.path.join(tmp, "so.f32")
        nsdf.tofile(ps)
        sdf = {"xyz": xyz, "nsdf": nsdf}
        for frame in (0, 40):
            info = run([SDF_PROBE, px, ps, str(frame), po])
            sdf[f"frame{frame}"] = np.fromfile(po, dtype=np.float32).reshape(info["funcs"], n)
        np.savez_compressed(os.path.join(args.out, "ref_cuda_sdf.npz"), **sdf)
        meta["sdf_probe"] = {"n": n, "funcs": int(sdf["frame0"].shape[0]), "frames": [0, 40],
                             "functions": ["sdfSphere(p,0.9)", "sdfBox(p+(0.5,0.2,-0.4),0.1)", "sdfBox(..., r=0.04)", "sdfPlane(p)",
                                           "sdfCylinder(p+(0.5,0.2,-0.4),0.1)", "displacementPattern(p,s)", "sdfOpRound(tanh(s),0.04)",
                                           "sdfOpOnion(s,0.02)", "sdfOpSmoothSubtraction(sdfCylinder(..),tanh(s),0.05)",
                                           "manyCylinderCut(p,s)", "manySphere(p,s,false)", "manySphere(p,s,true)",
                                           "sdfOpSmoothUnion(s,sdfSphere(p,0.5),0.05)", "tanh(s)", "sdfOpSubtraction(s,sdfSphere(p,0.5))",
                                           "sdfOpIntersect(sdfOpUnion(s,sdfPlane(p)),sdfSphere(p,0.9))"]}
        print("sdf probe", meta["sdf_probe"]["funcs"], "functions x", n, "points x 2 frames", flush=True)

    if args.big:
        x = np.random.default_rng(0).uniform(-1, 1, (1000000, 3)).astype(np.float32)
        px, py = os.path.join(tmp, "x.f32"), os.path.join(tmp, "y.f32")
        x.tofile(px)
        meta["mlp_1M"] = run([DRIVER, "mlp", model_path("car_1"), px, py, "3"])
        iv, nm = oracle.view_matrices(0.0, 0.0)
        np.concatenate([iv.ravel(), nm.ravel()]).astype(np.float32).tofile(os.path.join(tmp, "cam.f32"))
        meta["frame_512_car_1"] = run([DRIVER, "render", model_path("car_1"), "512", "512", os.path.join(ROOT, "data", "matcaps", "blue.png"),
                                       os.path.join(tmp, "cam.f32"), "0", os.path.join(tmp, "o.u32"), "1"], timeout=1200)
        print("big", meta["mlp_1M"], meta["frame_512_car_1"], flush=True)
    json.dump(meta, open(os.path.join(args.out, "ref_cuda_meta.json"), "w"), indent=1)


if __name__ == "__main__":
    main()

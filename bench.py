#!/usr/bin/env python
"""bench.py -- headline benchmark of the neural-SDF hot path (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            # product arm (libnsdf.so on B200)
  python bench.py --impl reference --gpus N --steps K ...  # CPU arm: the oracle port on the host cores

A "step" is one 1920x1080 sphere-traced frame of plane_1.h5 with tetrahedron normals and matcap
shading (BASELINE config 2), inputs resident in HBM.  N > 1: one process per GPU (torchrun), every
rank renders its own view of the turntable (rank r: ry = r * 5.625 degrees) -- whole views per GPU,
no data-path collective, weak scaling.  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLOP_PER_QUERY = 14592            # 2 * (3*32 + 7*32*32 + 32*1), SURVEY 8(d)
MODEL = "plane_1"
W, H = 1920, 1080
MATCAP = "blue.png"
VIEW_STEP_DEG = 5.625


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "tflops_burst": d["bf16_tflops"], "tflops_sustained": d["bf16_tflops_sustained"],
                "source": "measured"}
    return {"hbm_gbs": 6650.0, "tflops_burst": 1590.0, "tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = sorted(int(r[0]) for r in self.rows if len(r) >= 6 and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) >= 6 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 6 and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def cpu_oracle_frame(rows=None, nthreads=0, ry=0.0):
    """One (sampled) frame of the workload on the host cores through the CPU oracle."""
    from oracle import oracle
    m = oracle.Model.from_h5(os.path.join(ROOT, "data", "neuralGeometries", MODEL + ".h5"))
    mc = oracle.load_matcap_png(os.path.join(ROOT, "data", "matcaps", MATCAP))
    iv, nm = oracle.view_matrices(0.0, ry)
    t0 = time.perf_counter()
    r = oracle.render(m, iv, nm, W, H, matcap=mc, want_stats=False, rows=rows, nthreads=nthreads)
    return time.perf_counter() - t0, r["queries"], oracle.hardware_threads()


REF_DRIVER = os.path.join(ROOT, "oracle", "_ref", "ref_driver")


def reference_cuda_frame(frames, model=MODEL, w=W, h=H, ry=0.0, timeout=900):
    """The UNMODIFIED reference's own CUDA render_kernel (oracle/_ref/ref_driver = the reference's translation units
    compiled from /root/reference by oracle/build_ref.sh + a headless main) on this box's GPU 0: `frames` timed frames
    after one warm-up call.  Returns the driver's JSON or None when it is not built / there is no GPU."""
    import tempfile
    import numpy as np
    if not os.path.exists(REF_DRIVER):
        return None
    from oracle import oracle
    iv, nm = oracle.view_matrices(0.0, ry)
    with tempfile.TemporaryDirectory() as td:
        cam = os.path.join(td, "cam.f32")
        np.concatenate([np.asarray(iv, np.float32).ravel(), np.asarray(nm, np.float32).ravel()]).tofile(cam)
        cmd = [REF_DRIVER, "render", os.path.join(ROOT, "data", "neuralGeometries", model + ".h5"), str(w), str(h),
               os.path.join(ROOT, "data", "matcaps", MATCAP), cam, "0", os.path.join(td, "out.u32"), str(frames)]
        try:
            r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
        except (OSError, subprocess.TimeoutExpired):
            return None
        if r.returncode != 0:
            return None
        for ln in r.stdout.splitlines():
            if ln.startswith("{"):
                return json.loads(ln)
    return None


def run_reference(args):
    """--impl reference.  The reference (daviesthomas/cudaNeuralRender) has NO CPU implementation of this path: its
    implementation IS CUDA.  When oracle/_ref/ref_driver exists (the reference's own, unmodified CUDA translation units,
    built by oracle/build_ref.sh) and a GPU is present, this arm times THAT on the same box: whole 1080p plane_1 frames
    through the reference's render_kernel (host loop of <= 6000 iterations x {thrust scan, D2H, createBatch, 9 CUTLASS
    GemmBatched, singleMarch}).  Its sceneSDF is what HEAD evaluates (manySphere); the native arm reports the same scene
    as `head_scene`.  Otherwise it times the CPU oracle (a line-by-line port) with all host threads on 1/8 of the rows."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = {"workload": f"{MODEL}.h5 {W}x{H} sphere-traced frame, FD normals + matcap {MATCAP}, max_steps=6000"}
    base = {"impl": "reference", "metric": "1080p frames/sec (plane_1, FD normals + matcap)", "unit": "frames/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "gpu_launches": 0}
    have_gpu = False
    try:
        import torch
        have_gpu = torch.cuda.is_available()
    except Exception:
        pass
    ref = reference_cuda_frame(max(1, args.steps)) if (have_gpu and not args.cpu_port) else None
    if ref is not None:
        fps = 1e3 / ref["mean_ms"]
        cfg["note"] = "reference's own CUDA kernels (unmodified, sm_100 build), scene = manySphere as at HEAD, one process, GPU 0"
        line = dict(base, value=fps, ms_per_step=ref["mean_ms"], config=cfg,
                    cpu_baseline={"value": fps, "unit": "frames/s", "cores": 1, "kind": "reference",
                                  "sample": f"{ref['frames']} whole {W}x{H} frames through the reference's render_kernel on this "
                                            f"box's GPU 0 after one warm-up frame (best {ref['best_ms']:.1f} ms, first {ref['first_call_ms']:.1f} ms)"},
                    e2e={"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0})
        print(json.dumps(line), flush=True)
        return
    stride = 8
    for _ in range(args.warmup):
        cpu_oracle_frame(rows=(0, H, stride))
    ts, qs, cores = [], 0, 0
    for _ in range(args.steps):
        dt, q, cores = cpu_oracle_frame(rows=(0, H, stride))
        ts.append(dt)
        qs += q
    total = sum(ts)
    fps = args.steps / (total * stride)
    sample = f"rows y%{stride}==0 of the {W}x{H} frame per step (1/{stride} of the rays), scaled by {stride}"
    cfg["note"] = "reference has no CPU path and its CUDA build is not available here; oracle port (scene = tanh) timed"
    line = dict(base, value=fps, ms_per_step=1e3 * total / args.steps * stride, config=cfg, queries_per_s=qs / total,
                cpu_baseline={"value": fps, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample},
                e2e={"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0})
    print(json.dumps(line), flush=True)


def run_native(args):
    import ctypes as C
    import numpy as np
    import torch
    import torch.distributed as dist
    from cudaneuralrender_b200 import host
    from cudaneuralrender_b200 import _capi as capi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    L = capi.lib()
    nn = host.NeuralNetwork()
    assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", MODEL + ".h5"))
    matcap = host.Image()
    assert matcap.loadPNG(os.path.join(ROOT, "data", "matcaps", MATCAP))
    host.copyStaticSettings(host.COLOR_MATCAP, 3)
    host.updateViewMatrices((0.0, rank * VIEW_STEP_DEG, 0.0), (0.0, 0.0, -2.0), 0)
    host.setSceneOptions(host.SCENE_TANH, args.max_steps, 0, 0)

    frame = torch.zeros(W * H, dtype=torch.int32, device="cuda")
    stats = torch.zeros(6, dtype=torch.int64, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2
    K, Wm = args.steps, args.warmup

    def step(flags=0):
        host._state.flags = flags
        host.render_kernel(frame, W, H, nn, matcap)

    for _ in range(Wm):
        step()
    # ---- untimed accounting pass: Q_frame, hits, rays handed to the latency warps
    stats.zero_()
    host._state.flags = 0
    host.render_kernel(frame, W, H, nn, matcap, stats=stats)
    torch.cuda.synchronize()
    sAB = stats.cpu().numpy().copy()
    q_frame, hits, evicted = int(sAB[0]), int(sAB[1]), int(sAB[3])

    # ---- timed region: K steps, device time per step, L2 flushed between steps
    sampler = ClockSampler(local)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    barrier()
    if rank == 0:
        sampler.start()
    t_wall0 = time.perf_counter()
    for i in range(K):
        flush.zero_()
        ev[i][0].record()
        step()
        ev[i][1].record()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.stop() if rank == 0 else None
    per_step_ms = [a.elapsed_time(b) for a, b in ev]
    my_ms = sum(per_step_ms)
    total_ms = max_over_ranks(my_ms)
    ms_per_step = total_ms / K
    fps = world * K / (total_ms * 1e-3)
    q_total = sum_over_ranks(float(q_frame))

    # ---- A/B: the earlier two-launch schedule (throughput kernel, then latency kernel), same frame
    evp = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(min(K, 5))]
    for e in evp:
        flush.zero_()
        e[0].record()
        step(capi.RENDER_TWO_KERNELS)
        e[1].record()
    torch.cuda.synchronize()
    host._state.flags = 0
    t_two = sum(e[0].elapsed_time(e[1]) for e in evp) / len(evp)

    # ---- mixed-precision tracer (opt-in NSDF_RENDER_TC_FARFIELD): far field on the tensor cores (fp16x3 rung), fp32 near
    # the surface.  Same frame, timed the same way, parity against the bit-exact frame measured here.
    mixed = None
    if args.mixed:
        exact_img = frame.cpu().numpy().view(np.uint32).copy()
        st_m = torch.zeros(6, dtype=torch.int64, device="cuda")
        host._state.flags = capi.RENDER_TC_FARFIELD
        host.render_kernel(frame, W, H, nn, matcap, stats=st_m)
        for _ in range(2):
            host.render_kernel(frame, W, H, nn, matcap)
        evm = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(min(K, 10))]
        for e in evm:
            flush.zero_()
            e[0].record()
            host.render_kernel(frame, W, H, nn, matcap)
            e[1].record()
        torch.cuda.synchronize()
        host._state.flags = 0
        mixed_img = frame.cpu().numpy().view(np.uint32)
        t_mixed = sum(e[0].elapsed_time(e[1]) for e in evm) / len(evm)
        h0, h1 = exact_img != 0, mixed_img != 0
        both = h0 & h1
        dch = np.max(np.stack([np.abs(((exact_img >> s) & 255).astype(np.int32) - ((mixed_img >> s) & 255).astype(np.int32))
                               for s in (0, 8, 16)]), axis=0)
        sm = st_m.cpu().numpy()
        mixed = {"ms_per_frame": t_mixed, "frames_per_s": 1e3 / t_mixed, "queries": int(sm[0]), "rays_finished_in_fp32": int(sm[3]),
                 "tflops_algorithmic": int(sm[0]) * FLOP_PER_QUERY / (t_mixed * 1e-3) / 1e12,
                 "hit_mask_mismatches": int((h0 != h1).sum()), "hit_pixels": int(h0.sum()),
                 "hit_px_within_1_255": float((dch[both] <= 1).mean()) if both.any() else 1.0,
                 "note": "far field (t >= 1e-4) on tcgen05 fp16x3 (|SDF err| <= 1e-4), accept test / normals / shading fp32; "
                         "not bit-identical, so not the headline"}
        host.render_kernel(frame, W, H, nn, matcap)   # leave the exact frame behind
        torch.cuda.synchronize()

    # ---- the scene the reference evaluates at HEAD (manySphere), for a like-for-like ratio against --impl reference
    host.setSceneOptions(host.SCENE_MANYSPHERE, args.max_steps, 0, 0)
    st_h = torch.zeros(6, dtype=torch.int64, device="cuda")
    host.render_kernel(frame, W, H, nn, matcap, stats=st_h)
    evh = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(min(K, 5))]
    for e in evh:
        flush.zero_()
        e[0].record()
        host.render_kernel(frame, W, H, nn, matcap)
        e[1].record()
    torch.cuda.synchronize()
    t_head = sum(e[0].elapsed_time(e[1]) for e in evh) / len(evh)
    head_scene = {"ms_per_frame": t_head, "frames_per_s": 1e3 / t_head, "queries": int(st_h[0].item()), "hit_pixels": int(st_h[1].item()),
                  "note": "scene = manySphere (src/volumeRender_kernel.cu:221), what the reference's HEAD renders; bit-exact path"}
    host.setSceneOptions(host.SCENE_TANH, args.max_steps, 0, 0)
    host.render_kernel(frame, W, H, nn, matcap)
    torch.cuda.synchronize()

    # ---- end to end through the C ABI with host buffers (pinned): matcap H2D + frame D2H every step
    mc_host = torch.from_numpy(matcap.hostData.view(np.int32)).pin_memory()
    out_host = torch.empty(W * H, dtype=torch.int32).pin_memory()
    st = capi.RenderStats()
    opts = capi.RenderOpts(host.COLOR_MATCAP, host.SCENE_TANH, args.max_steps, 0, 1, 0, 0)

    def e2e_step():
        capi.check(L.nsdf_render_host(nn.handle(), C.byref(host._state.cam), mc_host.data_ptr(), matcap.shape.x,
                                      matcap.shape.y, C.byref(opts), out_host.data_ptr(), W, H, C.byref(st)),
                   "nsdf_render_host")
    for _ in range(max(1, Wm)):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(K):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e_fps = world * K / e2e_s
    h2d = matcap.size() * 4 + 116 + 28
    d2h = W * H * 4 + 48
    assert int(st.queries) == q_frame, "e2e call rendered a different frame"

    # ---- dense sweep (BASELINE config 3): 256^3 lattice through car_1, points generated in-kernel
    dense = None
    if args.dense:
        nd = host.NeuralNetwork()
        assert nd.load(os.path.join(ROOT, "data", "neuralGeometries", "car_1.h5"))
        res = 256
        yb = torch.empty(res ** 3, dtype=torch.float32, device="cuda")
        dense = {}
        for label, prec in (("fp32", host.PREC_FP32),) + tuple(args.dense_tc):
            try:
                for _ in range(2):
                    nd.forward_lattice(res, out=yb, precision=prec)
            except capi.NsdfError:
                continue
            torch.cuda.synchronize()
            es = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(5)]
            for a, b in es:
                flush.zero_()
                a.record()
                nd.forward_lattice(res, out=yb, precision=prec)
                b.record()
            torch.cuda.synchronize()
            ms = min(a.elapsed_time(b) for a, b in es)
            dense[label] = {"queries_per_s": res ** 3 / (ms * 1e-3), "ms": ms,
                            "tflops": res ** 3 * FLOP_PER_QUERY / (ms * 1e-3) / 1e12}

    if world > 1:
        dist.barrier()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    pk = peaks()
    traffic = None
    try:   # dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel, from the committed ncu --set full capture
        tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))["trace_coop_kernel"]
        traffic = tj["dram_bytes_read"] + tj["dram_bytes_write"]
    except Exception:
        pass
    achieved = q_frame * FLOP_PER_QUERY / (ms_per_step * 1e-3) / 1e12
    hbm_bytes = 4 * W * H + 4 * hits + nn.getNumWeightParams() * 4 + nn.getNumBiasParams() * 4
    line = {
        "metric": "1080p frames/sec (plane_1, FD normals + matcap)", "value": fps, "unit": "frames/s", "n_gpus": world,
        "steps": K, "warmup": Wm, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{MODEL}.h5 {W}x{H} sphere-traced frame, FD normals + matcap {MATCAP}, scene=tanh, "
                               f"max_steps={args.max_steps}",
                   "parallelism": f"views{world}" if world > 1 else "single", "views_per_step": world,
                   "l2": "256 MiB memset between timed steps (outside the per-step events)",
                   "precision": "fp32 FFMA rung (bit-exact vs oracle)"},
        "queries_per_s": q_total * K / (total_ms * 1e-3), "queries_per_frame": q_frame, "hit_pixels": hits,
        "evicted_rays": evicted,
        "e2e": {"value": e2e_fps, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": 3 * K,   # per step: tile_probe_kernel, tile_sort_kernel, trace_coop_kernel (+ one 16-byte memset)
        "kernels": {"trace_coop_kernel (+ tile_probe_kernel, tile_sort_kernel)": {"ms": ms_per_step, "queries": q_frame},
                    "two_launch_schedule_ms": t_two},
        "roofline": {"kernel": "trace_coop_kernel", "bound": "tensor", "achieved": achieved,
                     "peak": pk["tflops_burst"], "unit": "TFLOP/s", "frac": achieved / pk["tflops_burst"],
                     "traffic": traffic, "peak_source": pk["source"],
                     "note": "algorithmic FLOP = queries x 14592 against the measured bf16 tensor peak; this is the bit-exact fp32 FFMA "
                             "rung, whose own ceiling is 74.4 TFLOP/s (frac_of_fp32_alu below); traffic = DRAM bytes per launch from "
                             "profiles/ncu_traffic.json (the frame stays in L2 within the launch)",
                     "frac_of_fp32_alu": achieved / 74.4,
                     "hbm": {"algorithmic_bytes": hbm_bytes, "achieved_gbs": hbm_bytes / (ms_per_step * 1e-3) / 1e9,
                             "peak_gbs": pk["hbm_gbs"]}},
        "clocks": clocks, "wall_s": t_wall,
    }
    line["head_scene"] = head_scene
    if mixed:
        line["mixed_precision"] = mixed
    if world == 1 and args.ref_cuda:
        rc = reference_cuda_frame(2)
        if rc is not None:
            line["reference_cuda_same_box"] = {"ms_per_frame": rc["mean_ms"], "frames_per_s": 1e3 / rc["mean_ms"],
                                               "speedup_vs_head_scene": rc["mean_ms"] / t_head,
                                               "note": "the reference's own unmodified CUDA kernels (oracle/_ref/ref_driver), same frame, "
                                                       "same GPU, 2 frames after a warm-up"}
    if dense:
        line["dense_256cubed"] = dense
    if world == 1 and not args.no_cpu:
        dt, q, cores = cpu_oracle_frame()
        line["cpu_baseline"] = {"value": 1.0 / dt, "unit": "frames/s", "cores": cores, "kind": "port",
                                "sample": f"one full {W}x{H} frame ({q} queries) through oracle/nsdf_oracle.cpp, all host threads",
                                "queries_per_s": q / dt}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_tiles(args):
    """BASELINE config 4: one 3840x2160 frame of 3a3d4a90....obj.h5, 64x64 tiles round-robin over the ranks,
    every rank's tracer storing its tiles straight into rank 0's frame buffer over NVLink (peer-mapped d_out).
    Strong scaling: value = frames/s of the whole job."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from cudaneuralrender_b200 import host, multigpu

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(x, op):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=op)
        return float(t.item())

    model, Wt, Ht = "3a3d4a90a2db90b4203936772104a82d.obj", 3840, 2160
    nn = host.NeuralNetwork()
    assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", model + ".h5"))
    matcap = host.Image()
    assert matcap.loadPNG(os.path.join(ROOT, "data", "matcaps", MATCAP))
    host.copyStaticSettings(host.COLOR_MATCAP, 3)
    host.updateViewMatrices((0.0, 0.0, 0.0), (0.0, 0.0, -2.0), 0)
    from cudaneuralrender_b200 import _capi as capi
    mixed = os.environ.get("NSDF_BENCH_TILES_MIXED", "0") == "1"   # opt-in: the mixed-precision tracer on every rank
    host.setSceneOptions(host.SCENE_TANH, args.max_steps, capi.RENDER_TC_FARFIELD if mixed else 0, 0)
    frame = multigpu.PeerFrame(Wt, Ht, rank, world, root=0)
    stats = torch.zeros(6, dtype=torch.int64, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    K, Wm = args.steps, args.warmup

    for _ in range(Wm):
        multigpu.render_frame_tiled(nn, matcap, Wt, Ht, frame)
    stats.zero_()
    multigpu.render_frame_tiled(nn, matcap, Wt, Ht, frame, stats=stats)
    barrier()
    my_q = float(stats[0].item())
    q_frame = reduce(my_q, dist.ReduceOp.SUM if world > 1 else None)
    q_max = reduce(my_q, dist.ReduceOp.MAX if world > 1 else None)
    hits = reduce(float(stats[1].item()), dist.ReduceOp.SUM if world > 1 else None)

    # parity of the partition: the gathered frame must equal a single-GPU render of the same frame, bytewise
    equal = None
    if rank == 0:
        gathered = frame.to_numpy()
        single = torch.zeros(Wt * Ht, dtype=torch.int32, device="cuda")
        host.render_kernel(single, Wt, Ht, nn, matcap)
        torch.cuda.synchronize()
        equal = bool(np.array_equal(gathered.ravel(), single.cpu().numpy().view(np.uint32)))
        if mixed:   # the mixed mode is not schedule-independent bit for bit: report the hit-mask agreement instead
            equal = bool(np.array_equal(gathered.ravel() != 0, single.cpu().numpy().view(np.uint32) != 0))

    sampler = ClockSampler(local)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    barrier()
    if rank == 0:
        sampler.start()
    t0 = time.perf_counter()
    for i in range(K):
        flush.zero_()
        ev[i][0].record()
        multigpu.render_frame_tiled(nn, matcap, Wt, Ht, frame)
        ev[i][1].record()
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None
    my_ms = sum(a.elapsed_time(b) for a, b in ev)
    total_ms = reduce(my_ms, dist.ReduceOp.MAX if world > 1 else None)   # a frame is done when its slowest rank is
    min_ms = reduce(my_ms, dist.ReduceOp.MIN if world > 1 else None)

    # end to end: frame on the HOST of rank 0 every step (render -> barrier -> D2H on the root)
    barrier()
    t0 = time.perf_counter()
    for _ in range(K):
        multigpu.render_frame_tiled(nn, matcap, Wt, Ht, frame)
        barrier()
        if rank == 0:
            frame.to_numpy()
    barrier()
    e2e_s = reduce(time.perf_counter() - t0, dist.ReduceOp.MAX if world > 1 else None)

    if rank == 0:
        pk = peaks()
        ms = total_ms / K
        achieved = q_frame * FLOP_PER_QUERY / (ms * 1e-3) / 1e12
        line = {"metric": "4K frames/sec (3a3d4a90.obj, FD normals + matcap), tile-partitioned", "value": K / (total_ms * 1e-3),
                "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": Wm, "ms_per_step": ms, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": f"{model}.h5 {Wt}x{Ht} sphere-traced frame, FD normals + matcap {MATCAP}, scene=tanh, "
                                       f"max_steps={args.max_steps}",
                           "parallelism": f"tiles{world}: 64x64 tiles round-robin, peer stores into GPU 0 over NVLink",
                           "l2": "256 MiB memset between timed steps (outside the per-step events)",
                           "precision": "mixed: tcgen05 fp16x3 far field + fp32 near surface (hit masks compared)" if mixed
                                        else "fp32 FFMA rung (bit-exact vs oracle)"},
                "queries_per_frame": q_frame, "queries_per_s": q_frame * K / (total_ms * 1e-3), "hit_pixels": hits,
                "tile_balance_queries": q_frame / world / q_max if q_max else None,
                "rank_time_min_over_max": min_ms / total_ms if total_ms else None,
                "tiles_equal_single_gpu": equal,
                "e2e": {"value": K / e2e_s, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": Wt * Ht * 4},
                "gpu_launches": K * world,
                "roofline": {"kernel": "trace_coop_kernel", "bound": "tensor", "achieved": achieved,
                             "peak": pk["tflops_burst"] * world, "unit": "TFLOP/s",
                             "frac": achieved / (pk["tflops_burst"] * world), "traffic": None, "peak_source": pk["source"],
                             "note": "whole job; peak = N x one GPU's measured bf16 peak; fp32 ALU rung (own ceiling 74 TFLOP/s per GPU)"},
                "clocks": clocks, "wall_s": wall}
        print(json.dumps(line), flush=True)
    frame.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--max-steps", type=int, default=6000, dest="max_steps")
    ap.add_argument("--no-dense", action="store_false", dest="dense")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-mixed", action="store_false", dest="mixed")
    ap.add_argument("--cpu-port", action="store_true", help="--impl reference: time the CPU oracle port even if the reference's "
                                                            "CUDA build (oracle/_ref) is present")
    ap.add_argument("--no-ref-cuda", action="store_false", dest="ref_cuda")
    ap.add_argument("--workload", default="frame1080", choices=["frame1080", "frame4k_tiles"],
                    help="frame1080: BASELINE config 2 (default; N>1 = one view per GPU, weak scaling).  "
                         "frame4k_tiles: BASELINE config 4, ONE 3840x2160 frame tile-partitioned over the N GPUs, "
                         "gathered on GPU 0 over NVLink (strong scaling)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "native" else args.warmup
    args.dense_tc = (("tc_fp16", 1), ("tc_bf16", 2), ("tc_fp16x3", 3), ("tc_fp16_acc16", 5))
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "frame4k_tiles":
        run_tiles(args)
    else:
        run_native(args)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""bench.py -- headline benchmark of the neural-SDF hot path (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            # product arm (libnsdf.so on B200)
  python bench.py --impl reference --gpus N --steps K ...  # reference arm: the reference's OWN CUDA (oracle/_ref/ref_driver)

A "step" is one pass over a FIXED view set: the 64-view turntable (ry = k * 5.625 degrees) of plane_1.h5
at 1920x1080 with tetrahedron normals and matcap shading (BASELINE config 2), rendered with what the unmodified reference
renders at HEAD (sceneSDF = manySphere, reference math) -- the same frames, the same config string in both arms.
A rank's views of a step go through ONE call of nsdf_render_views (one persistent launch: the serial tail of one view hides
behind the bulk of the others).  N > 1: one process per GPU (torchrun); view v goes to rank v % N, so every N renders the
SAME 64 frames per step (strong scaling, no data-path collective).  `value` = frames/s of the whole job.

Every line also carries the other BASELINE configs as legs: `frame4k_tiles` (C4: one 3840x2160 frame, 64x64 tiles round-robin
over the N GPUs, peer stores into GPU 0 over NVLink), `turntable_c5` (C5: 64 views x 5 shapes), and at N = 1 `c1` (car_1
512x512 frame, 1 000 000-point batch from an HBM B x 3 buffer), `dense_256cubed` (C3), `tanh_scene`, `mixed_precision`,
`parity_vs_oracle`, `parity_vs_reference`, `cpu_baseline`.  Prints ONE JSON line on rank 0.
"""
import argparse
import hashlib
import json
import os
import re
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLOP_PER_QUERY = 14592            # 2 * (3*32 + 7*32*32 + 32*1), SURVEY 8(d)
FP32_ALU_TFLOPS = 74.4            # 148 SMs x 128 FMA lanes x 2 x 1.965 GHz: the ceiling of the bit-exact fp32 rung
MODEL = "plane_1"
W, H = 1920, 1080
MATCAP = "blue.png"
N_VIEWS = 64
VIEW_STEP_DEG = 5.625
SHAPES = ["car_1", "plane_1", "plane_2", "plane_3", "3a3d4a90a2db90b4203936772104a82d.obj"]
WORKLOAD = (f"{MODEL}.h5 {W}x{H} sphere-traced frames, {N_VIEWS}-view turntable ry=k*{VIEW_STEP_DEG}deg per step, FD normals + "
            f"matcap {MATCAP}, scene=manySphere (reference HEAD), max_steps=6000")
METRIC = "1080p frames/sec (plane_1, FD normals + matcap)"
REF_DRIVER = os.path.join(ROOT, "oracle", "_ref", "ref_driver")


def geom(name):
    return os.path.join(ROOT, "data", "neuralGeometries", name + ".h5")


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "tflops_burst": d["bf16_tflops"], "tflops_sustained": d["bf16_tflops_sustained"],
                "source": "measured"}
    return {"hbm_gbs": 6650.0, "tflops_burst": 1590.0, "tflops_sustained": 1400.0, "source": "fallback"}


def kernel_sass_sha(kernel):
    """sha of the SASS of the library's `kernel` instantiation that the bench launches (instruction text without addresses'
    encodings): profiles/ncu_traffic.json is only quoted when it was captured from the same machine code.  None without cuobjdump."""
    want = {"trace_bulk4_kernel": "trace_bulk4_kernelILi2E", "trace_coop_kernel": "trace_coop_kernelE"}.get(kernel, kernel)
    lib = os.path.join(ROOT, "cudaneuralrender_b200", "libnsdf.so")
    try:
        out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, timeout=120).stdout
    except Exception:
        return None
    h, inside, n = hashlib.sha256(), False, 0
    for line in out.splitlines():
        if "Function :" in line:
            inside = want in line
            continue
        if inside:
            m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(.*?)\s*/\* 0x[0-9a-f]+ \*/", line)
            if m:
                h.update(m.group(1).encode() + b"\n")
                n += 1
    return h.hexdigest()[:16] if n else None


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


# ---- checker / baseline legs (the only places that touch oracle/) -------------------------------------------------------
def cpu_oracle_frame(rows=None, nthreads=0, ry=0.0, scene=None, model=MODEL, w=W, h=H):
    """One (sampled) frame on the host cores through the CPU oracle (deterministic arithmetic)."""
    from oracle import oracle
    m = oracle.Model.from_h5(geom(model))
    mc = oracle.load_matcap_png(os.path.join(ROOT, "data", "matcaps", MATCAP))
    iv, nm = oracle.view_matrices(0.0, ry)
    t0 = time.perf_counter()
    r = oracle.render(m, iv, nm, w, h, matcap=mc, want_stats=False, rows=rows, nthreads=nthreads,
                      scene=oracle.SCENE_MANYSPHERE if scene is None else scene)
    return time.perf_counter() - t0, r, oracle.hardware_threads()


def reference_cuda(args_list, timeout=900):
    """oracle/_ref/ref_driver = the UNMODIFIED reference's CUDA translation units (compiled from /root/reference by
    oracle/build_ref.sh) + a headless main.  Returns its JSON line, or None when it is not built / fails / there is no GPU."""
    if not os.path.exists(REF_DRIVER):
        return None
    try:
        r = subprocess.run([REF_DRIVER] + args_list, capture_output=True, text=True, timeout=timeout)
    except (OSError, subprocess.TimeoutExpired):
        return None
    if r.returncode != 0:
        return None
    for ln in r.stdout.splitlines():
        if ln.startswith("{"):
            return json.loads(ln)
    return None


def reference_cuda_frame(frames, model=MODEL, w=W, h=H, ry=0.0, td=None):
    import tempfile
    import numpy as np
    from oracle import oracle
    iv, nm = oracle.view_matrices(0.0, ry)
    with tempfile.TemporaryDirectory() as tmp:
        cam = os.path.join(tmp, "cam.f32")
        np.concatenate([np.asarray(iv, np.float32).ravel(), np.asarray(nm, np.float32).ravel()]).tofile(cam)
        return reference_cuda(["render", geom(model), str(w), str(h), os.path.join(ROOT, "data", "matcaps", MATCAP), cam, "0",
                               os.path.join(tmp, "out.u32"), str(frames)])


def run_reference(args):
    """--impl reference.  The reference (daviesthomas/cudaNeuralRender) has NO CPU implementation of this path: its
    implementation IS CUDA.  This arm times THAT on the same box: whole 1080p plane_1 frames of the same view set through the
    reference's own render_kernel (host loop of <= 6000 iterations x {thrust scan, D2H, createBatch, 9 CUTLASS GemmBatched,
    singleMarch}).  One view per step, cycling through the set in order (a frame takes > 1 s).  Without the built driver or a
    GPU it times the CPU oracle port (all host threads, 1/8 of the rows of each view)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = {"workload": WORKLOAD}
    base = {"impl": "reference", "metric": METRIC, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "gpu_launches": 0}
    have_gpu = False
    try:
        import torch
        have_gpu = torch.cuda.is_available()
    except Exception:
        pass
    if have_gpu and not args.cpu_port and os.path.exists(REF_DRIVER):
        times, first = [], None
        for _ in range(max(0, args.warmup - 1)):     # the driver's own first call inside each process is a warm-up frame already
            reference_cuda_frame(1, ry=0.0)
        for k in range(max(1, args.steps)):
            r = reference_cuda_frame(1, ry=(k % N_VIEWS) * VIEW_STEP_DEG)
            if r is None:
                times = []
                break
            times.append(r["mean_ms"])
            first = first or r["first_call_ms"]
        if times:
            ms = sum(times) / len(times)
            fps = 1e3 / ms
            cfg["note"] = ("reference's own CUDA kernels (unmodified, sm_100 build) on this box's GPU 0, one process; one view of the "
                           "set per step (views 0.." + str(len(times) - 1) + "), each timed after a warm-up frame of the same view")
            line = dict(base, value=fps, ms_per_step=ms, config=cfg,
                        cpu_baseline={"value": fps, "unit": "frames/s", "cores": 1, "kind": "reference",
                                      "sample": f"{len(times)} whole {W}x{H} frames (views 0..{len(times) - 1} of the set) through the "
                                                f"reference's render_kernel on GPU 0 (min {min(times):.1f} ms, max {max(times):.1f} ms)"},
                        e2e={"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0})
            print(json.dumps(line), flush=True)
            return
    stride = 8
    for _ in range(args.warmup):
        cpu_oracle_frame(rows=(0, H, stride))
    ts, qs, cores = [], 0, 0
    for k in range(args.steps):
        dt, r, cores = cpu_oracle_frame(rows=(0, H, stride), ry=(k % N_VIEWS) * VIEW_STEP_DEG)
        ts.append(dt)
        qs += r["queries"]
    total = sum(ts)
    fps = args.steps / (total * stride)
    cfg["note"] = "the reference's CUDA build is not available here (no GPU or oracle/_ref missing): CPU oracle port timed"
    line = dict(base, value=fps, ms_per_step=1e3 * total / args.steps * stride, config=cfg, queries_per_s=qs / total,
                cpu_baseline={"value": fps, "unit": "frames/s", "cores": cores, "kind": "port",
                              "sample": f"rows y%{stride}==0 of one view per step (1/{stride} of the rays), scaled by {stride}"},
                e2e={"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0})
    print(json.dumps(line), flush=True)


def compare_frames(mine, ref):
    import numpy as np
    hm, hr = mine != 0, ref != 0
    mism = hm != hr
    mism[-1, -1] = False           # the reference reads a stale SDF slot for the last pixel
    both = hm & hr
    both[-1, -1] = False
    ch = lambda v: np.stack([(v >> s) & 255 for s in (0, 8, 16)], -1).astype(np.int32)  # noqa: E731
    diff = np.abs(ch(mine) - ch(ref))[both]
    neq = mine != ref
    neq[-1, -1] = False
    return {"mask_mismatches": int(mism.sum()), "differing_pixels": int(neq.sum()), "hit_pixels": int(both.sum()),
            "frac_within_1_255": float((diff.max(1) <= 1).mean()) if diff.size else 1.0}


# ---- product arm ------------------------------------------------------------------------------------------------------
class Dist:
    def __init__(self):
        import torch
        self.torch = torch
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
        torch.cuda.set_device(self.local)
        if self.world > 1:
            import torch.distributed as dist
            self.dist = dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, x, op="max"):
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op={"max": self.dist.ReduceOp.MAX, "min": self.dist.ReduceOp.MIN, "sum": self.dist.ReduceOp.SUM}[op])
        return float(t.item())

    def close(self):
        if self.world > 1:
            self.dist.barrier()
            self.dist.destroy_process_group()


def timed(torch, fn, reps, flush=None):
    """Device time of fn() in ms, per repetition (events on the current stream, L2 flushed outside the events)."""
    out = []
    for _ in range(reps):
        if flush is not None:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        out.append((a, b))
    torch.cuda.synchronize()
    return [a.elapsed_time(b) for a, b in out]


def tiles_leg(D, args, host, multigpu, flush, steps):
    """BASELINE config 4: one 3840x2160 frame of 3a3d4a90....obj.h5 (reference HEAD scene), 64x64 tiles round-robin over the
    ranks, every rank's tracer storing its tiles straight into rank 0's frame buffer over NVLink (peer-mapped d_out)."""
    import numpy as np
    torch = D.torch
    model, Wt, Ht = SHAPES[4], 3840, 2160
    nn = host.NeuralNetwork()
    assert nn.load(geom(model))
    matcap = host.Image()
    assert matcap.loadPNG(os.path.join(ROOT, "data", "matcaps", MATCAP))
    host.copyStaticSettings(host.COLOR_MATCAP, 3)
    host.updateViewMatrices((0.0, 0.0, 0.0), (0.0, 0.0, -2.0), 0)
    host.setSceneOptions(host.SCENE_MANYSPHERE, args.max_steps, 0, 0, host.MATH_REFERENCE)
    frame = multigpu.PeerFrame(Wt, Ht, D.rank, D.world, root=0)
    stats = torch.zeros(6, dtype=torch.int64, device="cuda")
    for _ in range(2):
        multigpu.render_frame_tiled(nn, matcap, Wt, Ht, frame)
    multigpu.render_frame_tiled(nn, matcap, Wt, Ht, frame, stats=stats)
    D.barrier()
    my_q = float(stats[0].item())
    q_frame, q_max = D.reduce(my_q, "sum"), D.reduce(my_q, "max")
    equal = None
    if D.rank == 0:
        gathered = frame.to_numpy()
        single = torch.zeros(Wt * Ht, dtype=torch.int32, device="cuda")
        host.render_kernel(single, Wt, Ht, nn, matcap)
        torch.cuda.synchronize()
        equal = bool(np.array_equal(gathered.ravel(), single.cpu().numpy().view(np.uint32)))
        del single
    D.barrier()
    ms = timed(torch, lambda: multigpu.render_frame_tiled(nn, matcap, Wt, Ht, frame), steps, flush)
    D.barrier()
    mine = sum(ms)
    total, tmin = D.reduce(mine, "max"), D.reduce(mine, "min")     # a frame is done when its slowest rank is
    D.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):                                          # end to end: the frame on rank 0's HOST every step
        multigpu.render_frame_tiled(nn, matcap, Wt, Ht, frame)
        D.barrier()
        if D.rank == 0:
            frame.to_numpy()
    D.barrier()
    e2e_s = D.reduce(time.perf_counter() - t0, "max")
    frame.close()
    return {"workload": f"{model}.h5 {Wt}x{Ht} frame, scene=manySphere (reference HEAD), 64x64 tiles round-robin over {D.world} GPU(s), "
                        "peer stores into GPU 0 over NVLink", "scaling": "strong",
            "ms_per_frame": total / steps, "frames_per_s": steps / (total * 1e-3), "e2e_frames_per_s": steps / e2e_s,
            "queries_per_frame": q_frame, "tile_balance_queries": (q_frame / D.world / q_max) if q_max else None,
            "rank_time_min_over_max": (tmin / total) if total else None, "tiles_equal_single_gpu": equal, "steps": steps}


def turntable_c5_leg(D, args, host, flush, size):
    """BASELINE config 5: 64 views x the five shipped shapes, view v of shape s -> rank (s*64 + v) % N; one pass."""
    torch = D.torch
    matcap = host.Image()
    assert matcap.loadPNG(os.path.join(ROOT, "data", "matcaps", MATCAP))
    host.copyStaticSettings(host.COLOR_MATCAP, 3)
    host.setSceneOptions(host.SCENE_MANYSPHERE, args.max_steps, 0, 0, host.MATH_REFERENCE)
    nets = []
    for s in SHAPES:
        nn = host.NeuralNetwork()
        assert nn.load(geom(s))
        nets.append(nn)
    Wc, Hc = size
    jobs = [(s, v) for s in range(len(SHAPES)) for v in range(N_VIEWS)]
    mine = [j for i, j in enumerate(jobs) if i % D.world == D.rank]
    per_shape = [[v for s2, v in mine if s2 == s] for s in range(len(SHAPES))]
    cams = [host.view_cameras([(0.0, v * VIEW_STEP_DEG) for v in vs]) if vs else None for vs in per_shape]
    # The five shapes' calls run SIDE BY SIDE: one stream and one fifth of the SMs each (nsdf_tuning.max_ctas), so that one shape's
    # serial tail hides behind the others' bulk (tools/concurrent_probe.py: 8 views per shape 442 -> 600 frames/s, 64 views 673 -> 772).
    frames_c5 = [torch.zeros(max(len(vs), 1) * Wc * Hc, dtype=torch.int32, device="cuda") for vs in per_shape]
    streams = [torch.cuda.Stream() for _ in SHAPES]
    share = max(1, torch.cuda.get_device_properties(torch.cuda.current_device()).multi_processor_count // len(SHAPES))

    def one_pass():
        host.setTuning(max_ctas=share)
        start = torch.cuda.Event()
        start.record()
        for s in range(len(SHAPES)):                      # this rank's views of a shape: one persistent launch on the shape's stream
            if cams[s] is not None:
                streams[s].wait_event(start)
                with torch.cuda.stream(streams[s]):
                    host.render_views(frames_c5[s], Wc, Hc, nets[s], cams[s], matcap)
        for st in streams:
            torch.cuda.current_stream().wait_stream(st)
        host.setTuning()
    one_pass()
    D.barrier()
    ms = timed(torch, one_pass, 1, flush)[0]
    D.barrier()
    total = D.reduce(ms, "max")
    return {"workload": f"64 views (ry=k*{VIEW_STEP_DEG}deg) x 5 shapes at {Wc}x{Hc}, scene=manySphere (reference HEAD), view-parallel over "
                        f"{D.world} GPU(s)", "frames": len(jobs), "ms_total": total, "frames_per_s": len(jobs) / (total * 1e-3),
            "launch": f"per rank: five nsdf_render_views calls (one per shape) side by side on five streams, max_ctas = {share} each"}


def run_native(args):
    import ctypes as C
    import numpy as np
    from cudaneuralrender_b200 import host, multigpu
    from cudaneuralrender_b200 import _capi as capi

    D = Dist()
    torch = D.torch
    world, rank = D.world, D.rank
    L = capi.lib()
    nn = host.NeuralNetwork()
    assert nn.load(geom(MODEL))
    matcap = host.Image()
    assert matcap.loadPNG(os.path.join(ROOT, "data", "matcaps", MATCAP))
    host.copyStaticSettings(host.COLOR_MATCAP, 3)
    head = dict(scene=host.SCENE_MANYSPHERE, max_steps=args.max_steps, flags=0, evict_steps=0, math=host.MATH_REFERENCE)
    host.setSceneOptions(**head)

    stats = torch.zeros(6, dtype=torch.int64, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2
    K, Wm = args.steps, args.warmup
    my_views = multigpu.views_of_rank(N_VIEWS, rank, world)
    cams = host.view_cameras([(0.0, v * VIEW_STEP_DEG) for v in my_views])
    frames = torch.zeros(len(my_views) * W * H, dtype=torch.int32, device="cuda")    # this rank's frames of a step

    def set_view(v):
        host.updateViewMatrices((0.0, v * VIEW_STEP_DEG, 0.0), (0.0, 0.0, -2.0), 0)

    def step(st=None):
        host.render_views(frames, W, H, nn, cams, matcap, stats=st)

    for _ in range(Wm):
        step()
    # ---- untimed accounting pass: queries, hits of this rank's views
    stats.zero_()
    step(stats)
    torch.cuda.synchronize()
    sA = stats.cpu().numpy().copy()
    q_set = D.reduce(float(sA[0]), "sum")
    hits_set = D.reduce(float(sA[1]), "sum")

    # ---- timed region: K steps, device time per step, L2 flushed between steps
    sampler = ClockSampler(D.local)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    D.barrier()
    if rank == 0:
        sampler.start()
    t_wall0 = time.perf_counter()
    for i in range(K):
        flush.zero_()
        ev[i][0].record()
        step()
        ev[i][1].record()
    D.barrier()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.stop() if rank == 0 else None
    my_ms = sum(a.elapsed_time(b) for a, b in ev)
    total_ms = D.reduce(my_ms, "max")
    min_ms = D.reduce(my_ms, "min")
    ms_per_step = total_ms / K
    fps = N_VIEWS * K / (total_ms * 1e-3)

    # ---- end to end through the C ABI with HOST buffers (pinned): matcap H2D + frame D2H every frame
    mc_host = torch.from_numpy(matcap.hostData.view(np.int32)).pin_memory()
    out_host = torch.empty(len(my_views) * W * H, dtype=torch.int32).pin_memory()
    st = capi.RenderStats()
    opts = host._opts()

    def e2e_step():
        capi.check(L.nsdf_render_views_host(nn.handle(), cams, len(cams), mc_host.data_ptr(), matcap.shape.x, matcap.shape.y,
                                            C.byref(opts), out_host.data_ptr(), W, H, C.byref(st)), "nsdf_render_views_host")
    Ke = min(K, args.e2e_steps) if args.e2e_steps > 0 else K
    e2e_step()
    D.barrier()
    t0 = time.perf_counter()
    for _ in range(Ke):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = D.reduce(time.perf_counter() - t0, "max")
    e2e_fps = N_VIEWS * Ke / e2e_s
    h2d = matcap.size() * 4 * world + 128 * N_VIEWS       # the matcap once per rank and step, one camera per view
    d2h = W * H * 4 * N_VIEWS + 48 * world
    img_batch_view0 = frames[:W * H].cpu().numpy().view(np.uint32).reshape(H, W).copy() if rank == 0 else None

    # ---- legs every N carries: C4 (one 4K frame, tile-partitioned), C5 (64 views x 5 shapes) and the 4K turntable (4K frames by view)
    tiles = tiles_leg(D, args, host, multigpu, flush, max(3, min(K, 5))) if args.tiles else None
    c5 = turntable_c5_leg(D, args, host, flush, (512, 512)) if args.c5 else None
    tt4k = turntable4k_leg(D, args, host, multigpu, flush) if args.tt4k else None

    if rank != 0:
        D.close()
        return

    # ================= rank 0 only from here: single-GPU legs =================
    host.copyStaticSettings(host.COLOR_MATCAP, 3)
    host.setSceneOptions(**head)
    set_view(0)
    pk = peaks()
    line_extra = {}

    def frame_leg(net, w, h, reps=5, **opts_kw):
        buf = torch.zeros(w * h, dtype=torch.int32, device="cuda")
        s = torch.zeros(6, dtype=torch.int64, device="cuda")
        host.setSceneOptions(**dict(head, **opts_kw))
        host.render_kernel(buf, w, h, net, matcap, stats=s)
        host.render_kernel(buf, w, h, net, matcap)
        ms = timed(torch, lambda: host.render_kernel(buf, w, h, net, matcap), reps, flush)
        img = buf.cpu().numpy().view(np.uint32).reshape(h, w).copy()
        host.setSceneOptions(**head)
        sv = s.cpu().numpy()
        m = sum(ms) / len(ms)
        return {"ms_per_frame": m, "ms_min": min(ms), "frames_per_s": 1e3 / m, "queries": int(sv[0]), "hit_pixels": int(sv[1]),
                "queries_per_s": int(sv[0]) / (m * 1e-3)}, img

    # view 0 alone (the frame of BASELINE config 2 as round 1 quoted it), both scenes, both math modes
    single, img_ref_math = frame_leg(nn, W, H)
    single["note"] = "view 0 (rx=ry=0) alone, scene=manySphere, reference math"
    tanh_scene, _ = frame_leg(nn, W, H, scene=host.SCENE_TANH, math=host.MATH_DETERMINISTIC)
    tanh_scene["note"] = "view 0, scene=tanh(nSDF) (the shipped renders' variant; round 1's headline), deterministic math"
    line_extra["single_view"] = single
    line_extra["tanh_scene"] = tanh_scene

    # parity of the frame this bench times: against the reference binary's frame (golden, recorded on a B200 by
    # tools/make_ref_golden.py) and, in the deterministic arithmetic, against the CPU oracle's frame rendered below
    try:
        gmeta = json.load(open(os.path.join(ROOT, "tests", "golden", "ref_cuda_meta.json")))["frames"]
        gf = np.load(os.path.join(ROOT, "tests", "golden", "ref_cuda_frames.npz"))
        key = [k for k, m in gmeta.items() if m["model"] == MODEL and m["W"] == W and m["H"] == H and m["ry"] == 0.0 and m["frame"] == 0]
        if key:
            line_extra["parity_vs_reference"] = dict(compare_frames(img_batch_view0, gf[key[0] + "/image"]),
                                                     frame="view 0 as the timed batched step left it in HBM",
                                                     single_view_call_equal=bool(np.array_equal(img_batch_view0, img_ref_math)),
                                                     golden=f"tests/golden/ref_cuda_frames.npz:{key[0]} (reference's own render_kernel on a B200)")
    except Exception as e:   # noqa: BLE001
        line_extra["parity_vs_reference"] = {"unavailable": str(e)[:120]}

    # ---- mixed-precision tracer (opt-in NSDF_RENDER_TC_FARFIELD; tanh scene only): reported, not the headline
    if args.mixed:
        try:
            exact, img_e = frame_leg(nn, W, H, scene=host.SCENE_TANH, math=host.MATH_DETERMINISTIC)
            mixed, img_m = frame_leg(nn, W, H, scene=host.SCENE_TANH, math=host.MATH_DETERMINISTIC, flags=capi.RENDER_TC_FARFIELD)
            cmpm = compare_frames(img_m, img_e)
            line_extra["mixed_precision"] = dict(mixed, exact_ms_per_frame=exact["ms_per_frame"], **cmpm,
                                                 note="tanh scene; far field (t >= 1e-4) on tcgen05 fp16x3, accept test / normals / shading "
                                                      "fp32; not bit-identical and outside the RGB tolerance on part of the hit pixels")
        except capi.NsdfError as e:
            line_extra["mixed_precision"] = {"unavailable": str(e)}

    # ---- C1: car_1 512x512 frame, and the simpleInfer shape: 1 000 000 points from an HBM B x 3 buffer (variant A)
    if args.c1:
        car = host.NeuralNetwork()
        assert car.load(geom("car_1"))
        c1, _ = frame_leg(car, 512, 512)
        c1 = {"frame_512": dict(c1, note="car_1.h5 512x512, scene=manySphere (reference HEAD), matcap blue.png")}
        x = torch.from_numpy(np.random.default_rng(0).uniform(-1, 1, (1000000, 3)).astype(np.float32)).cuda().contiguous()
        y = torch.empty(1000000, dtype=torch.float32, device="cuda")
        for label, prec in (("fp32", host.PREC_FP32), ("tc_fp16x3", host.PREC_TC_FP16X3), ("tc_fp16", host.PREC_TC_FP16)):
            try:
                for _ in range(2):
                    car.forward_tensor(x, prec, out=y)
                ms = min(timed(torch, lambda: car.forward_tensor(x, prec, out=y), 5, flush))
                c1[f"batch_1M_{label}"] = {"ms": ms, "queries_per_s": 1e6 / (ms * 1e-3), "tflops": 1e6 * FLOP_PER_QUERY / (ms * 1e-3) / 1e12,
                                           "hbm_bytes": 16 * 1000000}
            except capi.NsdfError:
                pass
        xh = x.cpu().numpy()
        yh = car.forward_host(xh[:4096])
        t0 = time.perf_counter()
        yh = car.forward_host(xh)
        c1["batch_1M_fp32_host_buffers"] = {"ms": 1e3 * (time.perf_counter() - t0), "note": "nsdf_mlp_forward_host: H2D 12 MB + kernel + D2H 4 MB"}
        if args.ref_cuda and os.path.exists(REF_DRIVER):
            import tempfile
            with tempfile.TemporaryDirectory() as tmp:
                px, py = os.path.join(tmp, "x.f32"), os.path.join(tmp, "y.f32")
                xh.tofile(px)
                rc = reference_cuda(["mlp", geom("car_1"), px, py, "3"])
                if rc is not None:
                    yr = np.fromfile(py, dtype=np.float32)
                    c1["batch_1M_reference_cuda"] = {"ms": rc["best_ms"], "queries_per_s": rc["queries_per_s"], "first_call_ms": rc["first_call_ms"],
                                                     "bit_equal_to_product_fp32": bool(np.array_equal(yr.view(np.uint32), yh.view(np.uint32))),
                                                     "note": "the reference's nn.forward (9 CUTLASS GemmBatched launches) on the same points, same GPU"}
            rf = reference_cuda_frame(1, model="car_1", w=512, h=512)
            if rf is not None:
                c1["frame_512_reference_cuda"] = {"ms_per_frame": rf["mean_ms"], "frames_per_s": 1e3 / rf["mean_ms"]}
        line_extra["c1"] = c1

    # ---- C3: 256^3 lattice through car_1, points generated in-kernel (variant B), all rungs
    if args.dense:
        nd = host.NeuralNetwork()
        assert nd.load(geom("car_1"))
        res = 256
        yb = torch.empty(res ** 3, dtype=torch.float32, device="cuda")
        dense = {}
        for label, prec in (("fp32", host.PREC_FP32),) + tuple(args.dense_tc):
            try:
                for _ in range(2):
                    nd.forward_lattice(res, out=yb, precision=prec)
            except capi.NsdfError:
                continue
            ms = timed(torch, lambda: nd.forward_lattice(res, out=yb, precision=prec), 5, flush)
            m = sum(ms) / len(ms)
            dense[label] = {"queries_per_s": res ** 3 / (m * 1e-3), "ms": m, "ms_min": min(ms),
                            "tflops": res ** 3 * FLOP_PER_QUERY / (m * 1e-3) / 1e12,
                            "frac_of_tensor_peak": res ** 3 * FLOP_PER_QUERY / (m * 1e-3) / 1e12 / pk["tflops_burst"]}
        line_extra["dense_256cubed"] = dense

    # ---- the reference's own CUDA on the same box, same frame (view 0)
    if args.ref_cuda:
        rc = reference_cuda_frame(1)
        if rc is not None:
            line_extra["reference_cuda_same_box"] = {"ms_per_frame": rc["mean_ms"], "frames_per_s": 1e3 / rc["mean_ms"],
                                                     "speedup_view0_device_timed": rc["mean_ms"] / single["ms_per_frame"],
                                                     "note": "the reference's unmodified CUDA kernels (oracle/_ref/ref_driver), view 0, same GPU"}

    # ---- CPU baseline + oracle parity: one full frame (view 0, HEAD scene) through the oracle port on all host threads
    if not args.no_cpu:
        dt, r, cores = cpu_oracle_frame()
        line_extra["cpu_baseline"] = {"value": 1.0 / dt, "unit": "frames/s", "cores": cores, "kind": "port",
                                      "sample": f"one full {W}x{H} frame (view 0, scene=manySphere, {r['queries']} queries) through "
                                                "oracle/nsdf_oracle.cpp, all host threads",
                                      "queries_per_s": r["queries"] / dt}
        det, img_det = frame_leg(nn, W, H, reps=1, math=host.MATH_DETERMINISTIC)
        line_extra["parity_vs_oracle"] = {"equal": bool(np.array_equal(img_det, r["image"].reshape(H, W))),
                                          "queries_equal": bool(det["queries"] == r["queries"]), "queries": det["queries"],
                                          "note": "the timed frame's view 0 in NSDF_MATH_DETERMINISTIC against the CPU oracle's frame, all pixels"}

    # which of the library's kernels a step launches (csrc/nsdf_render.cu resolve_tuning): calls of >= 4 views and >= 12 M pixels run on the
    # four-rays-per-lane throughput tracer, smaller ones on the cooperative tracer; calls of < 12 views probe and sort the tiles first
    top_kernel = "trace_bulk4_kernel" if len(my_views) >= 4 and len(my_views) * W * H >= 12000000 else "trace_coop_kernel"
    launches_per_step = 1 if len(my_views) >= 12 else 3
    traffic, traffic_note = None, "no capture of the current tracer sources under profiles/ncu_traffic.json"
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        k = tj.get(top_kernel)
        if k and k.get("sass_sha") and k.get("sass_sha") == kernel_sass_sha(top_kernel):
            per_frame = (k["dram_bytes_read"] + k["dram_bytes_write"]) / k.get("frames_in_launch", 1)
            traffic = per_frame * len(my_views)          # one launch per step and rank = len(my_views) frames
            traffic_note = (f"dram bytes per launch ({len(my_views)} frames), scaled per frame from the ncu --set full capture "
                            f"{tj.get('capture', '')} of this kernel's machine code (equal SASS)")
        else:
            traffic_note = "profiles/ncu_traffic.json was captured from other machine code of this kernel than the library that is running holds: not quoted"
    except Exception:
        pass
    q_frame = q_set / N_VIEWS
    achieved = q_set * FLOP_PER_QUERY / (ms_per_step * 1e-3) / 1e12 / world     # per GPU
    hbm_bytes = 4 * W * H + 4 * hits_set / N_VIEWS + nn.getNumWeightParams() * 4 + nn.getNumBiasParams() * 4
    ms_per_frame = ms_per_step * world / N_VIEWS
    line = {
        "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": world, "steps": K, "warmup": Wm, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "parallelism": f"views{world}: view v -> rank v % {world}" if world > 1 else "single",
                   "frames_per_step": N_VIEWS, "launch": "nsdf_render_views: a rank's views of a step in one persistent launch", "l2": "256 MiB memset between timed steps (outside the per-step events)",
                   "precision": "fp32 FFMA rung, reference arithmetic order; NSDF_MATH_REFERENCE (rsqrtf / __saturatef as the reference)"},
        "ms_per_frame_per_gpu": ms_per_frame, "queries_per_s": q_set * K / (total_ms * 1e-3), "queries_per_frame_mean": q_frame,
        "hit_pixels_per_frame_mean": hits_set / N_VIEWS, "rank_time_min_over_max": min_ms / total_ms if total_ms else None,
        "e2e": {"value": e2e_fps, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": Ke},
        "gpu_launches": launches_per_step * K,   # per step and rank: (tile_probe_kernel, tile_sort_kernel,) the tracer over all the rank's views
        "roofline": {"kernel": top_kernel, "bound": "tensor", "achieved": achieved, "peak": pk["tflops_burst"],
                     "unit": "TFLOP/s", "frac": achieved / pk["tflops_burst"], "traffic": traffic, "traffic_note": traffic_note,
                     "peak_source": pk["source"], "frac_of_fp32_alu": achieved / FP32_ALU_TFLOPS,
                     "note": "per GPU; algorithmic FLOP = MLP evaluations (counted on the device) x 14592 over the step's device time, against "
                             "the measured bf16 tensor peak as SURVEY 8(d) prescribes; the kernel is the bit-exact fp32 FFMA rung whose own "
                             f"ceiling is {FP32_ALU_TFLOPS} TFLOP/s (frac_of_fp32_alu); HBM is not the bound (hbm below)",
                     "hbm": {"algorithmic_bytes_per_frame": hbm_bytes, "achieved_gbs": hbm_bytes / (ms_per_frame * 1e-3) / 1e9,
                             "peak_gbs": pk["hbm_gbs"]}},
        "clocks": clocks, "wall_s": t_wall,
    }
    if tiles:
        line["frame4k_tiles"] = tiles
    if c5:
        line["turntable_c5"] = c5
    if tt4k:
        line["turntable_4k"] = tt4k
    line.update(line_extra)
    print(json.dumps(line), flush=True)
    D.close()


def turntable4k_leg(D, args, host, multigpu, flush):
    """4K frames by view (the --workload turntable4k measurement as a leg of the default line): the 64-view turntable of the 4K model at
    3840x2160, view v -> rank v % N, a rank's views in one nsdf_render_views call; one warm-up pass, one timed pass."""
    torch = D.torch
    model, Wt, Ht = SHAPES[4], 3840, 2160
    nn = host.NeuralNetwork()
    assert nn.load(geom(model))
    matcap = host.Image()
    assert matcap.loadPNG(os.path.join(ROOT, "data", "matcaps", MATCAP))
    host.copyStaticSettings(host.COLOR_MATCAP, 3)
    host.setSceneOptions(host.SCENE_MANYSPHERE, args.max_steps, 0, 0, host.MATH_REFERENCE)
    my_views = multigpu.views_of_rank(N_VIEWS, D.rank, D.world)
    cams = host.view_cameras([(0.0, v * VIEW_STEP_DEG) for v in my_views])
    out = torch.zeros(len(my_views) * Wt * Ht, dtype=torch.int32, device="cuda")
    host.render_views(out, Wt, Ht, nn, cams, matcap)
    D.barrier()
    ms = timed(torch, lambda: host.render_views(out, Wt, Ht, nn, cams, matcap), 1, flush)[0]
    D.barrier()
    total = D.reduce(ms, "max")
    del out
    return {"workload": f"{model}.h5 {Wt}x{Ht}, {N_VIEWS}-view turntable ry=k*{VIEW_STEP_DEG}deg, scene=manySphere (reference HEAD), view v -> rank v % "
                        f"{D.world}, one nsdf_render_views call per rank", "frames": N_VIEWS, "ms_total": total,
            "frames_per_s": N_VIEWS / (total * 1e-3), "ms_per_frame_per_gpu": total * D.world / N_VIEWS}


def run_tiles(args):
    """--workload frame4k_tiles: BASELINE config 4 as the headline of the line (strong scaling over the N GPUs)."""
    from cudaneuralrender_b200 import host, multigpu
    D = Dist()
    flush = D.torch.empty(256 << 20, dtype=D.torch.uint8, device="cuda")
    sampler = ClockSampler(D.local)
    if D.rank == 0:
        sampler.start()
    leg = tiles_leg(D, args, host, multigpu, flush, args.steps)
    clocks = sampler.stop() if D.rank == 0 else None
    if D.rank == 0:
        pk = peaks()
        achieved = leg["queries_per_frame"] * FLOP_PER_QUERY / (leg["ms_per_frame"] * 1e-3) / 1e12
        line = {"metric": "4K frames/sec (3a3d4a90.obj, FD normals + matcap), tile-partitioned", "value": leg["frames_per_s"],
                "unit": "frames/s", "n_gpus": D.world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": leg["ms_per_frame"],
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": leg["workload"], "l2": "256 MiB memset between timed steps (outside the per-step events)"},
                "e2e": {"value": leg["e2e_frames_per_s"], "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 3840 * 2160 * 4},
                "gpu_launches": 3 * args.steps * D.world,
                "roofline": {"kernel": "trace_coop_kernel", "bound": "tensor", "achieved": achieved, "peak": pk["tflops_burst"] * D.world,
                             "unit": "TFLOP/s", "frac": achieved / (pk["tflops_burst"] * D.world), "traffic": None,
                             "peak_source": pk["source"], "frac_of_fp32_alu": achieved / (FP32_ALU_TFLOPS * D.world)},
                "clocks": clocks, "frame4k_tiles": leg}
        print(json.dumps(line), flush=True)
    D.close()


def run_turntable4k(args):
    """--workload turntable4k: the 64-view turntable of the 4K model (BASELINE config 4's shape and size, config 5's view set), view v ->
    rank v % N, a rank's views of a step in ONE nsdf_render_views call: 4K FRAMES spread over the GPUs by view (strong scaling; the
    frame4k_tiles workload splits ONE 4K frame by tiles instead)."""
    from cudaneuralrender_b200 import host, multigpu
    D = Dist()
    torch = D.torch
    model, Wt, Ht = SHAPES[4], 3840, 2160
    nn = host.NeuralNetwork()
    assert nn.load(geom(model))
    matcap = host.Image()
    assert matcap.loadPNG(os.path.join(ROOT, "data", "matcaps", MATCAP))
    host.copyStaticSettings(host.COLOR_MATCAP, 3)
    host.setSceneOptions(host.SCENE_MANYSPHERE, args.max_steps, 0, 0, host.MATH_REFERENCE)
    my_views = multigpu.views_of_rank(N_VIEWS, D.rank, D.world)
    cams = host.view_cameras([(0.0, v * VIEW_STEP_DEG) for v in my_views])
    out = torch.zeros(len(my_views) * Wt * Ht, dtype=torch.int32, device="cuda")
    pinned = torch.empty(len(my_views) * Wt * Ht, dtype=torch.int32, pin_memory=True)
    stats = torch.zeros(6, dtype=torch.int64, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    host.render_views(out, Wt, Ht, nn, cams, matcap, stats=stats)          # accounting pass (also a warm-up)
    torch.cuda.synchronize()
    q_set = D.reduce(float(stats[0].item()), "sum")
    for _ in range(max(args.warmup - 1, 2)):
        host.render_views(out, Wt, Ht, nn, cams, matcap)
    sampler = ClockSampler(D.local)
    if D.rank == 0:
        sampler.start()
    D.barrier()
    ms = timed(torch, lambda: host.render_views(out, Wt, Ht, nn, cams, matcap), args.steps, flush)
    D.barrier()
    total_ms = D.reduce(sum(ms), "max")
    min_ms = D.reduce(sum(ms), "min")
    clocks = sampler.stop() if D.rank == 0 else None
    # end to end: the rank's frames back in pinned host memory inside the timed region
    D.barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        host.render_views(out, Wt, Ht, nn, cams, matcap)
        pinned.copy_(out, non_blocking=True)
        torch.cuda.synchronize()
    e2e_s = D.reduce(time.perf_counter() - t0, "max")
    if D.rank == 0:
        pk = peaks()
        K = args.steps
        ms_per_step = total_ms / K
        achieved = q_set * FLOP_PER_QUERY / (ms_per_step * 1e-3) / 1e12 / D.world
        top_kernel = "trace_bulk4_kernel" if len(my_views) >= 4 else "trace_coop_kernel"
        line = {"metric": "4K frames/sec (3a3d4a90.obj, FD normals + matcap), 64-view turntable, views over the GPUs",
                "value": N_VIEWS * K / (total_ms * 1e-3), "unit": "frames/s", "n_gpus": D.world, "steps": K, "warmup": args.warmup,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic",
                "config": {"workload": f"{model}.h5 {Wt}x{Ht} sphere-traced frames, {N_VIEWS}-view turntable ry=k*{VIEW_STEP_DEG}deg per step, FD "
                                       f"normals + matcap {MATCAP}, scene=manySphere (reference HEAD), max_steps={args.max_steps}",
                           "parallelism": f"views{D.world}: view v -> rank v % {D.world}", "frames_per_step": N_VIEWS,
                           "l2": "256 MiB memset between timed steps (outside the per-step events)"},
                "ms_per_frame_per_gpu": ms_per_step * D.world / N_VIEWS, "queries_per_frame_mean": q_set / N_VIEWS,
                "rank_time_min_over_max": min_ms / total_ms if total_ms else None,
                "e2e": {"value": N_VIEWS * K / e2e_s, "unit": "frames/s", "h2d_bytes_per_step": 128 * N_VIEWS,
                        "d2h_bytes_per_step": Wt * Ht * 4 * N_VIEWS},
                "gpu_launches": K,
                "roofline": {"kernel": top_kernel, "bound": "tensor", "achieved": achieved, "peak": pk["tflops_burst"], "unit": "TFLOP/s",
                             "frac": achieved / pk["tflops_burst"], "traffic": None, "peak_source": pk["source"],
                             "frac_of_fp32_alu": achieved / FP32_ALU_TFLOPS},
                "clocks": clocks}
        print(json.dumps(line), flush=True)
    D.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--max-steps", type=int, default=6000, dest="max_steps")
    ap.add_argument("--e2e-steps", type=int, default=0, dest="e2e_steps", help="0 = as many as --steps")
    ap.add_argument("--no-dense", action="store_false", dest="dense")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-mixed", action="store_false", dest="mixed")
    ap.add_argument("--no-c1", action="store_false", dest="c1")
    ap.add_argument("--no-c5", action="store_false", dest="c5")
    ap.add_argument("--no-tiles", action="store_false", dest="tiles")
    ap.add_argument("--no-4k", action="store_false", dest="tt4k", help="skip the 4K turntable leg (4K frames by view)")
    ap.add_argument("--cpu-port", action="store_true", help="--impl reference: time the CPU oracle port even if the reference's "
                                                            "CUDA build (oracle/_ref) is present")
    ap.add_argument("--no-ref-cuda", action="store_false", dest="ref_cuda")
    ap.add_argument("--workload", default="turntable1080", choices=["turntable1080", "frame4k_tiles", "turntable4k"],
                    help="turntable1080: BASELINE config 2 over the fixed 64-view set (default).  frame4k_tiles: BASELINE config 4 "
                         "alone, ONE 3840x2160 frame tile-partitioned over the N GPUs (also a leg of every default line)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "native" else args.warmup
    args.dense_tc = (("tc_fp16", 1), ("tc_bf16", 2), ("tc_fp16x3", 3), ("tc_fp16_acc16", 5))
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "frame4k_tiles":
        run_tiles(args)
    elif args.workload == "turntable4k":
        run_turntable4k(args)
    else:
        run_native(args)


if __name__ == "__main__":
    main()

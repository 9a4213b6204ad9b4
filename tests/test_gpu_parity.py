"""GPU parity tests (-m gpu): the CUDA path, called through the C ABI, against the CPU oracle.

Bars: fp32 rungs and the renderer are BIT-EXACT against the oracle (integer comparison of the
float bit patterns / packed RGBA); tensor-core rungs within their stated bounds."""
import ctypes as C
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN, MATCAPS, MODELS, model_path

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env(gpu):
    import torch
    from cudaneuralrender_b200 import host
    from cudaneuralrender_b200 import _capi as capi
    from oracle import oracle
    return {"torch": torch, "host": host, "oracle": oracle, "capi": capi}


def _load(env, name):
    nn = env["host"].NeuralNetwork()
    assert nn.load(model_path(name))
    return nn, env["oracle"].Model.from_h5(model_path(name))


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _matrix(host, x):
    """Matrix(Shape(in, B)) holding the rows of x (element [b*in + c]), on the device."""
    m = host.Matrix(x.shape[1], x.shape[0])
    m.allocateMemory()
    m.hostData[:] = x.ravel()
    m.copyHostToDevice()
    return m


def test_native_library_is_loaded(env):
    from cudaneuralrender_b200 import _capi
    assert os.path.exists(_capi.LIB_PATH)
    assert _capi.lib().nsdf_device_sm_count() >= 100
    maps = open("/proc/self/maps").read()
    assert os.path.basename(_capi.LIB_PATH) in maps


def test_known_answers_on_gpu(env):
    """SURVEY App. B bit patterns through nsdf_mlp_forward (both fp32 evaluators)."""
    torch, host = env["torch"], env["host"]
    rows = json.load(open(os.path.join(GOLDEN, "kat_sdf.json")))["rows"]
    for name in MODELS:
        nn, _ = _load(env, name)
        mine = [r for r in rows if r["model"] == name]
        x = torch.tensor([r["point"] for r in mine], dtype=torch.float32, device="cuda")
        for prec in (host.PREC_FP32, host.PREC_FP32_WARP):
            y = nn.forward_tensor(x.contiguous(), prec).cpu().numpy()
            assert [int(b) for b in _bits(y)] == [int(r["bits"], 16) for r in mine], (name, prec)


@pytest.mark.parametrize("n", [0, 1, 31, 32, 33, 255, 256, 257, 65536, 2 ** 20, 1000003])
def test_fp32_mlp_bit_exact(env, n):
    """Random points, ragged sizes, and the multiples of 65 536 on which the reference's
    GemmBatched grid.z bug fails."""
    torch, host, oracle = env["torch"], env["host"], env["oracle"]
    nn, om = _load(env, "car_1")
    x = np.random.default_rng(n).uniform(-1, 1, (n, 3)).astype(np.float32)
    ref = oracle.mlp_forward(om, x).ravel()
    xd = torch.from_numpy(x).cuda().contiguous()
    for prec in (host.PREC_FP32, host.PREC_FP32_WARP):
        if prec == host.PREC_FP32_WARP and n > 2 ** 20:
            continue
        y = nn.forward_tensor(xd, prec).cpu().numpy()
        assert np.array_equal(_bits(y), _bits(ref)), prec
    if n:
        assert np.array_equal(_bits(nn.forward_host(x)), _bits(ref))


def test_fp32_thread_evaluator_variants_bit_exact(env, monkeypatch):
    """Every thread-per-point schedule of the fp32 rung (1, 2 points per thread in registers; 4 points per thread with the
    activations in shared memory at 128/256/384 threads) returns the oracle's bits, on ragged sizes too.  The variants exist in
    the development build only (make -C cudaneuralrender_b200/csrc dev, NSDF_LIB=dev); the product's one schedule is covered
    by test_fp32_bit_exact_ragged_sizes."""
    torch, host, oracle = env["torch"], env["host"], env["oracle"]
    if not host.dev_schedules_available():
        pytest.skip("development rungs are not in the product library")
    nn, om = _load(env, "plane_1")
    for n in (1, 1535, 1536, 1537, 300001):
        x = np.random.default_rng(n).uniform(-1, 1, (n, 3)).astype(np.float32)
        ref = oracle.mlp_forward(om, x).ravel()
        xd = torch.from_numpy(x).cuda().contiguous()
        for p, th in (("1", None), ("2", None), ("22", None), ("4", "128"), ("4", "256"), ("4", "384"), ("40", None), ("41", None),
                      (None, None)):
            for k, v in (("NSDF_FP32_P", p), ("NSDF_FP32_P4_THREADS", th)):
                if v is None:
                    monkeypatch.delenv(k, raising=False)
                else:
                    monkeypatch.setenv(k, v)
            y = nn.forward_tensor(xd, host.PREC_FP32).cpu().numpy()
            assert np.array_equal(_bits(y), _bits(ref)), (n, p, th)


def test_fp32_warp_evaluator_widths_bit_exact(env, monkeypatch):
    """The tracer's warp-cooperative evaluators on dense batches: 1/2/4/8 points per warp (lane = neuron) and the
    register-tiled 16-point form (mlp32_eval_warp16) return the oracle's bits, on ragged sizes and on every model."""
    torch, host, oracle = env["torch"], env["host"], env["oracle"]
    for name in MODELS:
        nn, om = _load(env, name)
        for n in (1, 15, 16, 17, 4095, 100003):
            x = np.random.default_rng(n).uniform(-1, 1, (n, 3)).astype(np.float32)
            ref = oracle.mlp_forward(om, x).ravel()
            xd = torch.from_numpy(x).cuda().contiguous()
            for r in (("79", "78", "77", "16", "8", "4", "2", "1") if host.dev_schedules_available() else ("4",)):
                monkeypatch.setenv("NSDF_FP32_WARP_R", r)    # read by the development build only; the product runs R = 4
                y = nn.forward_tensor(xd, host.PREC_FP32_WARP).cpu().numpy()
                assert np.array_equal(_bits(y), _bits(ref)), (name, n, r)
    monkeypatch.delenv("NSDF_FP32_WARP_R", raising=False)


def test_fp32_all_models_and_special_values(env):
    torch, host, oracle = env["torch"], env["host"], env["oracle"]
    rng = np.random.default_rng(5)
    x = rng.uniform(-1.5, 1.5, (40000, 3)).astype(np.float32)
    x[:8] = [[0, 0, 0], [-0.0, 0.0, -0.0], [1e-30, -1e-30, 1e-38], [1e4, -1e4, 1e4], [np.inf, 0, 0], [np.nan, 0, 0],
             [1, 1, 1], [-1, -1, -1]]
    for name in MODELS:
        nn, om = _load(env, name)
        y = nn.forward_tensor(torch.from_numpy(x).cuda().contiguous()).cpu().numpy()
        ref = oracle.mlp_forward(om, x).ravel()
        finite = np.isfinite(ref)
        assert np.array_equal(_bits(y[finite]), _bits(ref[finite])), name
        assert np.array_equal(np.isnan(y), np.isnan(ref))


def test_equal_inputs_equal_outputs_like_simpleInfer(env):
    """src/simpleInfer.cpp:112-147: 1 000 000 zero points, all outputs identical."""
    torch, host = env["torch"], env["host"]
    nn, _ = _load(env, "car_1")
    X = host.Matrix(host.Shape(3, 1000000))
    X.allocateMemory()
    X.copyHostToDevice()
    Y = nn.forward(X)
    assert (Y.shape.x, Y.shape.y) == (1, 1000000)
    Y.copyDeviceToHost()
    assert np.all(Y.hostData[:1000000] == Y[0])
    assert _bits(Y.hostData[:1])[0] == 0xBDF2F39C


def test_layered_operators_match_fused(env):
    """DenseLayer::forward chain (nsdf_dense_layer_forward) == fused kernel == oracle."""
    torch, host, oracle = env["torch"], env["host"], env["oracle"]
    nn, om = _load(env, "plane_2")
    n = 5000
    x = np.random.default_rng(2).uniform(-1, 1, (n, 3)).astype(np.float32)
    X = host.Matrix(host.Shape(3, n))
    X.allocateMemory()
    X.hostData[:] = x.ravel()
    X.copyHostToDevice()
    Z = nn.forward_layered(X)
    Z.copyDeviceToHost()
    ref = oracle.mlp_forward(om, x).ravel()
    assert np.array_equal(_bits(Z.hostData[:n]), _bits(ref))
    assert nn.layers[0].getXDim() == 3 and nn.layers[0].getYDim() == 32 and nn.getNumWeightParams() == 7296


def test_other_shapes_run_fused_and_render(env, tmp_path):
    """the reference's forward and renderer take ANY dense stack (src/layers/denseLayer.cu:180-227,
    src/volumeRender_kernel.cu:652-689).  Shapes other than (3|4) -> 32 x k -> 1 run through the generic fused evaluator
    (csrc/mlp_generic.cuh): ONE launch per forward, bits of the oracle, for widths 16 / 64 / 128, mixed widths, several
    outputs, 4 inputs; and their frames -- networks written with nsdf_h5_write_keras and loaded back -- equal the oracle's
    frames byte for byte (a 64- and a 128-wide net that embed car_1, a random 16-wide and a mixed-width net)."""
    torch, host, oracle, capi = env["torch"], env["host"], env["oracle"], env["capi"]
    for dims in ([3, 16, 8, 1], [5, 64, 64, 2], [3, 32, 1], [3, 32, 32, 32, 1], [3, 128, 128, 1], [4, 64, 16, 128, 1], [3, 1]):
        om = oracle.Model.random(dims, seed=len(dims))
        nn = host.NeuralNetwork()
        for i, (W, b) in enumerate(om.layers()):
            nn.addLayer(host.DenseLayer(f"dense_{i}", W.T, b, host.Tanh if i == len(dims) - 2 else host.ReLU))
        x = np.random.default_rng(0).uniform(-1, 1, (3001, dims[0])).astype(np.float32)
        y = nn.forward_tensor(torch.from_numpy(x).cuda().contiguous()).cpu().numpy()
        assert np.array_equal(_bits(y), _bits(oracle.mlp_forward(om, x).ravel())), dims
        z = nn.forward_layered(_matrix(host, x)).deviceData[: y.size].cpu().numpy()      # the per-layer operators agree
        assert np.array_equal(_bits(z), _bits(y)), dims

    def embed(layers, width):
        """car_1's 32-wide layers inside `width`-wide ones (extra neurons have zero weights and biases)."""
        out = []
        for i, (k, b) in enumerate(layers):
            kin = k.shape[0] if i == 0 else width
            kout = k.shape[1] if i == len(layers) - 1 else width
            K = np.zeros((kin, kout), np.float32)
            B = np.zeros(kout, np.float32)
            K[:k.shape[0], :k.shape[1]] = k
            B[:b.size] = b
            out.append((K, B))
        return out

    car = host.read_keras_h5(model_path("car_1"))
    rng = np.random.default_rng(3)
    nets = {"w64": embed(car, 64), "w128": embed(car[:4] + car[-1:], 128),
            "w16": [(0.6 * rng.standard_normal((3, 16)).astype(np.float32), 0.1 * rng.standard_normal(16).astype(np.float32)),
                    (0.4 * rng.standard_normal((16, 16)).astype(np.float32), 0.1 * rng.standard_normal(16).astype(np.float32)),
                    (0.3 * rng.standard_normal((16, 1)).astype(np.float32), np.array([0.05], np.float32))],
            "mixed": embed(car[:3], 64)[:2] + [(np.pad(car[2][0], ((0, 32), (0, 0))), car[2][1])] + car[3:]}
    matcap = oracle.load_matcap_png(os.path.join(MATCAPS, "blue.png"))
    iv, nm = oracle.view_matrices(0.0, 25.0)
    host.copyStaticSettings(host.COLOR_MATCAP, 3)
    host.copyViewMatrices(iv, nm, 0, 3)
    W, H = 120, 88
    for name, layers in nets.items():
        path = str(tmp_path / (name + ".h5"))
        host.write_keras_h5(path, layers)
        nn = host.NeuralNetwork()
        assert nn.load(path)
        assert capi.lib().nsdf_model_is_fused(nn.handle()) == 0
        om = oracle.Model.from_h5(path)
        for scene in (host.SCENE_TANH, host.SCENE_MANYSPHERE):
            host.setSceneOptions(scene, 6000)
            img, stats = host.render_host(nn, W, H, matcap)
            ref = oracle.render(om, iv, nm, W, H, matcap=matcap, scene=oracle.SCENE_TANH if scene == host.SCENE_TANH else oracle.SCENE_MANYSPHERE)
            assert np.array_equal(img, ref["image"]), (name, scene)
            assert stats["queries"] == ref["queries"] and stats["pixels"] == W * H, (name, scene)
        if name in ("w64", "w128"):
            assert int((ref["image"] != 0).sum()) > 200      # the embedded car is there
    host.setSceneOptions()


def test_lattice_matches_explicit_points(env):
    """BASELINE config 3 at reduced resolution: in-kernel lattice == explicit points == oracle."""
    torch, host, oracle = env["torch"], env["host"], env["oracle"]
    nn, om = _load(env, "car_1")
    res = 48
    y = nn.forward_lattice(res).cpu().numpy()
    i = np.arange(res, dtype=np.float32)
    s = np.float32(2.0) / np.float32(res)
    c = ((i + np.float32(0.5)).astype(np.float64) * np.float64(s) + np.float64(-1.0)).astype(np.float32)  # fmaf
    zz, yy, xx = np.meshgrid(c, c, c, indexing="ij")
    pts = np.stack([xx.ravel(), yy.ravel(), zz.ravel()], axis=1).astype(np.float32)
    ref = oracle.mlp_forward(om, pts).ravel()
    assert np.array_equal(_bits(y), _bits(ref))
    part = nn.forward_lattice(res, first=1000, count=777).cpu().numpy()
    assert np.array_equal(_bits(part), _bits(ref[1000:1777]))


# Stated bounds of the tensor-core rungs (absolute SDF error against the oracle; include/nsdf.h):
# FP16X3 meets the north star's fp32 bound of 1e-4, the single-pass rungs carry the looser ones.
TC_BOUNDS = {"PREC_TC_FP16X3": 1e-4, "PREC_TC_FP16": 4e-3, "PREC_TC_FP16_ACC16": 2e-2, "PREC_TC_BF16": 5e-2}


@pytest.mark.parametrize("name", MODELS)
def test_tc_rungs_within_stated_bounds(env, name):
    torch, host, oracle = env["torch"], env["host"], env["oracle"]
    nn, om = _load(env, name)
    x = np.random.default_rng(7).uniform(-1, 1, (200_000, 3)).astype(np.float32)
    x[:8] = [[0, 0, 0], [1, 1, 1], [-1, -1, -1], [1, -1, 1], [0.5, 0, 0], [0, -0.5, 0], [0, 0, 0.25], [1e-6, -1e-6, 1e-3]]
    ref = oracle.mlp_forward(om, x).ravel().astype(np.float64)
    xd = torch.from_numpy(x).cuda().contiguous()
    for pname, bound in TC_BOUNDS.items():
        y = nn.forward_tensor(xd, getattr(host, pname)).cpu().numpy().astype(np.float64)
        err = np.abs(y - ref)
        assert np.isfinite(y).all(), pname
        assert err.max() <= bound, (name, pname, err.max())


@pytest.mark.parametrize("kernel,cfgvar,cfgs", [("2", None, [0]), ("3", "NSDF_TC3_CFG", [0, 2, 4, 7]), ("4", "NSDF_TC4_CFG", [0, 1, 3])])
def test_tc_schedules_within_bounds(env, kernel, cfgvar, cfgs):
    """The A/B schedules of the tensor-core rungs (csrc/tc_mlp.cuh, tc_mlp3.cuh, tc_mlp4.cuh; selected by environment
    variables read at call time) compute the same network: each stays within the rung's stated bound, on a ragged size."""
    import os
    torch, host, oracle = env["torch"], env["host"], env["oracle"]
    if kernel != "2" and not host.dev_schedules_available():
        pytest.skip("A/B schedules are selected through the environment in the development build only")
    nn, om = _load(env, "plane_2")
    n = 300_007
    x = np.random.default_rng(11).uniform(-1, 1, (n, 3)).astype(np.float32)
    ref = oracle.mlp_forward(om, x).ravel().astype(np.float64)
    xd = torch.from_numpy(x).cuda().contiguous()
    old = {k: os.environ.get(k) for k in ("NSDF_TC_KERNEL", "NSDF_TC3_CFG", "NSDF_TC4_CFG")}
    try:
        os.environ["NSDF_TC_KERNEL"] = kernel
        for cfg in cfgs:
            if cfgvar:
                os.environ[cfgvar] = str(cfg)
            for pname in ("PREC_TC_FP16", "PREC_TC_BF16", "PREC_TC_FP16X3"):
                y = nn.forward_tensor(xd, getattr(host, pname)).cpu().numpy().astype(np.float64)
                assert np.isfinite(y).all()
                assert np.abs(y - ref).max() <= TC_BOUNDS[pname], (kernel, cfg, pname, np.abs(y - ref).max())
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


@pytest.mark.parametrize("n", [1, 127, 128, 129, 4097, 1000003])
def test_tc_ragged_sizes(env, n):
    """Tile tails: rows past n are computed on zeros and never stored."""
    torch, host, oracle = env["torch"], env["host"], env["oracle"]
    nn, om = _load(env, "car_1")
    x = np.random.default_rng(n).uniform(-1, 1, (n, 3)).astype(np.float32)
    ref = oracle.mlp_forward(om, x).ravel()
    xd = torch.from_numpy(x).cuda().contiguous()
    guard = torch.full((n + 256,), 12345.0, dtype=torch.float32, device="cuda")
    nn.forward_tensor(xd, host.PREC_TC_FP16X3, out=guard)
    g = guard.cpu().numpy()
    assert np.abs(g[:n].astype(np.float64) - ref).max() <= 1e-4
    assert (g[n:] == 12345.0).all()


def test_tc_lattice_matches_explicit_points(env):
    torch, host = env["torch"], env["host"]
    nn, _ = _load(env, "plane_1")
    res = 48
    idx = np.arange(res ** 3)
    c = (-1.0 + 2.0 * (np.stack([idx % res, (idx // res) % res, idx // (res * res)], 1) + 0.5) / res).astype(np.float32)
    # lattice points are generated with one fma in-kernel and may differ from numpy's by an ulp; a
    # single-pass rung rounds activations to 16 bits, so an ulp of input can move the output by a
    # rounding quantum: the tolerance is the rung's own bound there
    for prec, tol in ((host.PREC_TC_FP16, 4e-3), (host.PREC_TC_FP16X3, 1e-5)):
        a = nn.forward_lattice(res, precision=prec).cpu().numpy()
        b = nn.forward_tensor(torch.from_numpy(c).cuda().contiguous(), prec).cpu().numpy()
        assert np.abs(a - b).max() <= tol, prec


def _render_both(env, name, W, H, matcap_name=None, scene=0, max_steps=6000, rx=0.0, ry=0.0, flags=0, frame=0,
                 evict_steps=0):
    torch, host, oracle = env["torch"], env["host"], env["oracle"]
    nn, om = _load(env, name)
    matcap = oracle.load_matcap_png(os.path.join(MATCAPS, matcap_name)) if matcap_name else None
    iv, nm = oracle.view_matrices(rx, ry)
    host.copyStaticSettings(host.COLOR_MATCAP if matcap is not None else host.COLOR_FACING, 3)
    host.copyViewMatrices(iv, nm, frame, 3)
    host.setSceneOptions(scene, max_steps, flags, evict_steps)
    img, stats = host.render_host(nn, W, H, matcap)
    host.setSceneOptions()
    ref = oracle.render(om, iv, nm, W, H, matcap=matcap, scene=scene, max_steps=max_steps, frame=frame)
    return img, stats, ref


@pytest.mark.parametrize("name,W,H,matcap", [("car_1", 128, 128, "blue.png"), ("plane_1", 160, 90, "red.png"),
                                              ("plane_3", 97, 61, None), ("car_1", 64, 64, "ball-298x300.png"),
                                              ("3a3d4a90a2db90b4203936772104a82d.obj", 130, 70, "Normal.png")])
def test_frames_bit_exact(env, name, W, H, matcap):
    img, stats, ref = _render_both(env, name, W, H, matcap)
    hit = ref["hit"] != 0
    assert np.array_equal(img != 0, (ref["image"] != 0)), "hit masks differ"
    assert np.array_equal(img, ref["image"]), "RGBA differs"
    assert stats["queries"] == ref["queries"]
    assert stats["hit_pixels"] == int(hit.sum())
    assert stats["pixels"] == W * H


def test_frame_schedules_agree(env):
    """Simple kernel, ray-per-lane kernel, the default cooperative kernel, the two-launch schedule and the
    producer -> queue -> consumer kernel with aggressive hand-over thresholds (every ray crosses the queue)
    all produce the same bytes (schedule independence)."""
    host = env["host"]
    c = host.capi
    outs = []
    product = [(c.RENDER_SIMPLE, 0), (0, 0), (c.RENDER_NO_QUIET_SMS, 0), (c.RENDER_STAGED_STORES, 0)]
    dev = [(c.RENDER_NO_STRAGGLER, 0), (0, 4), (c.RENDER_TWO_KERNELS, 4), (c.RENDER_TWO_KERNELS, 0), (c.RENDER_QUEUE, 1),
           (c.RENDER_QUEUE, 37), (c.RENDER_QUEUE, 0), (c.RENDER_TWO_PASS, 0), (c.RENDER_TWO_PASS, 3), (c.RENDER_HYBRID, 0),
           (c.RENDER_HYBRID, 5)]
    for flags, ev in product + (dev if host.dev_schedules_available() else []):
        img, stats, ref = _render_both(env, "car_1", 96, 96, "blue.png", flags=flags, evict_steps=ev)
        assert np.array_equal(img, ref["image"]), (flags, ev)
        assert stats["queries"] == ref["queries"]
        outs.append(stats)
    assert outs[0]["evicted"] == 0
    if not host.dev_schedules_available():      # the product library answers the A/B flags with NSDF_EUNSUPPORTED, never silently
        with pytest.raises(c.NsdfError):
            _render_both(env, "car_1", 96, 96, "blue.png", flags=c.RENDER_TWO_PASS)


def test_express_lanes_bit_exact(env):
    """The cooperative tracer's express lanes (full grids only, so a frame larger than 148 CTAs x 256 ray slots) under
    nsdf_tuning: default thresholds, thresholds that send nearly every ray through the express queue, sparse express SMs,
    several rays per lane, the tail hand-over, express lanes off, tile priority forced on -- same bytes as the oracle, same
    query count."""
    host = env["host"]
    W, H = 256, 192
    ref = None
    try:
        for knobs in ({}, {"express_steps": 16}, {"express_steps": 32, "express_period": 37}, {"express_steps": 48, "express_from_bulk": 0},
                      {"express_steps": 24, "express_rays": 4, "express_solo_steps": 64}, {"express_steps": 16, "express_warps": 16},
                      {"express_steps": 32, "tail_lanes": 4, "tail_push_max": 4}, {"express_period": 0},
                      {"tile_priority_min_jobs": 1, "tile_priority_steps": 8, "express_steps": 32}):
            host.setTuning(**knobs)
            img, stats, r = _render_both(env, "plane_1", W, H, "blue.png")
            ref = ref or r
            assert np.array_equal(img, ref["image"]), knobs
            assert stats["queries"] == ref["queries"], knobs
            assert stats["pixels"] == W * H
    finally:
        host.setTuning()


def test_animation_net_frames_bit_exact(env, tmp_path):
    """4-input (x, y, z, t) nets, the reference's --animation mode (src/volumeRender_kernel.cu:524-545, t = frame number):
    a shipped shape with a fourth input row added to its first layer is written with nsdf_h5_write_keras, loaded through
    the product reader and rendered at several frame numbers by every schedule; frames equal the oracle's bytes, frame 0
    equals the 3-input shape's frame (the fourth product is exactly 0), later frames differ from it."""
    torch, host, oracle, capi = env["torch"], env["host"], env["oracle"], env["capi"]
    layers = host.read_keras_h5(model_path("car_1"))
    rng = np.random.default_rng(11)
    k0, b0 = layers[0]
    layers[0] = (np.concatenate([k0, (0.02 * rng.standard_normal((1, k0.shape[1]))).astype(np.float32)], axis=0), b0)
    path = str(tmp_path / "car_xyzt.h5")
    host.write_keras_h5(path, layers)
    nn = host.NeuralNetwork()
    assert nn.load(path) and nn.layers[0].getXDim() == 4
    om = oracle.Model.from_h5(path)
    matcap = oracle.load_matcap_png(os.path.join(MATCAPS, "blue.png"))
    iv, nm = oracle.view_matrices(0.0, 20.0)
    W, H = 112, 80
    host.copyStaticSettings(host.COLOR_MATCAP, 4)
    frames = {}
    try:
        for frame in (0, 3, 17):
            host.copyViewMatrices(iv, nm, frame, 4)
            ref = oracle.render(om, iv, nm, W, H, matcap=matcap, frame=frame)
            dev = (capi.RENDER_NO_STRAGGLER, capi.RENDER_TWO_PASS, capi.RENDER_HYBRID, capi.RENDER_QUEUE) if host.dev_schedules_available() else ()
            for flags in (0, capi.RENDER_SIMPLE, capi.RENDER_STAGED_STORES) + dev:
                host.setSceneOptions(host.SCENE_TANH, 6000, flags, 0)
                img, stats = host.render_host(nn, W, H, matcap)
                assert np.array_equal(img, ref["image"]), (frame, flags)
                assert stats["queries"] == ref["queries"]
            frames[frame] = ref["image"]
        img3, _, ref3 = _render_both(env, "car_1", W, H, "blue.png", ry=20.0)
        assert np.array_equal(frames[0], ref3["image"]) and np.array_equal(img3, frames[0])
        assert not np.array_equal(frames[3], frames[0]) and not np.array_equal(frames[17], frames[3])
        # the tensor-core far field (development build) is a 3-input path; the product library does not have it at all
        host.setSceneOptions(host.SCENE_TANH, 6000, capi.RENDER_TC_FARFIELD, 0)
        with pytest.raises(Exception):
            host.render_host(nn, W, H, matcap)
    finally:
        host.setSceneOptions()
        host.copyStaticSettings(host.COLOR_MATCAP, 3)
    # dense forward of the same net: layer-by-layer path, bits of the oracle
    x = rng.uniform(-1, 1, (4096, 4)).astype(np.float32)
    y = nn.forward_host(x)
    assert np.array_equal(_bits(y), _bits(oracle.mlp_forward(om, x).ravel()))


def _scene_programs(O):
    """Scene programs covering every primitive, modifier and combiner, among them the scenes sceneSDF holds commented out
    (src/volumeRender_kernel.cu:217-230)."""
    return {
        "round(tanh, 0.04) :220": [(O.PRIM_TANH, O.COMB_SET, []), (O.MOD_ROUND, 0, [0.04])],
        "displacementPattern :218": [(O.PRIM_TANH, O.COMB_SET, []), (O.MOD_DISPLACE, 0, [5.0, 0.05])],
        "sphere 0.9 :219": [(O.PRIM_SPHERE, O.COMB_SET, [0, 0, 0, 0.9])],
        "cylinder smooth-cut out of tanh :224-227": [(O.PRIM_TANH, O.COMB_SET, []),
                                                      (O.PRIM_CYLINDER, O.COMB_SMOOTH_SUBTRACT_FROM, [-0.5, -0.2, 0.4, 0.1, 0.1, 0.1, 0, 0.05])],
        "manySphere subtract": [(O.PRIM_NSDF, O.COMB_SET, []), (O.MOD_MANY_SPHERE, 0, [0.0])],
        "manyCylinderCut :222": [(O.PRIM_NSDF, O.COMB_SET, []), (O.MOD_MANY_CYLINDER_CUT, 0, [])],
        "box / plane / onion / hard booleans": [(O.PRIM_TANH, O.COMB_SET, []), (O.MOD_ONION, 0, [0.02]),
                                                 (O.PRIM_BOX, O.COMB_INTERSECT, [0.0, 0.1, 0.0, 0.6, 0.5, 0.7, 0.03]),
                                                 (O.PRIM_SPHERE, O.COMB_SUBTRACT, [0.3, 0.2, 0.4, 0.25]),
                                                 (O.PRIM_PLANE, O.COMB_UNION, [0.5]),
                                                 (O.PRIM_SPHERE, O.COMB_SUBTRACT_FROM, [0, 0, 0, 1.1]),
                                                 (O.PRIM_SPHERE, O.COMB_SMOOTH_UNION, [-0.4, 0.3, 0.2, 0.2, 0, 0, 0, 0.05]),
                                                 (O.PRIM_CYLINDER, O.COMB_SMOOTH_SUBTRACT, [0.1, 0.1, 0, 0.0, 0.08, 0.0, 0, 0.03])],
    }


def test_scene_program_points_bit_exact(env):
    """nsdf_scene_eval (the analytic SDF library of src/volumeRender_kernel.cu:67-196 composed at run time) against the
    oracle's restatement on random points and network outputs: same bits for every program."""
    torch, host, oracle = env["torch"], env["host"], env["oracle"]
    rng = np.random.default_rng(5)
    n = 100000
    p = rng.uniform(-1.3, 1.3, (n, 3)).astype(np.float32)
    ns = np.concatenate([rng.uniform(-0.5, 0.8, n - 6), [0.0, -0.0, 1e-7, -1e-7, 30.0, -30.0]]).astype(np.float32)
    dp, dn = torch.from_numpy(p).cuda(), torch.from_numpy(ns).cuda()
    try:
        for name, ops in _scene_programs(oracle).items():
            for frame in (0, 45):
                oracle.set_scene_program(ops)
                want = oracle.scene_program_eval(p, ns, frame)
                got = host.scene_eval(ops, dp, dn, frame).cpu().numpy()
                assert np.array_equal(_bits(got), _bits(want)), (name, frame, int((_bits(got) != _bits(want)).sum()))
    finally:
        oracle.set_scene_program([])
    bad = host.capi.scene_ops([(99, 0, [])])
    with pytest.raises(Exception):
        host.scene_eval(bad, dp, dn, 0)


def test_scene_program_frames_bit_exact(env):
    """Frames of program scenes equal the oracle's bytes (hit mask, RGBA, query count) for the default and the ray-per-lane
    schedules; the two compiled scenes rendered as programs equal their compiled frames."""
    torch, host, oracle, capi = env["torch"], env["host"], env["oracle"], env["capi"]
    nn, om = _load(env, "car_1")
    matcap = oracle.load_matcap_png(os.path.join(MATCAPS, "blue.png"))
    iv, nm = oracle.view_matrices(10.0, 25.0)
    W, H = 96, 72
    host.copyStaticSettings(host.COLOR_MATCAP, 3)
    host.copyViewMatrices(iv, nm, 30, 3)
    try:
        for name, ops in _scene_programs(oracle).items():
            heavy = "manyCylinderCut" in name
            oracle.set_scene_program(ops)
            ref = oracle.render(om, iv, nm, W, H, matcap=matcap, scene=oracle.SCENE_PROGRAM, frame=30,
                                max_steps=300 if heavy else 6000)
            dev = (capi.RENDER_NO_STRAGGLER, capi.RENDER_HYBRID) if host.dev_schedules_available() else ()
            for flags in ((0,) if heavy else (0, capi.RENDER_SIMPLE) + dev):
                host.setSceneProgram(ops, max_steps=300 if heavy else 6000, flags=flags)
                img, stats = host.render_host(nn, W, H, matcap)
                assert np.array_equal(img, ref["image"]), (name, flags)
                assert stats["queries"] == ref["queries"], (name, flags)
        for scene, ops in ((host.SCENE_TANH, [(oracle.PRIM_TANH, oracle.COMB_SET, [])]),
                           (host.SCENE_MANYSPHERE, [(oracle.PRIM_NSDF, oracle.COMB_SET, []), (oracle.MOD_MANY_SPHERE, 0, [1.0])])):
            host.setSceneOptions(scene)
            a, sa = host.render_host(nn, W, H, matcap)
            host.setSceneProgram(ops)
            b, sb = host.render_host(nn, W, H, matcap)
            assert np.array_equal(a, b) and sa["queries"] == sb["queries"]
    finally:
        host.setSceneOptions()
        oracle.set_scene_program([])


def test_tile_priority_order_is_invisible(env):
    """tile_priority.cuh: probing and re-ordering the tiles (forced on for this small frame through nsdf_tuning, also for a
    3-rank tile partition) must not change a byte of an exact frame or a query of its count."""
    torch, host, oracle, capi = env["torch"], env["host"], env["oracle"], env["capi"]
    try:
        for steps in (4, 32):
            host.setTuning(tile_priority_min_jobs=1, tile_priority_steps=steps)
            dev = (capi.RENDER_TWO_PASS, capi.RENDER_HYBRID, capi.RENDER_NO_STRAGGLER) if host.dev_schedules_available() else ()
            for flags in (0, capi.RENDER_STAGED_STORES) + dev:
                img, stats, ref = _render_both(env, "plane_1", 300, 200, "blue.png", flags=flags, ry=40.0)
                assert np.array_equal(img, ref["image"]), (steps, flags)
                assert stats["queries"] == ref["queries"]
        nn, om = _load(env, "car_1")
        iv, nm = oracle.view_matrices(0, 30)
        host.copyStaticSettings(host.COLOR_FACING, 3)
        host.copyViewMatrices(iv, nm, 0, 3)
        host.setSceneOptions(host.SCENE_TANH, 6000, 0, 0)
        W, H = 200, 136
        c = torch.full((W * H,), 0x55, dtype=torch.int32, device="cuda")
        for r in range(3):
            host.render_kernel(c, W, H, nn, tile_rank=r, tile_world=3)
        torch.cuda.synchronize()
        ref = oracle.render(om, iv, nm, W, H)
        assert np.array_equal(c.cpu().numpy().view(np.uint32).reshape(H, W), ref["image"])
    finally:
        host.setTuning()


def test_frame_rotated_camera_and_head_scene(env):
    img, stats, ref = _render_both(env, "car_1", 100, 80, "blue.png", rx=20.0, ry=135.0)
    assert np.array_equal(img, ref["image"])
    img, stats, ref = _render_both(env, "car_1", 80, 80, "blue.png", scene=1, frame=40)   # HEAD manySphere scene
    assert np.array_equal(img, ref["image"])
    assert stats["queries"] == ref["queries"]


def test_tc_farfield_frame(env):
    """Mixed-precision tracer (NSDF_RENDER_TC_FARFIELD, csrc/trace_tc.cuh): far-field marching on the tensor cores
    (fp16x3 rung, |SDF error| <= 1e-4), fp32 near the surface.  Not bit-identical by design; the north star asks for
    identical hit masks except at grazing silhouettes (mismatch fraction measured here) and reports RGB parity."""
    torch, host, capi = env["torch"], env["host"], env["capi"]
    if not host.dev_schedules_available():
        pytest.skip("the mixed-precision far field is out of the RGB tolerance and lives in the development build only")
    for name, W, H in (("car_1", 256, 256), ("plane_1", 480, 270)):
        nn, _ = _load(env, name)
        mc = host.Image()
        assert mc.loadPNG(os.path.join(MATCAPS, "blue.png"))
        host.copyStaticSettings(host.COLOR_MATCAP, 3)
        host.updateViewMatrices((0.0, 20.0, 0.0), (0.0, 0.0, -2.0), 0)
        frames, stats = [], []
        for flags in (0, capi.RENDER_TC_FARFIELD):
            host.setSceneOptions(host.SCENE_TANH, 6000, flags, 0)
            fr = torch.full((W * H,), 0x55555555, dtype=torch.int32, device="cuda")
            st = torch.zeros(6, dtype=torch.int64, device="cuda")
            host.render_kernel(fr, W, H, nn, mc, stats=st)
            torch.cuda.synchronize()
            frames.append(fr.cpu().numpy().view(np.uint32))
            stats.append(st.cpu().numpy())
        host.setSceneOptions(host.SCENE_TANH, 6000, 0, 0)
        ref, img = frames
        assert not np.any(img == 0x55555555), "a pixel was never written"
        hit0, hit1 = ref != 0, img != 0
        assert (hit0 != hit1).sum() <= max(2, 1e-3 * hit0.sum()), (name, int((hit0 != hit1).sum()))
        both = hit0 & hit1
        d = np.max(np.stack([np.abs(((ref >> s) & 255).astype(np.int32) - ((img >> s) & 255).astype(np.int32)) for s in (0, 8, 16)]), axis=0)
        assert (d[both] <= 1).mean() >= 0.85, (name, float((d[both] <= 1).mean()))
        assert stats[1][5] == W * H and stats[1][3] >= hit1.sum()      # every pixel covered; every hit went through the fp32 path
        # the scene is unsupported in this mode rather than silently different
    host.setSceneOptions(host.SCENE_MANYSPHERE, 6000, capi.RENDER_TC_FARFIELD, 0)
    fr = torch.zeros(64 * 64, dtype=torch.int32, device="cuda")
    nn, _ = _load(env, "car_1")
    with pytest.raises(capi.NsdfError):
        host.render_kernel(fr, 64, 64, nn, None)
    host.setSceneOptions(host.SCENE_TANH, 6000, 0, 0)


def test_frame_step_cap(env):
    for cap in (1, 7, 64):
        img, stats, ref = _render_both(env, "plane_1", 96, 64, None, max_steps=cap)
        assert np.array_equal(img, ref["image"]), cap
        assert stats["queries"] == ref["queries"]


def test_frame_determinism_and_tiles(env):
    """Same frame twice -> identical bytes; rendering the 64x64 tiles of ranks 0..G-1 into one
    buffer equals the single-rank frame (the multi-GPU partition, on one device)."""
    torch, host, oracle = env["torch"], env["host"], env["oracle"]
    nn, om = _load(env, "car_1")
    W, H = 200, 136
    iv, nm = oracle.view_matrices(0, 30)
    host.copyStaticSettings(host.COLOR_FACING, 3)
    host.copyViewMatrices(iv, nm, 0, 3)
    a = torch.zeros(W * H, dtype=torch.int32, device="cuda")
    b = torch.zeros_like(a)
    host.render_kernel(a, W, H, nn)
    host.render_kernel(b, W, H, nn)
    assert torch.equal(a, b)
    for G in (2, 3, 8):
        c = torch.full_like(a, 0x55)
        st = torch.zeros(6, dtype=torch.int64, device="cuda")
        for r in range(G):
            host.render_kernel(c, W, H, nn, stats=st, tile_rank=r, tile_world=G)
        assert torch.equal(a, c), G
        assert int(st[5]) == W * H
    ref = oracle.render(om, iv, nm, W, H)
    assert np.array_equal(a.cpu().numpy().view(np.uint32).reshape(H, W), ref["image"])


def test_full_size_frame_properties(env):
    """BASELINE config 2 size (1920x1080 plane_1, the configuration the numbers are quoted on, where tile priority and the
    express lanes are fully engaged): determinism across schedules, query accounting, and -- in both scenes -- rows 508..571
    of the frame compared with the oracle pixel for pixel."""
    torch, host, oracle = env["torch"], env["host"], env["oracle"]
    nn, om = _load(env, "plane_1")
    W, H = 1920, 1080
    matcap_np = oracle.load_matcap_png(os.path.join(MATCAPS, "blue.png"))
    mc = host.Image(host.Shape(matcap_np.shape[1], matcap_np.shape[0]))
    mc.allocateMemory()
    mc.hostData[:] = matcap_np.ravel()
    mc.copyHostToDevice()
    iv, nm = oracle.view_matrices()
    host.copyStaticSettings(host.COLOR_MATCAP, 3)
    host.copyViewMatrices(iv, nm, 0, 3)
    outs = []
    for flags in (0, host.capi.RENDER_SIMPLE):
        host.setSceneOptions(host.SCENE_TANH, 6000, flags)
        out = torch.zeros(W * H, dtype=torch.int32, device="cuda")
        st = torch.zeros(6, dtype=torch.int64, device="cuda")
        host.render_kernel(out, W, H, nn, mc, stats=st)
        torch.cuda.synchronize()
        outs.append((out, st.cpu().numpy()))
    host.setSceneOptions()
    assert torch.equal(outs[0][0], outs[1][0])
    assert outs[0][1][0] == outs[1][1][0]
    q, hits, rays = int(outs[0][1][0]), int(outs[0][1][1]), int(outs[0][1][2])
    assert rays == W * H and abs(q - 25034278) / 25034278 < 5e-3 and abs(hits - 70375) < 300
    img = outs[0][0].cpu().numpy().view(np.uint32).reshape(H, W)
    assert int((img != 0).sum()) == hits
    band = (508, 572, 1)
    ref = oracle.render(om, iv, nm, W, H, matcap=matcap_np, rows=band)
    assert np.array_equal(img[508:572], ref["image"][508:572])
    assert int((ref["image"][508:572] != 0).sum()) > 1000          # the band crosses the shape
    # the scene the reference renders at HEAD, same band
    host.setSceneOptions(host.SCENE_MANYSPHERE, 6000)
    out = torch.zeros(W * H, dtype=torch.int32, device="cuda")
    host.render_kernel(out, W, H, nn, mc)
    host.setSceneOptions()
    ref = oracle.render(om, iv, nm, W, H, matcap=matcap_np, rows=band, scene=oracle.SCENE_MANYSPHERE)
    assert np.array_equal(out.cpu().numpy().view(np.uint32).reshape(H, W)[508:572], ref["image"][508:572])


def test_full_size_4k_band_against_oracle(env):
    """BASELINE config 4 size (3840x2160, 3a3d4a90....obj): one 64-row band of the frame and of its 3-way tile partition
    against the oracle, pixel for pixel."""
    torch, host, oracle = env["torch"], env["host"], env["oracle"]
    nn, om = _load(env, "3a3d4a90a2db90b4203936772104a82d.obj")
    W, H = 3840, 2160
    matcap_np = oracle.load_matcap_png(os.path.join(MATCAPS, "blue.png"))
    mc = host.Image(host.Shape(matcap_np.shape[1], matcap_np.shape[0]))
    mc.allocateMemory()
    mc.hostData[:] = matcap_np.ravel()
    mc.copyHostToDevice()
    iv, nm = oracle.view_matrices()
    host.copyStaticSettings(host.COLOR_MATCAP, 3)
    host.copyViewMatrices(iv, nm, 0, 3)
    host.setSceneOptions(host.SCENE_TANH, 6000)
    a = torch.zeros(W * H, dtype=torch.int32, device="cuda")
    host.render_kernel(a, W, H, nn, mc)
    c = torch.full_like(a, 0x55)
    for r in range(3):
        host.render_kernel(c, W, H, nn, mc, tile_rank=r, tile_world=3)
    assert torch.equal(a, c)
    band = (1024, 1088, 1)
    ref = oracle.render(om, iv, nm, W, H, matcap=matcap_np, rows=band)
    img = a.cpu().numpy().view(np.uint32).reshape(H, W)
    assert np.array_equal(img[1024:1088], ref["image"][1024:1088])
    assert int((ref["image"][1024:1088] != 0).sum()) > 1000


def test_batched_views_and_staged_stores(env):
    """nsdf_render_views (n views in ONE persistent launch) == n calls of nsdf_render, bytewise and in the query count, in both
    scenes; the staged 16-byte tile stores (NSDF_RENDER_STAGED_STORES, always on for tile partitions) == the direct stores, for a
    frame whose width is not a multiple of 4 as well; per-view frame numbers animate the manySphere scene per view."""
    torch, host, oracle = env["torch"], env["host"], env["oracle"]
    nn, om = _load(env, "plane_1")
    W, H = 200, 136
    views = [(0.0, 0.0), (10.0, 45.0), (-20.0, 200.0), (0.0, 300.0), (5.0, 123.0)]
    frames_no = [0, 3, 40, 0, 7]
    host.copyStaticSettings(host.COLOR_FACING, 3)
    for scene in (host.SCENE_TANH, host.SCENE_MANYSPHERE):
        host.setSceneOptions(scene, 6000)
        ref = torch.zeros(len(views) * W * H, dtype=torch.int32, device="cuda")
        st1 = torch.zeros(6, dtype=torch.int64, device="cuda")
        for v, (rx, ry) in enumerate(views):
            host.updateViewMatrices((rx, ry, 0.0), (0.0, 0.0, -2.0), frames_no[v])
            host.render_kernel(ref[v * W * H:(v + 1) * W * H], W, H, nn, stats=st1)
        out = torch.full_like(ref, 0x55)
        st2 = torch.zeros(6, dtype=torch.int64, device="cuda")
        host.render_views(out, W, H, nn, host.view_cameras(views, frames_no), stats=st2)
        assert torch.equal(out, ref), scene
        keep = [0, 1, 2, 4, 5]     # queries, hit pixels, rays, unfinished, pixels ("evicted" counts express hand-overs: schedule-dependent)
        assert torch.equal(st1[keep], st2[keep])
        # views of a batch, tile-partitioned three ways (staged stores)
        out3 = torch.full_like(ref, 0x55)
        for r in range(3):
            host.render_views(out3, W, H, nn, host.view_cameras(views, frames_no), tile_rank=r, tile_world=3)
        assert torch.equal(out3, ref), scene
        # one view against the oracle
        iv, nm = oracle.view_matrices(10.0, 45.0)
        o = oracle.render(om, iv, nm, W, H, scene=oracle.SCENE_TANH if scene == host.SCENE_TANH else oracle.SCENE_MANYSPHERE, frame=3)
        assert np.array_equal(out[W * H:2 * W * H].cpu().numpy().view(np.uint32).reshape(H, W), o["image"])
    for (W2, H2) in ((200, 136), (203, 77), (64, 64)):
        host.setSceneOptions(host.SCENE_MANYSPHERE, 6000)
        host.updateViewMatrices((0.0, 30.0, 0.0), (0.0, 0.0, -2.0), 0)
        a = torch.zeros(W2 * H2, dtype=torch.int32, device="cuda")
        host.render_kernel(a, W2, H2, nn)
        host.setSceneOptions(host.SCENE_MANYSPHERE, 6000, host.capi.RENDER_STAGED_STORES)
        b = torch.full_like(a, 0x55)
        host.render_kernel(b, W2, H2, nn)
        assert torch.equal(a, b), (W2, H2)
    host.setSceneOptions()


def test_throughput_tracer_for_many_views(env):
    """Calls of >= 4 views and >= 12 M pixels run on trace_bulk4_kernel (four rays per lane, ring queue to cooperative consumers):
    frames and query counts equal the cooperative tracer's (nsdf_tuning.bulk4 = 0) and one nsdf_render per view, in both scenes
    and both math modes, with matcap shading, with staged stores, and with the ring queue and the consumer SMs forced into
    heavy use (rays leave after 24 steps; every 4th SM consumes from the start)."""
    torch, host, oracle = env["torch"], env["host"], env["oracle"]
    nn, om = _load(env, "plane_1")
    matcap = host.Image()
    assert matcap.loadPNG(os.path.join(MATCAPS, "blue.png"))
    matcap_np = oracle.load_matcap_png(os.path.join(MATCAPS, "blue.png"))
    W, H = 512, 384
    views = [(0.0, 0.0), (10.0, 45.0), (-20.0, 200.0), (0.0, 300.0), (5.0, 123.0)]
    frames_no = [0, 3, 40, 0, 7]
    keep = [0, 1, 2, 4, 5]
    host.copyStaticSettings(host.COLOR_MATCAP, 3)
    for scene, math in ((host.SCENE_MANYSPHERE, host.MATH_REFERENCE), (host.SCENE_TANH, host.MATH_DETERMINISTIC),
                        (host.SCENE_MANYSPHERE, host.MATH_DETERMINISTIC)):
        host.setSceneOptions(scene, 6000, 0, 0, math)
        ref = torch.zeros(len(views) * W * H, dtype=torch.int32, device="cuda")
        st1 = torch.zeros(6, dtype=torch.int64, device="cuda")
        for v, (rx, ry) in enumerate(views):
            host.updateViewMatrices((rx, ry, 0.0), (0.0, 0.0, -2.0), frames_no[v])
            host.render_kernel(ref[v * W * H:(v + 1) * W * H], W, H, nn, matcap, stats=st1)
        cams = host.view_cameras(views, frames_no)
        for knobs in (dict(bulk4=0), dict(bulk4=1), dict(bulk4=1, bulk4_evict_steps=24, bulk4_period=4, bulk4_drain=64),
                      dict(bulk4=1, bulk4_unroll=4, bulk4_refill_min=1), dict(bulk4=1, bulk4_unroll=1, bulk4_drain=128)):
            host.setTuning(**knobs)
            out = torch.full_like(ref, 0x55)
            st2 = torch.zeros(6, dtype=torch.int64, device="cuda")
            host.render_views(out, W, H, nn, cams, matcap, stats=st2)
            host.setTuning()
            assert torch.equal(out, ref), (scene, math, knobs)
            assert torch.equal(st1[keep], st2[keep]), (scene, math, knobs)
            if knobs.get("bulk4_evict_steps") and not host.dev_schedules_available():
                assert int(st2[3]) > 1000      # the ring queue really carried rays
        # staged stores on top of the throughput tracer (forced: the automatic choice wants >= 12 M pixels per call)
        host.setSceneOptions(scene, 6000, host.capi.RENDER_STAGED_STORES, 0, math)
        host.setTuning(bulk4=1)
        out = torch.full_like(ref, 0x55)
        host.render_views(out, W, H, nn, cams, matcap)
        host.setTuning()
        assert torch.equal(out, ref), (scene, math, "staged")
        if math == host.MATH_DETERMINISTIC:     # one view of the batch against the oracle
            iv, nm = oracle.view_matrices(10.0, 45.0)
            o = oracle.render(om, iv, nm, W, H, matcap=matcap_np, frame=3,
                              scene=oracle.SCENE_TANH if scene == host.SCENE_TANH else oracle.SCENE_MANYSPHERE)
            assert np.array_equal(out[W * H:2 * W * H].cpu().numpy().view(np.uint32).reshape(H, W), o["image"])
    host.setSceneOptions()
    host.copyStaticSettings(host.COLOR_FACING, 3)


def test_calls_side_by_side_with_max_ctas(env):
    """nsdf_tuning.max_ctas caps a call's persistent grid; calls of different model handles on different streams then run side by side
    (both tracers are cooperative launches of their own grid) and write the bytes the calls write alone."""
    torch, host = env["torch"], env["host"]
    nets = [_load(env, n)[0] for n in ("plane_1", "car_1", "plane_2")]
    W, H = 384, 320
    host.copyStaticSettings(host.COLOR_FACING, 3)
    host.setSceneOptions(host.SCENE_MANYSPHERE, 6000, 0, 0, host.MATH_REFERENCE)
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    for views, knobs in (([(0.0, 0.0), (0.0, 90.0)], {}), ([(0.0, 10.0 * k) for k in range(8)], dict(bulk4=1))):
        cams = host.view_cameras(views)
        n = len(views) * W * H
        alone = [torch.zeros(n, dtype=torch.int32, device="cuda") for _ in nets]
        host.setTuning(**knobs) if knobs else host.setTuning()
        for nn, o in zip(nets, alone):
            host.render_views(o, W, H, nn, cams)
        torch.cuda.synchronize()
        outs = [torch.full((n,), 0x55, dtype=torch.int32, device="cuda") for _ in nets]
        streams = [torch.cuda.Stream() for _ in nets]
        host.setTuning(max_ctas=sms // len(nets), **knobs)
        for _ in range(3):
            for nn, o, st in zip(nets, outs, streams):
                with torch.cuda.stream(st):
                    host.render_views(o, W, H, nn, cams)
        torch.cuda.synchronize()
        host.setTuning()
        for o, a in zip(outs, alone):
            assert torch.equal(o, a), (len(views), knobs)
    host.setSceneOptions()


def test_error_behaviour(env):
    torch, host = env["torch"], env["host"]
    from cudaneuralrender_b200 import _capi
    L = _capi.lib()
    h = C.c_void_p(None)
    assert L.nsdf_model_load_h5(b"/nonexistent/model.h5", C.byref(h)) == 10003
    with pytest.raises(FileNotFoundError):
        host.NeuralNetwork().load("/nonexistent/model.h5")
    nn, _ = _load(env, "car_1")
    assert L.nsdf_mlp_forward(nn.handle(), None, None, 5, 0, None) == 10001
    assert L.nsdf_mlp_forward(nn.handle(), None, None, 0, 0, None) == 0
    x = torch.zeros(30, device="cuda")
    y = torch.zeros(10, device="cuda")
    assert L.nsdf_mlp_forward(nn.handle(), x.data_ptr(), y.data_ptr(), 10, 99, None) == 10001
    assert L.nsdf_dense_layer_forward(x.data_ptr(), x.data_ptr(), y.data_ptr(), y.data_ptr(), 1, 2, 3, 1, 0, None) == 10002


def test_render_beside_a_busy_stream(env):
    """The tracer's warps wait on one another (queues, termination counters), so it is launched cooperatively: with another
    kernel occupying SMs on a second stream the launch waits for room instead of spinning on CTAs that cannot become resident.
    Frames rendered beside a stream of long-running kernels equal the frame rendered alone."""
    torch, host = env["torch"], env["host"]
    nn, _ = _load(env, "plane_1")
    W, H = 640, 360
    host.copyStaticSettings(host.COLOR_FACING, 3)
    host.updateViewMatrices((0.0, 30.0, 0.0), (0.0, 0.0, -2.0), 0)
    host.setSceneOptions(host.SCENE_MANYSPHERE, 6000)
    alone = torch.zeros(W * H, dtype=torch.int32, device="cuda")
    host.render_kernel(alone, W, H, nn)
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    a = torch.randn(4096, 4096, device="cuda")
    outs = []
    with torch.cuda.stream(side):
        for _ in range(40):               # ~100 ms of GEMMs that fill the machine
            a = (a @ a) * 1e-3
    for _ in range(6):                    # renders on the default stream while the side stream is busy
        o = torch.full((W * H,), 0x55, dtype=torch.int32, device="cuda")
        host.render_kernel(o, W, H, nn)
        outs.append(o)
    torch.cuda.synchronize()
    for o in outs:
        assert torch.equal(o, alone)
    # the throughput tracer (trace_bulk4_kernel) is a cooperative launch as well: a batch of views beside the busy stream
    views = [(0.0, 30.0), (0.0, 75.0), (10.0, 200.0), (0.0, 300.0)]
    cams = host.view_cameras(views)
    host.setTuning(bulk4=1)
    ref4 = torch.zeros(len(views) * W * H, dtype=torch.int32, device="cuda")
    host.render_views(ref4, W, H, nn, cams)
    torch.cuda.synchronize()
    assert torch.equal(ref4[:W * H], alone)
    outs = []
    with torch.cuda.stream(side):
        for _ in range(40):
            a = (a @ a) * 1e-3
    for _ in range(4):
        o = torch.full_like(ref4, 0x55)
        host.render_views(o, W, H, nn, cams)
        outs.append(o)
    torch.cuda.synchronize()
    host.setTuning()
    for o in outs:
        assert torch.equal(o, ref4)
    host.setSceneOptions()

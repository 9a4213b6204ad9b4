"""CPU tests (no GPU): the oracle against the golden vectors, the HDF5 readers against each other,
image I/O, and that libnsdf.so loads and exports every symbol include/nsdf.h declares."""
import ctypes as C
import json
import os
import re
import struct

import numpy as np
import pytest

from conftest import GOLDEN, MATCAPS, MODELS, ROOT, model_path

from oracle import h5mini, oracle


@pytest.fixture(scope="module")
def kat():
    with open(os.path.join(GOLDEN, "kat_sdf.json")) as f:
        return json.load(f)["rows"]


def test_oracle_known_answers(built, kat):
    """25 SDF values with bit patterns (CUTLASS SIMT order)."""
    models = {}
    for row in kat:
        m = models.setdefault(row["model"], oracle.Model.from_h5(model_path(row["model"])))
        y = oracle.mlp_forward(m, np.asarray([row["point"]], dtype=np.float32))[0, 0]
        bits = struct.unpack("<I", struct.pack("<f", float(y)))[0]
        assert bits == int(row["bits"], 16), (row, float(y))
        t = oracle.tanhf(np.asarray([y], dtype=np.float32))[0]
        assert abs(float(t) - row["tanh"]) <= 2e-7 * max(1.0, abs(row["tanh"])) + 6e-8


def test_oracle_matches_float64_numpy(built):
    """fp32 ascending-k evaluation stays within 1e-6 of an fp64 evaluation."""
    rng = np.random.default_rng(0)
    x = rng.uniform(-1, 1, (20000, 3)).astype(np.float32)
    for name in MODELS[:2]:
        m = oracle.Model.from_h5(model_path(name))
        a = x.astype(np.float64)
        layers = m.layers()
        for i, (W, b) in enumerate(layers):
            a = a @ W.astype(np.float64).T + b.astype(np.float64)
            if i < len(layers) - 1:
                a = np.maximum(a, 0)
        y = oracle.mlp_forward(m, x)
        assert np.max(np.abs(y - a)) < 1e-6


def test_oracle_equal_inputs_equal_outputs(built):
    """The only check the reference ships (src/simpleInfer.cpp:134-146)."""
    m = oracle.Model.from_h5(model_path("car_1"))
    y = oracle.mlp_forward(m, np.zeros((10000, 3), dtype=np.float32))
    assert np.all(y == y[0])
    assert y[0, 0].view(np.uint32) == 0xBDF2F39C


def test_oracle_threads_agree(built):
    m = oracle.Model.from_h5(model_path("plane_1"))
    x = np.random.default_rng(1).uniform(-1, 1, (5003, 3)).astype(np.float32)
    assert np.array_equal(oracle.mlp_forward(m, x, nthreads=1), oracle.mlp_forward(m, x, nthreads=4))


def test_oracle_tanh_accuracy(built):
    x = np.concatenate([np.linspace(-12, 12, 200001), np.random.default_rng(0).normal(0, 0.3, 100000)]).astype(np.float32)
    t = oracle.tanhf(x)
    ref = np.tanh(x.astype(np.float64))
    ulp = np.spacing(np.abs(ref).astype(np.float32)).astype(np.float64)
    assert np.max(np.abs(t - ref) / ulp) < 1.6
    assert np.all(oracle.tanhf(np.float32([0.0, -0.0])) == 0)
    assert oracle.tanhf(np.float32([np.inf]))[0] == 1 and oracle.tanhf(np.float32([-np.inf]))[0] == -1
    assert np.isnan(oracle.tanhf(np.float32([np.nan]))[0])
    small = np.float32([1e-6, -3e-7, 1e-4])
    assert np.array_equal(oracle.tanhf(small), small)   # tanh(x) == x near the surface


def test_h5_layout_golden(built):
    """raw-data offsets and dims, identical in all five files; both readers agree."""
    with open(os.path.join(GOLDEN, "h5_layout.json")) as f:
        gold = json.load(f)
    for name in MODELS:
        path = model_path(name)
        assert os.path.getsize(path) == gold["file_bytes"]
        f5 = h5mini.H5File(path)
        names = f5.root.list_object_names()
        assert names == [l["layer"] for l in gold["layers"]]
        raw = open(path, "rb").read()
        for l in gold["layers"]:
            inner = f5.root.children[l["layer"]].children[l["layer"]]
            k, b = inner.children["kernel:0"], inner.children["bias:0"]
            assert list(k.dims) == l["kernel_dims"] and k.data_offset == l["kernel_offset"]
            assert list(b.dims) == l["bias_dims"] and b.data_offset == l["bias_offset"]
            assert np.array_equal(k.data.ravel(), np.frombuffer(raw, "<f4", k.data.size, l["kernel_offset"]))


def _capi():
    from cudaneuralrender_b200 import _capi
    return _capi


def test_library_exports_every_declared_symbol(built):
    capi = _capi()
    L = capi.lib()
    hdr = open(os.path.join(ROOT, "include", "nsdf.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(nsdf_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 25
    for name in declared:
        assert hasattr(L, name), f"libnsdf.so does not export {name}"
    assert declared == set(capi.PROTOTYPES), declared ^ set(capi.PROTOTYPES)
    assert b"sm_100a" in L.nsdf_version()
    assert L.nsdf_error_string(10002).startswith(b"nsdf:")


def test_ctypes_structs_mirror_the_header(built):
    """The ctypes mirrors of nsdf_tuning / nsdf_render_opts / nsdf_render_stats (cudaneuralrender_b200/_capi.py) list the fields of
    include/nsdf.h in the same order (all int-sized or pointer members: order and count fix the layout)."""
    capi = _capi()
    hdr = open(os.path.join(ROOT, "include", "nsdf.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)

    def fields(name):
        body = re.search(r"typedef struct " + name + r"\s*\{(.*?)\}\s*" + name + r"\s*;", hdr, flags=re.S).group(1)
        out = []
        for decl in body.split(";"):
            decl = decl.strip()
            if decl:
                out.append(re.findall(r"([A-Za-z_][A-Za-z0-9_]*)\s*(?:\[[^\]]*\])?$", decl)[0])
        return out

    assert fields("nsdf_tuning") == [n for n, _ in capi.Tuning._fields_]
    assert fields("nsdf_render_opts") == [n for n, _ in capi.RenderOpts._fields_]
    assert fields("nsdf_render_stats") == [n for n, _ in capi.RenderStats._fields_]
    t = capi.Tuning()
    capi.lib().nsdf_tuning_default(C.byref(t))
    assert t.max_ctas == 0 and t.bulk4 == -1 and t.express_steps == 160      # first, a middle and the last member land where C wrote them


def test_product_h5_reader_matches_independent_reader(built):
    """libnsdf's C++ reader (include/nsdf_h5.hpp) vs oracle/h5mini.py, host only."""
    L = _capi().lib()
    for name in MODELS:
        n, cnt = C.c_int(0), C.c_size_t(0)
        dims = (C.c_int * 16)()
        assert L.nsdf_h5_read_keras(model_path(name).encode(), C.byref(n), dims, 16, None, 0, C.byref(cnt)) == 0
        assert n.value == 9 and list(dims[:10]) == [3] + [32] * 8 + [1] and cnt.value == 7553
        params = np.empty(cnt.value, dtype=np.float32)
        assert L.nsdf_h5_read_keras(model_path(name).encode(), None, None, 0,
                                    params.ctypes.data_as(C.POINTER(C.c_float)), params.size, None) == 0
        ref_params, ref_dims = h5mini.pack_model(h5mini.load_keras_mlp(model_path(name)))
        assert np.array_equal(params, ref_params) and list(ref_dims) == list(dims[:10])


def test_product_h5_reader_rejects_garbage(built, tmp_path):
    L = _capi().lib()
    p = tmp_path / "bad.h5"
    p.write_bytes(b"not an hdf5 file" * 100)
    assert L.nsdf_h5_read_keras(str(p).encode(), None, None, 0, None, 0, None) == 10003
    assert L.nsdf_h5_read_keras(str(tmp_path / "missing.h5").encode(), None, None, 0, None, 0, None) == 10003
    raw = bytearray(open(model_path("car_1"), "rb").read())
    q = tmp_path / "trunc.h5"
    q.write_bytes(bytes(raw[:20000]))   # cuts into the kernels
    assert L.nsdf_h5_read_keras(str(q).encode(), None, None, 0, None, 0, None) == 10003


def _random_layers(dims, seed):
    rng = np.random.default_rng(seed)
    return [(rng.standard_normal((dims[i], dims[i + 1])).astype(np.float32), rng.standard_normal(dims[i + 1]).astype(np.float32))
            for i in range(len(dims) - 1)]


def test_h5_writer_round_trips_shipped_files(built, tmp_path):
    """nsdf_h5_write_keras: a shipped file read, written and read again gives the same floats through the
    product reader AND through the independent Python parser."""
    from cudaneuralrender_b200 import host
    for name in MODELS:
        layers = host.read_keras_h5(model_path(name))
        out = str(tmp_path / (name + ".h5"))
        host.write_keras_h5(out, layers)
        back = host.read_keras_h5(out)
        indep = h5mini.load_keras_mlp(out)
        assert len(back) == len(layers) == len(indep) == 9
        for (k0, b0), (k1, b1), (k2, b2) in zip(layers, back, indep):
            assert np.array_equal(k0, k1) and np.array_equal(b0, b1)
            assert np.array_equal(k0, np.asarray(k2).reshape(k0.shape)) and np.array_equal(b0, np.asarray(b2).ravel())


def test_h5_layer_names_order_beyond_ten_layers(built, tmp_path):
    """Thirteen layers: the name sort of the reference (src/neuralNetwork.cpp:88-91) yields dense, dense_1, dense_10, ...;
    the product reader follows the root `layer_names` attribute and returns network order.  Other widths and a 4-input
    first layer go through the same files."""
    from cudaneuralrender_b200 import host
    dims = [4, 16, 64, 32, 128, 32, 32, 8, 32, 32, 32, 32, 32, 1]
    layers = _random_layers(dims, 7)
    out = str(tmp_path / "deep.h5")
    host.write_keras_h5(out, layers)
    back = host.read_keras_h5(out)
    assert [k.shape for k, _ in back] == [(dims[i], dims[i + 1]) for i in range(13)]
    for (k0, b0), (k1, b1) in zip(layers, back):
        assert np.array_equal(k0, k1) and np.array_equal(b0, b1)
    names = ["dense"] + [f"dense_{i}" for i in range(1, 13)]
    indep = h5mini.load_keras_mlp(out)                  # name-sorted, like the reference
    for nm, (k2, b2) in zip(sorted(names), indep):
        k0, b0 = layers[names.index(nm)]
        assert np.array_equal(k0, np.asarray(k2).reshape(k0.shape)) and np.array_equal(b0, np.asarray(b2).ravel())


def test_oracle_animation_net_frame0_equals_static_net(built, tmp_path):
    """Oracle, 4-input (x, y, z, t) nets (src/volumeRender_kernel.cu:524-545): with t = frame = 0 the fourth product of the
    first layer is exactly 0, so the frame equals the 3-input shape's; a later frame differs; the dense forward agrees with
    a float64-free numpy restatement of the same ascending-k chain on a few points."""
    from cudaneuralrender_b200 import host
    layers = host.read_keras_h5(model_path("car_1"))
    rng = np.random.default_rng(11)
    k0, b0 = layers[0]
    layers[0] = (np.concatenate([k0, (0.02 * rng.standard_normal((1, k0.shape[1]))).astype(np.float32)], axis=0), b0)
    path = str(tmp_path / "car_xyzt.h5")
    host.write_keras_h5(path, layers)
    om4, om3 = oracle.Model.from_h5(path), oracle.Model.from_h5(model_path("car_1"))
    assert list(om4.dims) == [4] + [32] * 8 + [1]
    iv, nm = oracle.view_matrices(0.0, 20.0)
    a = oracle.render(om4, iv, nm, 64, 48, frame=0)
    b = oracle.render(om3, iv, nm, 64, 48, frame=0)
    c = oracle.render(om4, iv, nm, 64, 48, frame=9)
    assert np.array_equal(a["image"], b["image"]) and a["queries"] == b["queries"]
    assert not np.array_equal(c["image"], a["image"])
    x = rng.uniform(-1, 1, (5, 4)).astype(np.float32)
    x[:, 3] = 9.0
    y = oracle.mlp_forward(om4, x).ravel()
    for i in range(5):
        act = x[i]
        for li, (k, bias) in enumerate(layers):
            z = np.zeros(k.shape[1], dtype=np.float32)
            for kk in range(k.shape[0]):   # ascending k, one rounding per fused multiply-add
                z = (k[kk].astype(np.float64) * np.float64(act[kk]) + z.astype(np.float64)).astype(np.float32)
            z = z + bias
            act = np.maximum(z, 0) if li < len(layers) - 1 else z
        assert act[0] == y[i]


def test_oracle_sinf_accuracy(built):
    """sinf_det (the deterministic sin both sides use for sdfOpDisplace, src/volumeRender_kernel.cu:103-110)."""
    xs = np.concatenate([np.linspace(-20, 20, 20001), np.linspace(-1000, 1000, 2001)]).astype(np.float32)
    got = np.array([oracle.sinf(float(x)) for x in xs], dtype=np.float64)
    assert np.max(np.abs(got - np.sin(xs.astype(np.float64)))) < 3e-7
    assert np.isnan(oracle.sinf(float("inf"))) and np.isnan(oracle.sinf(float("nan"))) and oracle.sinf(0.0) == 0.0


def test_oracle_scene_library_matches_float64_formulas(built):
    """The analytic SDF library (src/volumeRender_kernel.cu:67-174) as scene programs against float64 numpy restatements of
    the same formulas: primitives, hard and smooth booleans, modifiers, and the two composite scenes."""
    rng = np.random.default_rng(3)
    n = 4000
    p = rng.uniform(-1.2, 1.2, (n, 3)).astype(np.float32)
    ns = rng.uniform(-0.3, 0.6, n).astype(np.float32)
    P, N = p.astype(np.float64), ns.astype(np.float64)
    sat = lambda x: np.clip(x, 0.0, 1.0)
    def ssub(d1, d2, k):
        h = sat(0.5 - 0.5 * (d1 + d2) / k); return d1 * (1 - h) - d2 * h + k * h * (1 - h)
    def suni(d1, d2, k):
        h = sat(0.5 + 0.5 * (d2 - d1) / k); return d2 * (1 - h) + d1 * h - k * h * (1 - h)
    def box(q, b, r):
        d = np.abs(q) - b; return np.linalg.norm(np.maximum(d, 0), axis=1) + np.minimum(d.max(axis=1), 0) - r
    def cyl(q, c):
        return np.hypot(q[:, 0] - c[0], q[:, 1] - c[2]) - c[1]
    O = oracle
    cases = []
    c0 = np.array([0.2, -0.1, 0.3])
    cases.append(([(O.PRIM_SPHERE, O.COMB_SET, [0.2, -0.1, 0.3, 0.9])], np.linalg.norm(P - c0, axis=1) - 0.9))
    cases.append(([(O.PRIM_BOX, O.COMB_SET, [0.2, -0.1, 0.3, 0.4, 0.3, 0.2, 0.05])], box(P - c0, np.array([0.4, 0.3, 0.2]), 0.05)))
    cases.append(([(O.PRIM_PLANE, O.COMB_SET, [0.5])], P[:, 1] - 0.5))
    cases.append(([(O.PRIM_CYLINDER, O.COMB_SET, [-0.5, -0.2, 0.4, 0.1, 0.1, 0.1])],
                  cyl(P - np.array([-0.5, -0.2, 0.4]), [0.1, 0.1, 0.1])))
    cases.append(([(O.PRIM_TANH, O.COMB_SET, []), (O.MOD_ROUND, 0, [0.04])], np.tanh(N) - 0.04))
    cases.append(([(O.PRIM_TANH, O.COMB_SET, []), (O.MOD_ONION, 0, [0.02])], np.abs(np.tanh(N)) - 0.02))
    cases.append(([(O.PRIM_TANH, O.COMB_SET, []), (O.MOD_DISPLACE, 0, [5.0, 0.05])],
                  np.tanh(N) + np.sin(5 * P[:, 0]) * np.sin(5 * P[:, 1]) * np.sin(5 * P[:, 2]) * np.float64(np.float32(0.05))))
    sph = np.linalg.norm(P, axis=1) - 0.5
    for comb, f in ((O.COMB_UNION, lambda s, d: np.minimum(s, d)), (O.COMB_INTERSECT, lambda s, d: np.maximum(s, d)),
                    (O.COMB_SUBTRACT, lambda s, d: np.maximum(s, -d)), (O.COMB_SUBTRACT_FROM, lambda s, d: np.maximum(d, -s)),
                    (O.COMB_SMOOTH_UNION, lambda s, d: suni(s, d, 0.05)), (O.COMB_SMOOTH_SUBTRACT, lambda s, d: ssub(s, d, 0.05)),
                    (O.COMB_SMOOTH_SUBTRACT_FROM, lambda s, d: ssub(d, s, 0.05))):
        k = np.float64(np.float32(0.05))
        cases.append(([(O.PRIM_TANH, O.COMB_SET, []), (O.PRIM_SPHERE, comb, [0, 0, 0, 0.5, 0, 0, 0, 0.05])], f(np.tanh(N), sph)))
    # manyCylinderCut :156-174 in float64
    s = N.copy(); cp = P.copy(); cp[:, 1] -= 0.5
    for i in range(300):
        if i % 20 == 0:
            cp[:, 1] += 0.1; cp[:, 0] = P[:, 0] + 0.9
        s = ssub(s, cyl(cp, [0.02, 0.02, 0.02]), 0.01); cp[:, 0] -= 0.1
    cases.append(([(O.PRIM_NSDF, O.COMB_SET, []), (O.MOD_MANY_CYLINDER_CUT, 0, [])], s))
    # manySphere(p, nSDF, false) :176-196, frame 90
    s = N.copy(); cp = P.copy(); cp[:, 1] -= 0.6; cp[:, 2] += -0.7 + 90 * 2 * 0.7 / 360
    for i in range(9):
        if i % 3 == 0:
            cp[:, 1] += 0.4; cp[:, 0] = P[:, 0] + 0.5
        s = ssub(s, np.linalg.norm(cp, axis=1) - 0.1, 0.01); cp[:, 0] -= 0.4
    cases.append(([(O.PRIM_NSDF, O.COMB_SET, []), (O.MOD_MANY_SPHERE, 0, [0.0])], s))
    for ops, want in cases:
        oracle.set_scene_program(ops)
        got = oracle.scene_program_eval(p, ns, frame=90).astype(np.float64)
        assert np.max(np.abs(got - want)) < 2e-5, (ops, np.max(np.abs(got - want)))
    # the two compiled scenes are programs too
    oracle.set_scene_program([(O.PRIM_NSDF, O.COMB_SET, []), (O.MOD_MANY_SPHERE, 0, [1.0])])
    a = oracle.scene_program_eval(p[:200], ns[:200], frame=7)
    b = np.array([oracle.scene_sdf(1, p[i], ns[i], 7) for i in range(200)], dtype=np.float32)
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    oracle.set_scene_program([])


def test_h5_writer_rejects_bad_arguments(built, tmp_path):
    L = _capi().lib()
    dims = (C.c_int * 2)(3, 0)
    params = np.zeros(8, dtype=np.float32)
    pp = params.ctypes.data_as(C.POINTER(C.c_float))
    assert L.nsdf_h5_write_keras(str(tmp_path / "x.h5").encode(), pp, dims, 1) == 10001
    assert L.nsdf_h5_write_keras(None, pp, dims, 1) == 10001
    dims = (C.c_int * 2)(3, 1)
    assert L.nsdf_h5_write_keras(str(tmp_path / "no_such_dir" / "x.h5").encode(), pp, dims, 1) == 10003


def test_png_decoder_matches_pil(built):
    L = _capi().lib()
    for fn in sorted(os.listdir(MATCAPS)):
        if not fn.endswith(".png"):
            continue
        path = os.path.join(MATCAPS, fn)
        w, h = C.c_int(0), C.c_int(0)
        assert L.nsdf_png_load(path.encode(), None, 0, C.byref(w), C.byref(h)) == 0
        tex = np.empty(w.value * h.value, dtype=np.uint32)
        assert L.nsdf_png_load(path.encode(), tex.ctypes.data, tex.size, C.byref(w), C.byref(h)) == 0
        ref = oracle.load_matcap_png(path)
        assert ref.shape == (h.value, w.value)
        assert np.array_equal(tex.reshape(ref.shape), ref), fn


def test_png_ppm_writers(built, tmp_path):
    from PIL import Image as PILImage
    L = _capi().lib()
    rng = np.random.default_rng(3)
    w, h = 37, 21
    tex = rng.integers(0, 2 ** 32, size=(h, w), dtype=np.uint64).astype(np.uint32)
    p = str(tmp_path / "a.png")
    assert L.nsdf_png_save(p.encode(), tex.ctypes.data, w, h, 0) == 0
    im = np.asarray(PILImage.open(p).convert("RGBA"), dtype=np.uint32)
    back = (im[..., 3] << 24) | (im[..., 2] << 16) | (im[..., 1] << 8) | im[..., 0]
    assert np.array_equal(back, tex)
    # Image::savePNG default: the texel stream reversed (180-degree turn), image.cu:84-98
    assert L.nsdf_png_save(p.encode(), tex.ctypes.data, w, h, 1) == 0
    im = np.asarray(PILImage.open(p).convert("RGBA"), dtype=np.uint32)
    back = (im[..., 3] << 24) | (im[..., 2] << 16) | (im[..., 1] << 8) | im[..., 0]
    assert np.array_equal(back, tex[::-1, ::-1])
    # own decoder reads own encoder
    out = np.empty(w * h, dtype=np.uint32)
    ww, hh = C.c_int(0), C.c_int(0)
    assert L.nsdf_png_load(p.encode(), out.ctypes.data, out.size, C.byref(ww), C.byref(hh)) == 0
    assert np.array_equal(out.reshape(h, w), tex[::-1, ::-1])
    q = str(tmp_path / "a.ppm")
    assert L.nsdf_ppm_save(q.encode(), tex.ctypes.data, w, h) == 0
    raw = open(q, "rb").read()
    hdr = f"P6\n{w}\n{h}\n255\n".encode()   # helper_image.h:265-274
    assert raw.startswith(hdr) and len(raw) == len(hdr) + 3 * w * h
    rgb = np.frombuffer(raw[len(hdr):], dtype=np.uint8).reshape(h, w, 3)
    assert np.array_equal(rgb[..., 0], tex & 0xFF) and np.array_equal(rgb[..., 2], (tex >> 16) & 0xFF)


def test_camera_helper_matches_oracle(built):
    capi = _capi()
    L = capi.lib()
    for rx, ry, t in [(0, 0, (0, 0, -2)), (30, 45, (0, 0, -2)), (-10, 200, (0.1, -0.2, -3)), (0, 5.625, (0, 0, -2))]:
        cam = capi.Camera()
        L.nsdf_camera_from_view(rx, ry, t[0], t[1], t[2], 7, C.byref(cam))
        iv, nm = oracle.view_matrices(rx, ry, t)
        assert cam.frame == 7
        assert np.array_equal(np.array(cam.inv_view[:], dtype=np.float32), iv)   # same formulae -> same bits
        np.testing.assert_allclose(np.array(cam.normal[:], dtype=np.float32), nm, atol=2e-6)
    cam = capi.Camera()
    L.nsdf_camera_from_view(0, 0, 0, 0, -2, 0, C.byref(cam))
    assert list(cam.inv_view[:]) == [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 2]
    assert list(cam.normal[:]) == [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, -2, 0, 0, 0, 1]


def test_oracle_render_small_frame(built):
    """Renderer restatement: determinism, thread invariance, query accounting, max_steps rule."""
    m = oracle.Model.from_h5(model_path("car_1"))
    iv, nm = oracle.view_matrices()
    a = oracle.render(m, iv, nm, 64, 64, nthreads=1)
    b = oracle.render(m, iv, nm, 64, 64, nthreads=4)
    assert np.array_equal(a["image"], b["image"]) and a["queries"] == b["queries"]
    hits = int(a["hit"].sum())
    assert 200 < hits < 2000
    assert a["queries"] == int(a["steps"].sum()) + 4 * hits
    assert np.all((a["image"] != 0) <= (a["hit"] != 0))
    assert np.all((a["image"][a["hit"] != 0] >> 24) == 255)   # facing colour has alpha 255
    c = oracle.render(m, iv, nm, 64, 64, max_steps=8)
    assert int(c["steps"].max()) <= 8 and int(c["hit"].sum()) < hits
    # HEAD scene (manySphere union) shows more surface than the tanh scene
    d = oracle.render(m, iv, nm, 64, 64, scene=oracle.SCENE_MANYSPHERE)
    assert int(d["hit"].sum()) > hits


def test_oracle_workload_characterisation(built):
    """SURVEY App. C planning numbers for car_1 512x512 (independent numpy probe):
    2 416 iterations, 3 126 923 queries, 37 897 hit pixels -- allow the probe's tanh/rsqrt slack."""
    m = oracle.Model.from_h5(model_path("car_1"))
    iv, nm = oracle.view_matrices()
    r = oracle.render(m, iv, nm, 512, 512)
    assert abs(int(r["hit"].sum()) - 37897) <= 40
    assert abs(r["queries"] - 3126923) / 3126923 < 2e-3


def test_h5_chunked_and_filtered_datasets(built, tmp_path):
    """the reference reads whatever libhdf5 reads (src/neuralNetwork.cpp:85-151).  Chunked datasets (version-1
    chunk B-tree), with and without the shuffle + deflate filter pipeline, load through the product reader bit for bit; a
    deflate stream with dynamic Huffman blocks (zlib's own compressor) decodes through the same inflate the matcap PNGs use."""
    import zlib
    from cudaneuralrender_b200 import host
    layers = host.read_keras_h5(model_path("plane_3"))
    rng = np.random.default_rng(5)
    layers.insert(3, (rng.standard_normal((32, 32)).astype(np.float32), rng.standard_normal(32).astype(np.float32)))  # 10 layers
    for storage in (1, 2):
        out = str(tmp_path / f"chunked{storage}.h5")
        host.write_keras_h5(out, layers, storage)
        raw = open(out, "rb").read()
        assert b"TREE\x01\x00" in raw                     # a raw-data chunk B-tree node is in the file
        back = host.read_keras_h5(out)
        assert len(back) == len(layers)
        for (k0, b0), (k1, b1) in zip(layers, back):
            assert k0.shape == k1.shape and np.array_equal(k0.view(np.uint32), k1.view(np.uint32))
            assert np.array_equal(b0.view(np.uint32), b1.view(np.uint32))
    # the decoder behind the deflate filter (include/nsdf_inflate.hpp, shared with the PNG reader) on streams made by zlib's own
    # compressor: stored, fixed-Huffman and dynamic-Huffman blocks, shuffled float data like an h5py gzip chunk
    import ctypes as C
    from cudaneuralrender_b200 import _capi as capi
    k = layers[2][0]
    shuffled = np.ascontiguousarray(k.view(np.uint8).reshape(-1, 4).T).tobytes()
    for payload in (shuffled, open(model_path("plane_3"), "rb").read(), b"", b"a" * 70000, bytes(rng.integers(0, 256, 5000, dtype=np.uint8))):
        for level in (0, 1, 6, 9):
            z = zlib.compress(payload, level)
            out = (C.c_ubyte * max(1, len(payload)))()
            n_out = C.c_size_t(0)
            rc = capi.lib().nsdf_zlib_inflate(z, len(z), out, len(payload), C.byref(n_out))
            assert rc == 0 and n_out.value == len(payload) and bytes(out[:len(payload)]) == payload, (len(payload), level)
    assert capi.lib().nsdf_zlib_inflate(b"\x78\x9c\xff\xff", 4, out, 1, C.byref(n_out)) != 0     # garbage is an error, not a crash

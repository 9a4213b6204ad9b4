"""GPU tests of the drop-in executables (apps/): the reference's simpleInfer self-check and a --single frame
of neuralSDFRenderer, compared with the CPU oracle."""
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import MATCAPS, ROOT, model_path

pytestmark = pytest.mark.gpu


def test_simple_infer_app(gpu):
    out = subprocess.run([os.path.join(ROOT, "apps", "simpleInfer"), model_path("car_1"), "65536"], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    # 65 536 equal inputs (a batch size on which the reference's GemmBatched launch fails)
    m = re.search(r"All 65536 outputs equal: raw (\S+), tanh (\S+)", out.stdout)
    assert m, out.stdout
    assert np.float32(m.group(1)) == np.float32(-1.186287105e-01)      # SURVEY App. B, car_1 at the origin
    assert abs(float(m.group(2)) - (-1.180753484e-01)) < 1e-6
    m = re.search(r"sdf\(0.1, 0.2, 0.3\) = (\S+) \(raw (\S+)\)", out.stdout)
    assert m and np.float32(m.group(2)) == np.float32(-1.369275451e-01)


@pytest.mark.parametrize("matcap", ["blue.png", None])
def test_renderer_app_single_frame(gpu, tmp_path, matcap):
    from PIL import Image
    from oracle import oracle
    W, H = 112, 80
    prefix = str(tmp_path) + "/"
    cmd = [os.path.join(ROOT, "apps", "neuralSDFRenderer"), "-i", model_path("car_1"), "-o", prefix, "-W", str(W), "-H", str(H),
           "-ry", "30", "--single", "--ppm", "--stats", "--scene", "tanh", "--math", "deterministic"]   # the oracle's arithmetic
    if matcap:
        cmd += ["-M", os.path.join(MATCAPS, matcap)]
    out = subprocess.run(cmd, capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    assert re.search(r"volumeRender, Throughput = \S+ MTexels/s, Time = \S+ s, Size = %d Texels, NumDevsUsed = 1, Workgroup = 64" % (W * H),
                     out.stdout), out.stdout
    om = oracle.Model.from_h5(model_path("car_1"))
    iv, nm = oracle.view_matrices(0.0, 30.0)
    mc = oracle.load_matcap_png(os.path.join(MATCAPS, matcap)) if matcap else None
    ref = oracle.render(om, iv, nm, W, H, matcap=mc)
    q = int(re.search(r"(\d+) MLP queries", out.stdout).group(1))
    assert q == ref["queries"]
    img = ref["image"].reshape(H, W)
    rgba = np.stack([img & 255, (img >> 8) & 255, (img >> 16) & 255, (img >> 24) & 255], -1).astype(np.uint8)
    png = np.asarray(Image.open(prefix + "car_1.h5.png").convert("RGBA"))
    assert np.array_equal(png, rgba[::-1, ::-1, :])          # Image::savePNG turns the frame by 180 degrees
    ppm = open(prefix + "car_1.h5.ppm", "rb").read()
    head = f"P6\n{W}\n{H}\n255\n".encode()
    assert ppm.startswith(head) and ppm[len(head):] == rgba[:, :, :3].tobytes()


def test_renderer_app_scene_program(gpu, tmp_path):
    """--scene displace: sceneSDF's commented `displacementPattern` (src/volumeRender_kernel.cu:218) selected at run time
    through the C++ mirror's setSceneProgram; query count and frame equal the oracle's."""
    from PIL import Image
    from oracle import oracle
    W, H = 96, 64
    prefix = str(tmp_path) + "/"
    cmd = [os.path.join(ROOT, "apps", "neuralSDFRenderer"), "-i", model_path("plane_1"), "-o", prefix, "-W", str(W), "-H", str(H),
           "-ry", "15", "--single", "--stats", "--math", "deterministic", "--scene", "displace", "-M", os.path.join(MATCAPS, "blue.png")]
    out = subprocess.run(cmd, capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    om = oracle.Model.from_h5(model_path("plane_1"))
    iv, nm = oracle.view_matrices(0.0, 15.0)
    mc = oracle.load_matcap_png(os.path.join(MATCAPS, "blue.png"))
    oracle.set_scene_program([(oracle.PRIM_TANH, oracle.COMB_SET, []), (oracle.MOD_DISPLACE, 0, [5.0, 0.05])])
    try:
        ref = oracle.render(om, iv, nm, W, H, matcap=mc, scene=oracle.SCENE_PROGRAM)
    finally:
        oracle.set_scene_program([])
    assert int(re.search(r"(\d+) MLP queries", out.stdout).group(1)) == ref["queries"]
    img = ref["image"].reshape(H, W)
    rgba = np.stack([img & 255, (img >> 8) & 255, (img >> 16) & 255, (img >> 24) & 255], -1).astype(np.uint8)
    png = np.asarray(Image.open(prefix + "plane_1.h5.png").convert("RGBA"))
    assert np.array_equal(png, rgba[::-1, ::-1, :])
    bad = subprocess.run(cmd[:-4] + ["--scene", "nonsense"], capture_output=True, text=True)
    assert bad.returncode != 0


def test_renderer_app_turntable_batches(gpu, tmp_path):
    """--views N --batch B: the turntable of src/main.cpp:470-478 rendered in batches of B views per persistent launch
    (nsdf_render_views), PNG files written by host threads; every frame equals the oracle's frame of that view."""
    from PIL import Image
    from oracle import oracle
    W, H, N = 96, 64, 5
    prefix = str(tmp_path) + "/"
    cmd = [os.path.join(ROOT, "apps", "neuralSDFRenderer"), "-i", model_path("car_1"), "-o", prefix, "-W", str(W), "-H", str(H),
           "--views", str(N), "--batch", "2", "--scene", "manysphere", "--math", "deterministic", "--stats",
           "-M", os.path.join(MATCAPS, "blue.png")]
    out = subprocess.run(cmd, capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    assert out.stdout.count("saving frame:") == N
    om = oracle.Model.from_h5(model_path("car_1"))
    mc = oracle.load_matcap_png(os.path.join(MATCAPS, "blue.png"))
    for i in range(N):
        iv, nm = oracle.view_matrices(0.0, 360.0 * i / N)
        ref = oracle.render(om, iv, nm, W, H, matcap=mc, scene=oracle.SCENE_MANYSPHERE, frame=i)["image"].reshape(H, W)
        rgba = np.stack([ref & 255, (ref >> 8) & 255, (ref >> 16) & 255, (ref >> 24) & 255], -1).astype(np.uint8)
        png = np.asarray(Image.open(prefix + "%03d.png" % i).convert("RGBA"))
        assert np.array_equal(png, rgba[::-1, ::-1, :]), i

"""GPU side of the multi-GPU layer.  With one GPU: the peer-frame object at world 1 and the tile partition
rendered rank by rank into one buffer.  With >= 2 GPUs: bench.py --workload frame4k_tiles under torchrun, which
checks the gathered frame bytewise against a single-GPU render."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from conftest import MATCAPS, ROOT, model_path

pytestmark = pytest.mark.gpu


def _setup():
    from cudaneuralrender_b200 import host
    nn = host.NeuralNetwork()
    assert nn.load(model_path("car_1"))
    mc = host.Image()
    assert mc.loadPNG(os.path.join(MATCAPS, "blue.png"))
    host.copyStaticSettings(host.COLOR_MATCAP, 3)
    host.updateViewMatrices((0.0, 30.0, 0.0), (0.0, 0.0, -2.0), 0)
    host.setSceneOptions(host.SCENE_TANH, 6000, 0, 0)
    return host, nn, mc


def test_peer_frame_world1_and_emulated_ranks(gpu):
    from cudaneuralrender_b200 import multigpu
    host, nn, mc = _setup()
    W, H = 200, 136
    single = torch.zeros(W * H, dtype=torch.int32, device="cuda")
    host.render_kernel(single, W, H, nn, mc)
    torch.cuda.synchronize()
    ref = single.cpu().numpy().view(np.uint32).reshape(H, W)
    f = multigpu.PeerFrame(W, H, 0, 1)
    multigpu.render_frame_tiled(nn, mc, W, H, f)
    torch.cuda.synchronize()
    assert np.array_equal(f.to_numpy(), ref)
    # three emulated ranks writing into the same (root) buffer, as the peers do over NVLink
    from cudaneuralrender_b200 import _capi as capi
    capi.check(capi.lib().nsdf_dev_memset(f._owned, 0xFF, W * H * 4), "memset")
    for r in range(3):
        host.render_kernel(None, W, H, nn, mc, tile_rank=r, tile_world=3, d_output_ptr=f.ptr)
    torch.cuda.synchronize()
    assert np.array_equal(f.to_numpy(), ref)
    f.close()


def test_multi_renderer_one_device(gpu):
    """nsdf_multi_* with a single device: same frames as nsdf_render / nsdf_render_views (runs on the 1-GPU box)."""
    from cudaneuralrender_b200 import multigpu
    from oracle import oracle
    host, nn, mc = _setup()
    host.setSceneOptions(host.SCENE_MANYSPHERE, 6000, 0, 0, host.MATH_REFERENCE)
    W, H = 200, 136
    views = [(0.0, 30.0), (10.0, 200.0), (0.0, 0.0)]
    ref = torch.zeros(len(views) * W * H, dtype=torch.int32, device="cuda")
    host.render_views(ref, W, H, nn, host.view_cameras(views), mc)
    mr = multigpu.MultiRenderer(nn, [0])
    matcap_np = oracle.load_matcap_png(os.path.join(MATCAPS, "blue.png"))
    out = torch.full_like(ref, 0x55)
    st = mr.render(out, W, H, host.view_cameras(views), matcap_np)
    assert torch.equal(out, ref) and st["pixels"] == len(views) * W * H
    one = torch.full((W * H,), 0x55, dtype=torch.int32, device="cuda")
    mr.render(one, W, H, host.view_cameras(views[:1]), matcap_np)
    assert torch.equal(one, ref[:W * H])
    mr.close()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_multi_renderer_over_all_gpus(gpu):
    """One process, all GPUs of the box: a frame partitioned by tiles (peer stores over NVLink from inside the tracer) and a
    turntable partitioned by views are bytewise equal to the single-GPU frames."""
    from cudaneuralrender_b200 import multigpu
    from oracle import oracle
    host, nn, mc = _setup()
    host.setSceneOptions(host.SCENE_MANYSPHERE, 6000, 0, 0, host.MATH_REFERENCE)
    matcap_np = oracle.load_matcap_png(os.path.join(MATCAPS, "blue.png"))
    n = torch.cuda.device_count()
    mr = multigpu.MultiRenderer(nn, list(range(n)))
    for (W, H) in ((1000, 520), (203, 77)):
        views = [(0.0, 5.625 * k) for k in range(2 * n + 1)]
        ref = torch.zeros(len(views) * W * H, dtype=torch.int32, device="cuda")
        host.render_views(ref, W, H, nn, host.view_cameras(views), mc)
        out = torch.full_like(ref, 0x55)
        st = mr.render(out, W, H, host.view_cameras(views), matcap_np)
        assert torch.equal(out, ref) and st["pixels"] == len(views) * W * H, (W, H)
        if W * H >= 500000:      # every device's share on the throughput tracer (trace_bulk4.cuh), forced through the tuning struct
            host.setTuning(bulk4=1)
            out = torch.full_like(ref, 0x55)
            st = mr.render(out, W, H, host.view_cameras(views), matcap_np)
            host.setTuning()
            assert torch.equal(out, ref) and st["pixels"] == len(views) * W * H, (W, H, "bulk4")
        one = torch.full((W * H,), 0x55, dtype=torch.int32, device="cuda")
        st = mr.render(one, W, H, host.view_cameras(views[:1]), matcap_np)
        assert torch.equal(one, ref[:W * H]) and st["pixels"] == W * H, (W, H)
    mr.close()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_tiles_over_two_gpus_match_single(gpu):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "bench.py"), "--gpus", "2", "--workload", "frame4k_tiles", "--steps", "2",
           "--warmup", "3"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][-1])
    leg = line["frame4k_tiles"]
    assert line["n_gpus"] == 2 and leg["tiles_equal_single_gpu"] is True
    assert leg["tile_balance_queries"] > 0.9

"""nsdf_render_noalloc (include/nsdf.h): the allocation-free variant of nsdf_render, called through the C-ABI and compared with
nsdf_render itself.  The entry point has no counterpart in the reference; the bar
is the one of the other frame tests: the same bytes.  Named test_zz_* so that pytest collects it last."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import MATCAPS, model_path

pytestmark = pytest.mark.gpu

ETOOLARGE = 10005


def _render(fn, nn, cam, mat, opts, out, W, H, stats, name):
    from cudaneuralrender_b200 import _capi as capi
    mat_ptr, mw, mh = (mat.deviceData.data_ptr(), mat.shape.x, mat.shape.y) if mat is not None else (None, 0, 0)
    capi.check(fn(nn.handle(), C.byref(cam), mat_ptr, mw, mh, C.byref(opts), out.data_ptr(), W, H, stats.data_ptr(), None), name)


def test_noalloc_entry_point_matches_nsdf_render(gpu):
    import torch
    from cudaneuralrender_b200 import host
    from cudaneuralrender_b200 import _capi as capi
    L = capi.lib()
    nn = host.NeuralNetwork()
    assert nn.load(model_path("plane_1"))
    matcap = host.Image()
    assert matcap.loadPNG(os.path.join(MATCAPS, "blue.png"))
    keep = [0, 1, 2, 4, 5]          # queries, hit pixels, rays, unfinished, pixels ("evicted" is schedule-dependent)

    # two calls in a row, without a synchronisation between them: different views, sizes, scenes, colourings and math modes
    cases = [dict(W=640, H=360, view=(0.0, 30.0), frame=0, scene=capi.SCENE_MANYSPHERE, math=capi.MATH_REFERENCE, mat=matcap),
             dict(W=203, H=157, view=(10.0, 200.0), frame=7, scene=capi.SCENE_TANH, math=capi.MATH_DETERMINISTIC, mat=None)]
    expect, got = [], []
    for c in cases:
        opts = capi.RenderOpts()
        L.nsdf_render_opts_default(C.byref(opts))
        opts.scene, opts.math = c["scene"], c["math"]
        opts.color_type = host.COLOR_MATCAP if c["mat"] is not None else host.COLOR_FACING
        c["opts"], c["cam"] = opts, host.view_cameras([c["view"]], [c["frame"]])[0]
        ref = torch.full((c["W"] * c["H"],), 0x55, dtype=torch.int32, device="cuda")
        st = torch.zeros(6, dtype=torch.int64, device="cuda")
        _render(L.nsdf_render, nn, c["cam"], c["mat"], opts, ref, c["W"], c["H"], st, "nsdf_render")
        expect.append((ref, st))
    torch.cuda.synchronize()
    for rounds in range(2):          # the second round starts from the scratch contents the first one left behind
        got = []
        for c in cases:
            out = torch.full((c["W"] * c["H"],), 0x55, dtype=torch.int32, device="cuda")
            st = torch.zeros(6, dtype=torch.int64, device="cuda")
            got.append((out, st))
        free_before = torch.cuda.mem_get_info()[0]
        for c, (out, st) in zip(cases, got):
            _render(L.nsdf_render_noalloc, nn, c["cam"], c["mat"], c["opts"], out, c["W"], c["H"], st, "nsdf_render_noalloc")
        torch.cuda.synchronize()
        free_after = torch.cuda.mem_get_info()[0]
        for i, ((ref, st1), (out, st2)) in enumerate(zip(expect, got)):
            assert torch.equal(out, ref), (rounds, i)
            assert torch.equal(st1[keep], st2[keep]), (rounds, i, st1.tolist(), st2.tolist())
            assert int((out != 0).sum()) > 1000          # a real frame, not two empty ones
        if rounds == 1:              # set-up (module load, function attributes) happened in round 0: nothing may be taken now
            assert free_after >= free_before, (free_before, free_after)

    # beyond the fixed maximum: a clear error before anything is launched, and the next call is unaffected
    c = cases[1]
    out = torch.full((c["W"] * c["H"],), 0x55, dtype=torch.int32, device="cuda")
    st = torch.zeros(6, dtype=torch.int64, device="cuda")
    rc = L.nsdf_render_noalloc(nn.handle(), C.byref(c["cam"]), None, 0, 0, C.byref(c["opts"]), out.data_ptr(), 7680, 4320,
                               st.data_ptr(), None)
    assert rc == ETOOLARGE
    assert b"maximum" in L.nsdf_error_string(rc)
    torch.cuda.synchronize()
    assert int((out != 0x55).sum()) == 0
    # largest supported frame class is accepted by the size check (3840 x 2160 = 60 x 34 tiles)
    assert 60 * 34 <= 2048
    _render(L.nsdf_render_noalloc, nn, c["cam"], None, c["opts"], out, c["W"], c["H"], st, "nsdf_render_noalloc")
    torch.cuda.synchronize()
    assert torch.equal(out, expect[1][0])
    # tile partitions and the A/B flags belong to nsdf_render: refused, not ignored
    bad = capi.RenderOpts()
    L.nsdf_render_opts_default(C.byref(bad))
    bad.tile_world, bad.tile_rank = 2, 1
    assert L.nsdf_render_noalloc(nn.handle(), C.byref(c["cam"]), None, 0, 0, C.byref(bad), out.data_ptr(), c["W"], c["H"],
                                 st.data_ptr(), None) == 10002

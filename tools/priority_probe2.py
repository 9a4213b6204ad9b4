import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudaneuralrender_b200 import _capi as capi, host
def run(nn, mc, W, H, flags, evict=0, iters=4):
    host.setSceneOptions(host.SCENE_TANH, 6000, flags, evict)
    fr = torch.zeros(W * H, dtype=torch.int32, device="cuda")
    for _ in range(2): host.render_kernel(fr, W, H, nn, mc)
    torch.cuda.synchronize(); ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); host.render_kernel(fr, W, H, nn, mc); b.record(); b.synchronize(); ts.append(a.elapsed_time(b))
    return min(ts), fr.cpu().numpy()
def setenv(**kw):
    for k, v in kw.items(): os.environ[k] = str(v)
for model, W, H in (("plane_1", 1920, 1080), ("3a3d4a90a2db90b4203936772104a82d.obj", 3840, 2160)):
    nn = host.NeuralNetwork(); assert nn.load(os.path.join(ROOT, "data", "neuralGeometries", model + ".h5"))
    mc = host.Image(); mc.loadPNG(os.path.join(ROOT, "data", "matcaps", "blue.png"))
    host.copyStaticSettings(host.COLOR_MATCAP, 3); host.updateViewMatrices((0, 0, 0), (0, 0, -2), 0)
    setenv(NSDF_TWO_PASS_MIN_PIXELS=1 << 30)
    for steps in (0, 32):
        setenv(NSDF_TILE_PRIORITY_STEPS=steps)
        row = []
        for period in (16, 8, 4):
            for ev in (32, 64, 128):
                setenv(NSDF_FUSED_CONSUMER_PERIOD=period, NSDF_FUSED_EVICT_HI=ev, NSDF_FUSED_EVICT_STEPS=ev)
                t, a = run(nn, mc, W, H, capi.RENDER_QUEUE)
                row.append(f"p{period}/e{ev}: {t:.2f}")
        print(model[:8], f"QUEUE schedule, probe_steps {steps}: " + " | ".join(row), flush=True)

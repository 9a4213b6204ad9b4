#!/bin/bash
# compute-sanitizer target (the reference has no sanitizer hooks; the tracer rests on hand-rolled inter-CTA queues).
#   tools/sanitize.sh memcheck|racecheck|synccheck|initcheck      ONE tool per GPU-box call (see B200_PROFILING.md)
# Renders a 128x128 frame (small grid: no express lanes) and a 320x240 frame (full grid: express lanes, tile hand-over off) of
# car_1 in the reference's scene, a 3-way staged tile partition and a 3-view batch, through the product library.
set -e
TOOL=${1:-memcheck}
cd "$(dirname "$0")/.."
compute-sanitizer --tool "$TOOL" --error-exitcode 9 python - <<'PY'
import os, sys
import torch
sys.path.insert(0, os.getcwd())
from cudaneuralrender_b200 import host
nn = host.NeuralNetwork(); assert nn.load("data/neuralGeometries/car_1.h5")
mc = host.Image(); assert mc.loadPNG("data/matcaps/blue.png")
host.copyStaticSettings(host.COLOR_MATCAP, 3)
host.setSceneOptions(host.SCENE_MANYSPHERE, 300, 0, 0, host.MATH_REFERENCE)   # 300 steps bound the run time under the tool
for (W, H) in ((128, 128), (320, 240)):
    host.updateViewMatrices((0.0, 30.0, 0.0), (0.0, 0.0, -2.0), 0)
    out = torch.zeros(W * H, dtype=torch.int32, device="cuda")
    host.render_kernel(out, W, H, nn, mc)
    part = torch.zeros(W * H, dtype=torch.int32, device="cuda")
    for r in range(3):
        host.render_kernel(part, W, H, nn, mc, tile_rank=r, tile_world=3)
    torch.cuda.synchronize()
    assert torch.equal(out, part)
    views = torch.zeros(3 * W * H, dtype=torch.int32, device="cuda")
    host.render_views(views, W, H, nn, host.view_cameras([(0.0, 30.0), (10.0, 90.0), (0.0, 200.0)]), mc)
    torch.cuda.synchronize()
    assert torch.equal(views[:W * H], out)
print("sanitize workload ok")
PY

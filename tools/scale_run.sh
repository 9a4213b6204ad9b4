#!/bin/bash
# Scaling run on one box: bench.py's default line at N GPUs (64-view 1080p turntable split over the ranks, strong scaling) --
# which carries the frame4k_tiles leg (BASELINE config 4: one 4K frame, tiles over the N GPUs, peer stores over NVLink) and the
# turntable_c5 leg -- plus NVLink data counters of GPU 0 around a burst of tile-partitioned 4K frames.
# usage: tools/scale_run.sh N        -> gpurun_out/scale_n$N.json, gpurun_out/nvlink_n$N.txt
n=${1:-2}
mkdir -p gpurun_out
if [ "$n" = 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n))"; fi
$L bench.py --gpus $n --steps 10 --warmup 3 --no-dense --no-cpu --no-ref-cuda --no-c1 --no-mixed > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
echo "bench rc=$?"
if [ "$n" != 1 ]; then
  nvidia-smi nvlink -gt d -i 0 > gpurun_out/nvlink_n${n}_before.txt 2>&1
  $L bench.py --gpus $n --workload frame4k_tiles --steps 50 --warmup 3 > gpurun_out/tiles_n$n.json 2> gpurun_out/tiles_n$n.err
  nvidia-smi nvlink -gt d -i 0 > gpurun_out/nvlink_n${n}_after.txt 2>&1
  python - "$n" <<'PY'
import re, sys, json
n = int(sys.argv[1])
def total(path, key):
    t = 0
    for l in open(path):
        m = re.search(r"Data %s:\s*(\d+)\s*KiB" % key, l)
        if m: t += int(m.group(1))
    return t
try:
    rx = total(f"gpurun_out/nvlink_n{n}_after.txt", "Rx") - total(f"gpurun_out/nvlink_n{n}_before.txt", "Rx")
    tx = total(f"gpurun_out/nvlink_n{n}_after.txt", "Tx") - total(f"gpurun_out/nvlink_n{n}_before.txt", "Tx")
    d = json.loads([l for l in open(f"gpurun_out/tiles_n{n}.json") if l.startswith("{")][-1])
    leg = d["frame4k_tiles"]
    frames = 2 + 1 + 1 + 2 * leg["steps"] + 3      # warm-ups, accounting, timed, end-to-end passes of tiles_leg (approximately)
    expect = 3840 * 2160 * 4 * (n - 1) / n / 1024
    print(json.dumps({"n_gpus": n, "gpu0_nvlink_rx_kib": rx, "gpu0_nvlink_tx_kib": tx, "tile_frames_rendered_about": frames,
                      "rx_kib_per_frame": rx / frames, "expected_payload_kib_per_frame": expect,
                      "ms_per_frame": leg["ms_per_frame"], "tiles_equal_single_gpu": leg["tiles_equal_single_gpu"]}))
except Exception as e:
    print("nvlink summary failed:", e)
PY
fi
python - "$n" <<'PY'
import json, sys
n = sys.argv[1]
try:
    d = json.loads([l for l in open(f"gpurun_out/scale_n{n}.json") if l.startswith("{")][-1])
    print("N", n, "value", round(d["value"], 2), d["unit"], "ms/step", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["value"], 2),
          "| 4K tiles", round(d["frame4k_tiles"]["ms_per_frame"], 2), "ms equal", d["frame4k_tiles"]["tiles_equal_single_gpu"],
          "balance", d["frame4k_tiles"]["tile_balance_queries"], "| C5", round(d["turntable_c5"]["frames_per_s"], 1), "frames/s")
except Exception as e:
    print("scale summary failed:", e)
PY

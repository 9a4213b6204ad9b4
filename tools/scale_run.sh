#!/bin/bash
# Scaling runs on one box: default workload (one 1080p view per GPU, weak) and BASELINE config 4 (one 4K frame, tiles, strong).
# usage: tools/scale_run.sh "1 2 4 8"   -> gpurun_out/scale_*.json
mkdir -p gpurun_out
for n in ${1:-1 2 4 8}; do
  if [ "$n" = 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n))"; fi
  $L bench.py --gpus $n --steps 10 --warmup 3 --no-dense --no-cpu --no-ref-cuda > gpurun_out/scale_views_n$n.json 2> gpurun_out/scale_views_n$n.err
  $L bench.py --gpus $n --workload frame4k_tiles --steps 10 --warmup 3 > gpurun_out/scale_tiles_n$n.json 2> gpurun_out/scale_tiles_n$n.err
  NSDF_BENCH_TILES_MIXED=1 $L bench.py --gpus $n --workload frame4k_tiles --steps 10 --warmup 3 > gpurun_out/scale_tiles_mixed_n$n.json 2> gpurun_out/scale_tiles_mixed_n$n.err
done
python - <<'PY'
import json, glob
for kind in ("views", "tiles", "tiles_mixed"):
    base = None
    for n in (1, 2, 4, 8):
        try:
            d = json.loads([l for l in open(f"gpurun_out/scale_{kind}_n{n}.json") if l.startswith("{")][-1])
        except Exception as e:
            print(kind, n, "missing", e); continue
        base = base or d["value"]
        print(kind, "n", n, "value", round(d["value"], 2), d["unit"], "ms", round(d["ms_per_step"], 2), "eff", round(d["value"] / base / n, 3),
              "equal" , d.get("tiles_equal_single_gpu"), "balance", d.get("tile_balance_queries"))
PY

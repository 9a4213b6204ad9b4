#!/usr/bin/env python
"""Summarises an .ncu-rep (read here, no GPU needed) into the text kept under profiles/.

  python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/rNN_ncu_<kernel>.txt
Prints the raw-page metrics the roofline uses and the hottest SASS lines of the source page.
"""
import csv
import io
import subprocess
import sys

KEYS = ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput",
        "sm__pipe_tensor_cycles_active", "sm__pipe_fma_cycles_active", "sm__pipe_alu_cycles_active",
        "sm__inst_executed_pipe_tc", "sm__inst_executed_pipe_uniform", "sm__warps_active.avg.pct", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__occupancy_limit", "smsp__issue_active.avg.pct",
        "smsp__inst_executed.sum", "sm__cycles_active.avg", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts.avg", "sm__throughput.avg.pct", "l1tex__data_pipe_tc_wavefronts_mem_shared.sum",
        "launch__shared_mem_per_block", "sm__cycles_elapsed.avg ", "lts__t_bytes.sum ")


def page(rep, name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    raw = page(rep, "raw")
    head, units = raw[0], raw[1]
    for row in raw[2:]:
        print("== raw page:", row[head.index("Kernel Name")])
        for k, v, u in zip(head, row, units):
            if any(k.startswith(s.strip()) for s in KEYS) and ".max." not in k and ".min." not in k and ".sum.pct" not in k:
                print(f"  {k} = {v} {u}")
    src = page(rep, "source")
    hi = next(i for i, r in enumerate(src) if r and r[0] == "Address")
    h = src[hi]
    ia, ie, isamp = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
    stalls = [i for i, k in enumerate(h) if k.startswith("stall_") and "Not Issued" not in k]
    body = [r for r in src[hi + 1:] if len(r) == len(h)]
    tot_e = sum(int(r[ie]) for r in body)
    tot_s = sum(int(r[isamp]) for r in body)
    print(f"== source page: {len(body)} SASS lines, {tot_e} warp instructions, {tot_s} stall samples")
    agg = {}
    for r in body:
        for i in stalls:
            agg[h[i]] = agg.get(h[i], 0) + int(r[i])
    print("  stall totals:", ", ".join(f"{k}={v}" for k, v in sorted(agg.items(), key=lambda x: -x[1])[:8]))
    print(f"  top {top} lines by samples:  [line] SASS | executed | samples | top stalls")
    order = sorted(range(len(body)), key=lambda n: -int(body[n][isamp]))[:top]
    for n in sorted(order):
        r = body[n]
        t = sorted(((int(r[i]), h[i]) for i in stalls), reverse=True)[:2]
        print(f"  [{n:4d}] {r[ia].strip()[:64]:64s} | {r[ie]:>10s} | {r[isamp]:>6s} | " + ", ".join(f"{k}={v}" for v, k in t if v))


if __name__ == "__main__":
    main()

#!/bin/bash
# development: frame time of the default tracer under a few scheduling knobs (GPU box)
run() { echo -n "$* : "; env "$@" timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu --no-dense 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('fps %.1f ms %.2f two-launch %.2f evicted %d e2e %.1f' % (d['value'], d['ms_per_step'], d['kernels']['two_launch_schedule_ms'], d['evicted_rays'], d['e2e']['value']))"; }
run NSDF_COOP_QUEUE=0
run NSDF_COOP_QUEUE=1
run NSDF_COOP_EVICT=128
run NSDF_COOP_EVICT=512
run NSDF_COOP_QUIET_PERIOD=8
run NSDF_COOP_QUIET_PERIOD=32
run NSDF_COOP_QUIET_PERIOD=32 NSDF_COOP_EVICT=128

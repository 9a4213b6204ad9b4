run() { echo -n "$* : "; env $1 timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu --no-dense --max-steps $2 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('fps %.1f ms %.2f two-launch %.2f evicted %d q %d' % (d['value'], d['ms_per_step'], d['kernels']['two_launch_schedule_ms'], d['evicted_rays'], d['queries_per_frame']))"; }
for ms in 6000 2048 1024 512 256; do run NSDF_COOP_QUEUE=0 $ms; done
run NSDF_COOP_QUEUE=1 1024

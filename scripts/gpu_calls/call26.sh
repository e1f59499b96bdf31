mkdir -p gpurun_out
for i in 1 2 3; do timeout 600 python bench.py --workload greedy --only --no-cpu --page-lines 0 --predict-lines 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('greedy value %.4g e2e %.4g ms %.1f kms %.1f launches %d clocks %s' % (d['value'], d['e2e']['value'], d['ms_per_step'], r['kernel_ms_per_step'], d['gpu_launches'], d['clocks']['sm_mhz']))
"; done

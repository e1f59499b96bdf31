# round 2, call 36: (a) are the three differing parity rows of the competitive leg the same with the full merge copy
# (COR_NO_REACH_TRUNC=1) and one-block push (COR_PUSH_PARTS=1)?  (b) ncu --set full of the push / select kernels at head
mkdir -p gpurun_out
for sw in "COR_NO_REACH_TRUNC=1 COR_PUSH_PARTS=1 COR_NO_SIZED_RETRY=1" "COR_X=1"; do
  echo "== $sw"
  env $sw timeout 900 python bench.py --workload beam_competitive --only --no-cpu --steps 1 --warmup 1 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); p = d['parity']
print(p['parity_rows_checked'], p['rows_identical'], p['rows_differing_at_logged_near_tie'], p['rows_differing_otherwise'])
for r in p.get('differing_rows', []): print(r)"
done
COR_NO_GRAPH=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:beam_push_kernel -s 40 -c 3 -o gpurun_out/r02_ncu_push_head -f python scripts/beam_probe.py beam_competitive 64 1 > gpurun_out/ncu_push_head.log 2>&1; tail -1 gpurun_out/ncu_push_head.log | cut -c1-120
COR_NO_GRAPH=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:beam_select_small_kernel -s 40 -c 3 -o gpurun_out/r02_ncu_select_head -f python scripts/beam_probe.py beam_competitive 64 1 > gpurun_out/ncu_select_head.log 2>&1; tail -1 gpurun_out/ncu_select_head.log | cut -c1-120
ls -la gpurun_out/*.ncu-rep

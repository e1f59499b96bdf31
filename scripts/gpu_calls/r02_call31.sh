# round 2, call 31: retry waves sized from what the failed lines report (need-sorted, power-of-two waves) against the old
# "an eighth of the lines per wave" rule (COR_NO_SIZED_RETRY=1): parity tests, then A/B
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_beam.py tests/test_gpu_loop_graph.py tests/test_gpu_timed_shape.py tests/test_gpu_reference_pin.py -m gpu -x -q 2>&1 | tail -4
for w in "beam_competitive 256" "beam_competitive 512" "beam_stress 64"; do
  echo "== $w: eighth-of-the-lines retry, then sized retry"
  COR_NO_SIZED_RETRY=1 timeout 300 python scripts/beam_probe.py $w 2 2>&1 | tail -1
  timeout 300 python scripts/beam_probe.py $w 2 2>&1 | tail -1
done
for sw in 1 0; do
  echo "== bench competitive bracket (2048 lines, 72 GB), COR_NO_SIZED_RETRY=$sw"
  if [ $sw = 1 ]; then export COR_NO_SIZED_RETRY=1; else unset COR_NO_SIZED_RETRY; fi
  timeout 900 python bench.py --workload beam_competitive --only --no-cpu --no-parity --steps 3 --warmup 3 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); print({k: d.get(k) for k in ('metric','value','ms_per_step','gpu_launches')}, d.get('e2e'))"
done

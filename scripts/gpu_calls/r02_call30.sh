# round 2, call 30: beam_push_kernel with several blocks per line (sort per block, merge in slices, last finisher commits)
# + one child per thread in the collect: parity tests, then A/B against one block per line (COR_PUSH_PARTS=1)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_beam.py tests/test_gpu_loop_graph.py tests/test_gpu_timed_shape.py tests/test_gpu_reference_pin.py -m gpu -x -q 2>&1 | tail -4
for w in "beam_competitive 256" "beam_competitive 512" "beam_stress 64" "beam_competitive 32"; do
  echo "== $w: one block per line, then parts"
  COR_PUSH_PARTS=1 timeout 300 python scripts/beam_probe.py $w 2 2>&1 | tail -1
  timeout 300 python scripts/beam_probe.py $w 2 2>&1 | tail -1
done
echo "== page leg"
timeout 600 python bench.py --page-only 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); print(json.dumps(d.get('page', d))[:1500])"

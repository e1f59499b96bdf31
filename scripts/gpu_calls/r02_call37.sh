# round 2, call 37: pop in rounds of 4 chunks (ids and record reads fetched together), bitonic sort with warp-shuffle steps,
# near-tie log with an absolute term at the threshold cuts: parity tests, probes (before: 288 / 412 / 187 ms, page beam 7.7-8.1 ms), bench leg parity
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_beam.py tests/test_gpu_loop_graph.py tests/test_gpu_timed_shape.py tests/test_gpu_reference_pin.py tests/test_gpu_seq2seq.py -m gpu -x -q 2>&1 | tail -3
for w in "beam_competitive 256" "beam_competitive 512" "beam_stress 64" "beam_competitive 32"; do
  timeout 300 python scripts/beam_probe.py $w 3 2>&1 | tail -1 | cut -c1-75,230-400
done
timeout 600 python bench.py --page-only 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); p = d.get('page', d); print('page greedy', p['ms_per_call'], 'beam', p['beam']['ms_per_call'])"
timeout 900 python bench.py --workload beam_competitive --only --no-cpu 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); p = d['parity']
print(round(d['value']), d['ms_per_step'], p['parity_rows_checked'], p['rows_identical'], p['rows_differing_at_logged_near_tie'], p['rows_differing_otherwise'])
for r in p.get('differing_rows', []): print(r)"

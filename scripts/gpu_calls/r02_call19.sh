# round 2, call 19: row compaction in the fast-greedy loop graph
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests/test_gpu_loop_graph.py tests/test_gpu_core.py tests/test_gpu_timed_shape.py tests/test_gpu_seq2seq.py tests/test_gpu_reference_pin.py -m gpu -x -q -s 2>&1 | grep -E "greedy graph|passed|failed|Error|assert" | tail -12 ) | tee gpurun_out/r02_c19_tests.log
python - <<'PY'
import json, sys, os, argparse
sys.path.insert(0, os.getcwd())
import bench
args = argparse.Namespace(gemm_mode=0)
print(json.dumps(bench.measure_early_exit(args, 0)))
PY
timeout 600 python bench.py --only --workload greedy --no-cpu --strong-greedy-lines 0 --strong-beam-lines 0 > gpurun_out/r02_c19_greedy.json 2>> gpurun_out/r02_c19.err
python -c "import json;d=json.load(open('gpurun_out/r02_c19_greedy.json'));print('greedy', round(d['value']), round(d['e2e']['value']), d['ms_per_step'], d['parity'], d['gpu_launches'], d['host_launch_calls'])"
tail -3 gpurun_out/r02_c19.err

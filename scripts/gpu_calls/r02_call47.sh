# round 2, call 47: 128x32-tile GEMM with a fourth MMA issuer warp (before: 7.2 us at K = 512, 11.8 us at K = 1536, page 6.75 ms)
# GEMM tests, warm kernel time (before: 7.84 us at K = 512, 13.7 us at K = 1536), page leg, the suites that use the kernel
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -2
for shape in "32 2048 512" "32 2048 1536" "256 2048 1536"; do
  timeout 600 ncu --metrics gpu__time_duration.sum --cache-control none --clock-control none -k regex:gemm_tc_kernel --csv python scripts/gemm_probe.py 3 $shape 2>&1 | grep gemm_tc | awk -F'","' -v s="$shape" '{print s, $NF}' | tail -2
done
timeout 600 python bench.py --page-only 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); p = d.get('page', d); print('page greedy', p['ms_per_call'], 'beam', p['beam']['ms_per_call'])"
timeout 1500 python -m pytest tests/test_gpu_core.py tests/test_gpu_loop_graph.py tests/test_gpu_beam.py tests/test_gpu_seq2seq.py tests/test_gpu_timed_shape.py tests/test_gpu_reference_pin.py -m gpu -x -q 2>&1 | tail -4

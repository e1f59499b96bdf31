# round 2, call 8: push kernel with 1024 threads per line; attention beside the lower-cell GEMM (register-only side kernel)
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_beam.py tests/test_gpu_loop_graph.py tests/test_gpu_timed_shape.py -m gpu -x -q 2>&1 | tail -5 ) | tee gpurun_out/r02_c08_tests.log
python scripts/beam_probe.py beam_competitive 512 2 | tee gpurun_out/r02_c08_comp512.json
python scripts/beam_probe.py beam_stress 64 2 | tee gpurun_out/r02_c08_stress64.json
( COR_OVERLAP=1 COR_ATTN_SIDE=1 timeout 900 python -m pytest tests/test_gpu_core.py tests/test_gpu_timed_shape.py -m gpu -x -q 2>&1 | tail -5 ) | tee gpurun_out/r02_c08_tests_overlap.log
for mode in base overlap side; do
  case $mode in base) E="";; overlap) E="COR_OVERLAP=1";; side) E="COR_OVERLAP=1 COR_ATTN_SIDE=1";; esac
  for wl in greedy beam; do
    env $E timeout 600 python bench.py --only --workload $wl --no-cpu --no-parity --strong-greedy-lines 0 --strong-beam-lines 0 > gpurun_out/r02_c08_${wl}_$mode.json 2>> gpurun_out/r02_c08.err
    python -c "import json;d=json.load(open('gpurun_out/r02_c08_${wl}_$mode.json'));print('$wl $mode', round(d['value']), round(d['e2e']['value']), d['ms_per_step'], d['clocks'])"
  done
done
tail -5 gpurun_out/r02_c08.err

# round 2, call 15: LSTM epilogue with TMA box stores from a staging tile: parity, then A/B against direct stores
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests/test_gpu_core.py tests/test_gpu_timed_shape.py tests/test_gpu_beam.py tests/test_gpu_gemm.py tests/test_gpu_reference_pin.py -m gpu -x -q 2>&1 | tail -8 ) | tee gpurun_out/r02_c15_tests.log
for mode in tma direct; do
  case $mode in tma) E="";; direct) E="COR_NO_EPI_TMA=1";; esac
  for wl in greedy beam; do
    env $E timeout 600 python bench.py --only --workload $wl --no-cpu --no-parity --strong-greedy-lines 0 --strong-beam-lines 0 > gpurun_out/r02_c15_${wl}_$mode.json 2>> gpurun_out/r02_c15.err
    python -c "import json;d=json.load(open('gpurun_out/r02_c15_${wl}_$mode.json'));print('$wl $mode', round(d['value']), round(d['e2e']['value']), d['ms_per_step'], d['roofline']['achieved'], d['roofline']['kernel_ms_per_step'], d['clocks'])"
  done
done
tail -5 gpurun_out/r02_c15.err

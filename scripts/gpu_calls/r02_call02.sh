# round 2, call 2: beam search as a WHILE graph -- parity tests, page latency and C2 beam throughput, with the host loop beside it
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_beam.py tests/test_gpu_reference_pin.py tests/test_gpu_timed_shape.py -m gpu -x -q 2>&1 | tail -5 ) | tee gpurun_out/r02_c02_tests.log
for g in 0 1; do
  COR_NO_GRAPH=$g timeout 300 python bench.py --page-only --page-lines 32 > gpurun_out/r02_c02_page32_nograph$g.json 2>> gpurun_out/r02_c02.err; cut -c1-700 gpurun_out/r02_c02_page32_nograph$g.json; echo
  COR_NO_GRAPH=$g timeout 600 python bench.py --only --workload beam --no-cpu --no-parity --strong-greedy-lines 0 --strong-beam-lines 0 > gpurun_out/r02_c02_beam_nograph$g.json 2>> gpurun_out/r02_c02.err; cut -c1-300 gpurun_out/r02_c02_beam_nograph$g.json; echo
done
tail -5 gpurun_out/r02_c02.err

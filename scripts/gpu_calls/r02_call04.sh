# round 2, call 4: loop graphs with the finished-line counter: tests, greedy A/B, launch list of one beamed page
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_loop_graph.py tests/test_gpu_gemm.py "tests/test_gpu_core.py::test_small_magnitude_weights_variant_k" -m gpu -x -q -s 2>&1 | tail -15 ) | tee gpurun_out/r02_c04_tests.log
for g in 0 1; do
  COR_NO_GRAPH=$g timeout 600 python bench.py --only --workload greedy --no-cpu --no-parity --strong-greedy-lines 0 --strong-beam-lines 0 > gpurun_out/r02_c04_greedy_nograph$g.json 2>> gpurun_out/r02_c04.err; cut -c1-300 gpurun_out/r02_c04_greedy_nograph$g.json; echo
done
COR_NO_GRAPH=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r02_c04_launches_page32_beam.csv python bench.py --page-only --page-lines 32 > gpurun_out/r02_c04_ncu_page.log 2>&1
tail -2 gpurun_out/r02_c04_ncu_page.log | cut -c1-300
tail -5 gpurun_out/r02_c04.err

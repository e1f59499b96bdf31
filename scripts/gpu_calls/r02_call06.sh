# round 2, call 6: SWITCH node for the GEMM tile shape + one-block select for a page of lines
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_loop_graph.py tests/test_gpu_beam.py tests/test_gpu_reference_pin.py tests/test_gpu_timed_shape.py -m gpu -x -q 2>&1 | tail -8 ) | tee gpurun_out/r02_c06_tests.log
python scripts/page_beam_probe.py 32 10 | tee gpurun_out/r02_c06_page_beam.json
COR_NO_SHAPE_SWITCH=1 python scripts/page_beam_probe.py 32 10 | tee gpurun_out/r02_c06_page_beam_noswitch.json
COR_NO_SELECT_SMALL=1 python scripts/page_beam_probe.py 32 10 | tee gpurun_out/r02_c06_page_beam_nosel.json
python scripts/page_beam_probe.py 64 10 | tee gpurun_out/r02_c06_page64_beam.json
python scripts/page_beam_probe.py 8 10 | tee gpurun_out/r02_c06_page8_beam.json
timeout 600 python bench.py --only --workload beam_competitive --no-cpu --strong-greedy-lines 0 --strong-beam-lines 0 > gpurun_out/r02_c06_comp.json 2>> gpurun_out/r02_c06.err; cut -c1-400 gpurun_out/r02_c06_comp.json; echo
COR_NO_SHAPE_SWITCH=1 timeout 600 python bench.py --only --workload beam_competitive --no-cpu --no-parity --strong-greedy-lines 0 --strong-beam-lines 0 > gpurun_out/r02_c06_comp_noswitch.json 2>> gpurun_out/r02_c06.err; cut -c1-400 gpurun_out/r02_c06_comp_noswitch.json; echo
tail -5 gpurun_out/r02_c06.err

# round 2, call 14: line kernel (expand + push + pop per line) and gather-with-scan for a page of lines
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_beam.py tests/test_gpu_loop_graph.py tests/test_gpu_reference_pin.py tests/test_gpu_seq2seq.py -m gpu -x -q 2>&1 | tail -8 ) | tee gpurun_out/r02_c14_tests.log
for n in 8 32 64; do python scripts/page_beam_probe.py $n 10 | tee gpurun_out/r02_c14_page${n}_beam.json; done
COR_NO_LINE_KERNEL=1 python scripts/page_beam_probe.py 32 10 | tee gpurun_out/r02_c14_page32_beam_noline.json
COR_NO_GRAPH=1 python scripts/page_beam_probe.py 32 10 | tee gpurun_out/r02_c14_page32_beam_hostloop.json
python scripts/beam_probe.py beam_competitive 64 2 | tee gpurun_out/r02_c14_comp64.json
COR_NO_LINE_KERNEL=1 python scripts/beam_probe.py beam_competitive 64 2 | tee gpurun_out/r02_c14_comp64_noline.json

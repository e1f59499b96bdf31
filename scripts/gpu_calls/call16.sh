set -x
mkdir -p gpurun_out

timeout 300 python -m pytest tests/test_gpu_core.py -m gpu -x -q -k "page_sized or c2_size" 2>&1 | tail -15 | tee gpurun_out/gpu_tests_page.log
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -12 | tee gpurun_out/gpu_tests.log
for n in 8 32; do timeout 300 python bench.py --page-only --page-lines $n > gpurun_out/page6_$n.json 2> gpurun_out/page6.err; cat gpurun_out/page6_$n.json; done
tail -3 gpurun_out/page6.err

set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -12 | tee gpurun_out/gpu_tests.log
for n in 8 32; do timeout 300 python bench.py --page-only --page-lines $n > gpurun_out/page5_$n.json 2> gpurun_out/page5.err; cat gpurun_out/page5_$n.json; done
COR_NO_RECUR=1 timeout 300 python bench.py --page-only --page-lines 32 > gpurun_out/page5_32_norecur.json 2>> gpurun_out/page5.err; cat gpurun_out/page5_32_norecur.json
tail -3 gpurun_out/page5.err

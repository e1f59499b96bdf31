set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/gpu_tests.log
for n in 8 32 64 128 256; do python bench.py --page-only --page-lines $n > gpurun_out/page4_$n.json 2>> gpurun_out/page4.err; done
COR_NO_NARROW=1 python bench.py --page-only --page-lines 32 > gpurun_out/page4_wide_32.json 2>> gpurun_out/page4.err
COR_NO_NARROW=1 python bench.py --page-only --page-lines 256 > gpurun_out/page4_wide_256.json 2>> gpurun_out/page4.err
cat gpurun_out/page4_*.json
ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 1200 --csv --log-file gpurun_out/launches_page32_narrow.csv python bench.py --page-only --page-lines 32 > gpurun_out/ncu_p4.log 2>&1

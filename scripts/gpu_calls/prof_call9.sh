set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
for n in 8 32 64; do python bench.py --page-only --page-lines $n > gpurun_out/page2_$n.json 2>> gpurun_out/page2.err; done
COR_SKINNY_MAX_ROWS=128 python bench.py --page-only --page-lines 128 > gpurun_out/page2_128s.json 2>> gpurun_out/page2.err
COR_SKINNY_MAX_ROWS=32 python bench.py --page-only --page-lines 64 > gpurun_out/page2_64tc.json 2>> gpurun_out/page2.err
cat gpurun_out/page2_*.json
ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 1200 --csv --log-file gpurun_out/launches_page32_skinny.csv python bench.py --page-only --page-lines 32 > gpurun_out/ncu_p2.log 2>&1

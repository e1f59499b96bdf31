set -x
mkdir -p gpurun_out
for n in 8 32; do COR_SKINNY_MAX_ROWS=0 python bench.py --page-only --page-lines $n > gpurun_out/page3_tc_$n.json 2>> gpurun_out/page3.err; done
cat gpurun_out/page3_*.json
COR_SKINNY_MAX_ROWS=0 ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 1200 --csv --log-file gpurun_out/launches_page32_tc.csv python bench.py --page-only --page-lines 32 > gpurun_out/ncu_p3.log 2>&1

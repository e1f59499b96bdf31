set -x
mkdir -p gpurun_out
for n in 8 32 64 128; do python bench.py --page-only --page-lines $n > gpurun_out/page_$n.json 2>> gpurun_out/page.err; done
cat gpurun_out/page_*.json
ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 1200 --csv --log-file gpurun_out/launches_page32.csv python bench.py --page-only --page-lines 32 > gpurun_out/ncu_p.log 2>&1
B="python bench.py --workload beam --only --steps 1 --warmup 3 --no-cpu"
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 600 -c 6 -o gpurun_out/prof_gemm_beam $B > gpurun_out/ncu_gb.log 2>&1

set -x
mkdir -p gpurun_out
P="python bench.py --page-only --page-lines 32"
COR_PD_NOCOOP=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:page_decoder_kernel -s 2 -c 1 -o gpurun_out/prof_pagedec $P > gpurun_out/ncu_pd.log 2>&1; tail -4 gpurun_out/ncu_pd.log
COR_PD_NOCOOP=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_p18_page32.csv $P > gpurun_out/ncu_pg.log 2>&1
F="python bench.py --workload greedy --only --lines 16384 --steps 1 --warmup 3 --no-cpu --page-lines 0 --predict-lines 0"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 20 -c 4 -o gpurun_out/prof_gemm_p18_encL1 $F > gpurun_out/ncu_f.log 2>&1; tail -3 gpurun_out/ncu_f.log

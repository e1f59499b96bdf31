set -x
mkdir -p gpurun_out
export COR_NO_OVERLAP=1
F="python bench.py --workload greedy --only --lines 16384 --steps 1 --warmup 3 --no-cpu"
$F > gpurun_out/plain_f.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 3300 -c 1100 --csv --log-file gpurun_out/launches_p1_greedy_16384.csv $F > gpurun_out/ncu_g.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 1500 -c 8 -o gpurun_out/prof_gemm_fp16 $F > gpurun_out/ncu_f.log 2>&1
ls -la gpurun_out

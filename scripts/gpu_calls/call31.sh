set -x
mkdir -p gpurun_out
B="python bench.py --workload beam --only --steps 1 --warmup 3 --no-cpu --page-lines 0 --predict-lines 0"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:beam_expand_warp_kernel -s 150 -c 2 -o gpurun_out/prof_expand_p31 $B > gpurun_out/ncu_e.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:beam_gather_kernel -s 150 -c 2 -o gpurun_out/prof_gather_p31 $B > gpurun_out/ncu_ga.log 2>&1
ls -la gpurun_out/*p31*

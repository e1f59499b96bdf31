set -x
mkdir -p gpurun_out
B="python bench.py --workload beam --only --steps 1 --warmup 3 --no-cpu"
F="python bench.py --workload greedy --only --lines 16384 --steps 1 --warmup 3 --no-cpu"
$B > gpurun_out/plain_b.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel.*LstmEpi -s 600 -c 4 -o gpurun_out/prof_gemm_beam $B > gpurun_out/ncu_gb.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:attention_tma -s 100 -c 3 -o gpurun_out/prof_attn_p9 $F > gpurun_out/ncu_a.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:beam_expand_warp -s 50 -c 2 -o gpurun_out/prof_expand $B > gpurun_out/ncu_e.log 2>&1
ls -la gpurun_out

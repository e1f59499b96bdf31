set -x
mkdir -p gpurun_out
G="python bench.py --workload greedy --only --lines 4096 --steps 1 --warmup 3 --no-cpu"
B="python bench.py --workload beam --only --beam-lines 2048 --steps 1 --warmup 3 --no-cpu"
F="python bench.py --workload greedy --only --lines 16384 --steps 1 --warmup 3 --no-cpu"
$G > gpurun_out/plain_g.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 4300 -c 1500 --csv --log-file gpurun_out/launches_fp16_greedy_4096.csv $G > gpurun_out/ncu_g.log 2>&1
$B > gpurun_out/plain_b.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 3900 -c 1400 --csv --log-file gpurun_out/launches_fp16_beam_2048.csv $B > gpurun_out/ncu_b.log 2>&1
$F > gpurun_out/plain_f.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel.*LstmEpi -s 900 -c 3 -o gpurun_out/prof_lstm_fp16 $F > gpurun_out/ncu_f.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:attention_kernel -s 200 -c 2 -o gpurun_out/prof_attn $F > gpurun_out/ncu_a.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel.*StoreEpi -s 600 -c 4 -o gpurun_out/prof_store $F > gpurun_out/ncu_s.log 2>&1
ls -la gpurun_out

set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/gpu_tests.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1 | tee gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_p23.json 2> gpurun_out/bench_p23.err; cut -c1-160 gpurun_out/bench_p23.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_p23.json 2> gpurun_out/bench_ref_p23.err; cut -c1-120 gpurun_out/bench_ref_p23.json
F="python bench.py --workload greedy --only --lines 16384 --steps 1 --warmup 3 --no-cpu --page-lines 0 --predict-lines 0"
B="python bench.py --workload beam --only --steps 1 --warmup 3 --no-cpu --page-lines 0 --predict-lines 0"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 3300 -c 1400 --csv --log-file gpurun_out/launches_p23_greedy_16384.csv $F > gpurun_out/ncu_g.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 3400 -c 1400 --csv --log-file gpurun_out/launches_p23_beam_9472.csv $B > gpurun_out/ncu_b.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attention_tma_kernel -s 300 -c 3 -o gpurun_out/prof_attn_p23 $F > gpurun_out/ncu_a.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 2000 -c 6 -o gpurun_out/prof_gemm_p23 $F > gpurun_out/ncu_f.log 2>&1
ls gpurun_out | grep p23

set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/gpu_tests.log
timeout 900 python bench.py > gpurun_out/bench_p18.json 2> gpurun_out/bench_p18.err; cut -c1-200 gpurun_out/bench_p18.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_p18.json 2> gpurun_out/bench_ref_p18.err; cut -c1-300 gpurun_out/bench_ref_p18.json
rm -f gpurun_out/pd_timing.jsonl
for n in 8 32; do COR_PD_TIMING=gpurun_out/pd_timing.jsonl timeout 300 python bench.py --page-only --page-lines $n > gpurun_out/page8_$n.json 2>> gpurun_out/page8.err; done
tail -2 gpurun_out/pd_timing.jsonl
P="python bench.py --page-only --page-lines 32"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/launches_p18_page32.csv $P > gpurun_out/ncu_pg.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:page_decoder_kernel -s 2 -c 1 -o gpurun_out/prof_pagedec $P > gpurun_out/ncu_pd.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:lstm_recur_kernel -s 4 -c 2 -o gpurun_out/prof_recur $P > gpurun_out/ncu_rc.log 2>&1
F="python bench.py --workload greedy --only --lines 16384 --steps 1 --warmup 3 --no-cpu --page-lines 0 --predict-lines 0"
B="python bench.py --workload beam --only --steps 1 --warmup 3 --no-cpu --page-lines 0 --predict-lines 0"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 3300 -c 1400 --csv --log-file gpurun_out/launches_p18_greedy_16384.csv $F > gpurun_out/ncu_g.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 3400 -c 1400 --csv --log-file gpurun_out/launches_p18_beam_9472.csv $B > gpurun_out/ncu_b.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel.*LstmEpi -s 40 -c 3 -o gpurun_out/prof_gemm_p18_encL1 $F > gpurun_out/ncu_f.log 2>&1
ls -la gpurun_out | tail -15

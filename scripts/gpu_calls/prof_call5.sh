set -x
mkdir -p gpurun_out
nproc
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/gpu_tests.log
python bench.py > gpurun_out/bench_p9.json 2> gpurun_out/bench_p9.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
B="python bench.py --workload beam --only --steps 1 --warmup 3 --no-cpu"
F="python bench.py --workload greedy --only --lines 16384 --steps 1 --warmup 3 --no-cpu"
ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 1400 --csv --log-file gpurun_out/launches_p9_beam_9472.csv $B > gpurun_out/ncu_b.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 1400 --csv --log-file gpurun_out/launches_p9_greedy_16384.csv $F > gpurun_out/ncu_g.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 900 -c 6 -o gpurun_out/prof_gemm_p9 $F > gpurun_out/ncu_f.log 2>&1
ls -la gpurun_out

set -x
F="python bench.py --workload greedy --only --lines 16384 --steps 1 --warmup 3 --no-cpu"
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python bench.py --no-cpu > gpurun_out/bench_p8.json 2> gpurun_out/bench_p8.err && ncu --set full --clock-control none --import-source on -k regex:"attention" -s 300 -c 2 -o gpurun_out/prof_attn8 $F > gpurun_out/ncu_a8.log 2>&1

# round 2, call 1: whole GPU suite, default bench, reference arm, launch lists of beam + greedy steps
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv,noheader
( time timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 ) 2>&1 | tee gpurun_out/r02_c01_gpu_tests.log
( time timeout 1200 python bench.py > gpurun_out/r02_c01_bench.json 2> gpurun_out/r02_c01_bench.err ) 2>&1 | tail -4
cut -c1-1500 gpurun_out/r02_c01_bench.json; tail -5 gpurun_out/r02_c01_bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r02_c01_launches_beam.csv python bench.py --only --workload beam --steps 1 --warmup 3 --no-cpu --no-parity --strong-greedy-lines 0 --strong-beam-lines 0 > gpurun_out/r02_c01_ncu_beam.log 2>&1
tail -2 gpurun_out/r02_c01_ncu_beam.log | cut -c1-300

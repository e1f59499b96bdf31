set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/gpu_tests.log
timeout 900 python bench.py --no-cpu --page-lines 0 --predict-lines 0 > gpurun_out/bench_p14.json 2> gpurun_out/bench_p14.err; cat gpurun_out/bench_p14.json | cut -c1-300

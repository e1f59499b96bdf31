set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/gpu_tests.log
timeout 600 python scripts/measure_mode3.py > gpurun_out/mode3.log 2>&1; tail -2 gpurun_out/mode3.log
timeout 900 python bench.py > gpurun_out/bench_head.json 2> gpurun_out/bench_head.err; cat gpurun_out/bench_head.json | cut -c1-600
timeout 600 python bench.py --gemm-mode 3 --no-cpu --page-lines 0 --predict-lines 0 > gpurun_out/bench_mode3.json 2> gpurun_out/bench_mode3.err; cat gpurun_out/bench_mode3.json | cut -c1-400

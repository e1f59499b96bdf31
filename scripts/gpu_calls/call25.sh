set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/gpu_tests.log
timeout 900 python bench.py > gpurun_out/bench_p25.json 2> gpurun_out/bench_p25.err; cut -c1-160 gpurun_out/bench_p25.json

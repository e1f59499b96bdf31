set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/gpu_tests.log
timeout 900 python bench.py --no-cpu --page-lines 0 --predict-lines 0 > gpurun_out/bench_p37.json 2> gpurun_out/bench_p37.err; cut -c1-160 gpurun_out/bench_p37.json
timeout 900 python bench.py --no-cpu --page-lines 0 --predict-lines 0 > gpurun_out/bench_p37b.json 2> gpurun_out/bench_p37.err; cut -c1-160 gpurun_out/bench_p37b.json

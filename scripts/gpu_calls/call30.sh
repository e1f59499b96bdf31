set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_core.py -m gpu -x -q -s -k "page_path_agrees" 2>&1 | tail -5 | tee gpurun_out/gpu_tests_c.log
timeout 600 python bench.py --gemm-mode 3 --no-cpu --page-lines 0 --predict-lines 0 > gpurun_out/bench_mode3_p30.json 2> gpurun_out/bench_mode3_p30.err; cut -c1-160 gpurun_out/bench_mode3_p30.json

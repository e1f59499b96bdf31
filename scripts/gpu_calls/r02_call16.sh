# round 2, call 16: three-stage predict pipeline; whole GPU suite on the head state
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 ) 2>&1 | tee gpurun_out/r02_c16_gpu_tests.log
timeout 600 python bench.py --predict-only > gpurun_out/r02_c16_predict.json 2> gpurun_out/r02_c16.err; cat gpurun_out/r02_c16_predict.json | cut -c1-600; tail -3 gpurun_out/r02_c16.err

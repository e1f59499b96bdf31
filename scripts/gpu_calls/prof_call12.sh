set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/gpu_tests.log
python bench.py --predict-only > gpurun_out/predict1.json 2> gpurun_out/predict1.err
cat gpurun_out/predict1.json
tail -3 gpurun_out/predict1.err

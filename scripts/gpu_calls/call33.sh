mkdir -p gpurun_out
timeout 600 python bench.py --predict-only --predict-lines 8192 > gpurun_out/predict_p33.json 2> gpurun_out/predict_p33.err; cat gpurun_out/predict_p33.json; tail -3 gpurun_out/predict_p33.err

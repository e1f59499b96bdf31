# round 2, call 10: training step parity (loss, gradients, Adam) against the torch float64 oracle
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_train.py -m gpu -x -q -s 2>&1 | tail -25 ) | tee gpurun_out/r02_c10_train_tests.log
( timeout 300 python -m pytest tests/test_gpu_core.py -m gpu -x -q -s -k "variant_k" 2>&1 | tail -5 ) | tee gpurun_out/r02_c10_variant_k.log

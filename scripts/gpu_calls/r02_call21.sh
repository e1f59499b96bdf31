# round 2, call 21: training with block-per-row attention kernels, parallel regulariser, decoder beside encoder layers; whole suite + smoke
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_train.py -m gpu -x -q -s 2>&1 | grep -E "worst|passed|failed|Error|val_loss" | tail -12 ) | tee gpurun_out/r02_c21_train_tests.log
timeout 900 python bench.py --workload train --steps 10 --warmup 3 > gpurun_out/r02_c21_train.json 2> gpurun_out/r02_c21_train.err; cut -c1-330 gpurun_out/r02_c21_train.json; tail -3 gpurun_out/r02_c21_train.err
( time timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 ) 2>&1 | tee gpurun_out/r02_c21_gpu_tests.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3

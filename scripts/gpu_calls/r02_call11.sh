# round 2, call 11: training: graph-recorded step parity, the train() loop, the C5 bench row
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_train.py -m gpu -x -q -s 2>&1 | tail -25 ) | tee gpurun_out/r02_c11_train_tests.log
( time timeout 900 python bench.py --workload train --steps 10 --warmup 3 > gpurun_out/r02_c11_train.json 2> gpurun_out/r02_c11_train.err ) 2>&1 | tail -3
cut -c1-1800 gpurun_out/r02_c11_train.json; tail -5 gpurun_out/r02_c11_train.err
COR_NO_GRAPH=1 timeout 900 python bench.py --workload train --steps 10 --warmup 3 --no-cpu > gpurun_out/r02_c11_train_nograph.json 2>> gpurun_out/r02_c11_train.err; cut -c1-300 gpurun_out/r02_c11_train_nograph.json

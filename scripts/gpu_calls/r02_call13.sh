# round 2, call 13 (2 GPUs): the bench as the driver launches it for N = 2, the training row over NCCL, the data-parallel parity test
mkdir -p gpurun_out
nvidia-smi -L
( timeout 600 python -m pytest tests/test_gpu_train.py -m gpu -x -q -s -k "data_parallel" 2>&1 | tail -6 ) | tee gpurun_out/r02_c13_dp_test.log
( time timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r02_c13_bench_n2.json 2> gpurun_out/r02_c13_bench_n2.err ) 2>&1 | tail -3
cut -c1-400 gpurun_out/r02_c13_bench_n2.json; tail -4 gpurun_out/r02_c13_bench_n2.err
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 2 --workload train --steps 10 --warmup 3 > gpurun_out/r02_c13_train_n2.json 2> gpurun_out/r02_c13_train_n2.err ) 2>&1 | tail -3
cut -c1-400 gpurun_out/r02_c13_train_n2.json; tail -4 gpurun_out/r02_c13_train_n2.err
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29543 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r02_c13_ref_n2.json 2> gpurun_out/r02_c13_ref_n2.err ) 2>&1 | tail -3
cut -c1-300 gpurun_out/r02_c13_ref_n2.json

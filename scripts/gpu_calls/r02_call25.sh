# round 2, call 25 (8 GPUs): the bench as the driver launches it for N = 8 (weak `value` + the strong fixed jobs), the training row over NCCL
mkdir -p gpurun_out
nvidia-smi -L | wc -l
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r02_bench_final_n8.json 2> gpurun_out/r02_bench_final_n8.err ) 2>&1 | grep real
cut -c1-300 gpurun_out/r02_bench_final_n8.json; tail -3 gpurun_out/r02_bench_final_n8.err
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29552 bench.py --gpus 8 --workload train --steps 20 --warmup 3 > gpurun_out/r02_bench_final_train_n8.json 2> gpurun_out/r02_bench_final_train_n8.err ) 2>&1 | grep real
cut -c1-300 gpurun_out/r02_bench_final_train_n8.json; tail -3 gpurun_out/r02_bench_final_train_n8.err

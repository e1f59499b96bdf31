mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/bench_p29_n8.json 2> gpurun_out/bench_p29_n8.err; cut -c1-200 gpurun_out/bench_p29_n8.json; tail -2 gpurun_out/bench_p29_n8.err

# round 2, call 44 (2 GPUs): the head state on two ranks -- the data-parallel training test, then the default bench under torchrun
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_train.py -m gpu -x -q -k "two_gpus or data_parallel" 2>&1 | tail -3
( time timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r02_bench_head_n2.json 2> gpurun_out/r02_bench_head_n2.err ) 2>&1 | grep real
python -c "
import json;d=json.load(open('gpurun_out/r02_bench_head_n2.json'))
print('beam', round(d['value']), round(d['e2e']['value']), d['ms_per_step'], d['n_gpus'])
g=d['also']['greedy']; print('greedy', round(g['value']), g['ms_per_step'])
print({k:(v['seconds'], round(v['chars_per_s'])) for k,v in d.get('strong',{}).items() if isinstance(v,dict)})
print('train', d['also'].get('train_c5',{}).get('ms_per_step'))"
tail -2 gpurun_out/r02_bench_head_n2.err

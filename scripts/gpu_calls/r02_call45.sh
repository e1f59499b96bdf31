# round 2, call 45 (8 GPUs): default bench at head under torchrun
mkdir -p gpurun_out
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29581 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r02_bench_head_n8.json 2> gpurun_out/r02_bench_head_n8.err ) 2>&1 | grep real
python -c "
import json;d=json.load(open('gpurun_out/r02_bench_head_n8.json'))
print('beam', round(d['value']), round(d['e2e']['value']), d['ms_per_step'], d['n_gpus'])
g=d['also']['greedy']; print('greedy', round(g['value']), g['ms_per_step'])
print({k:(v['seconds'], round(v['chars_per_s'])) for k,v in d.get('strong',{}).items() if isinstance(v,dict)})"
tail -2 gpurun_out/r02_bench_head_n8.err

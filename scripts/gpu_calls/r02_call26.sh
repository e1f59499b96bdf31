# round 2, call 26 (8 GPUs): default bench again -- strong legs warmed up through the collectives, int16 / pinned result return
mkdir -p gpurun_out
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r02_bench_final_n8_b.json 2> gpurun_out/r02_bench_final_n8_b.err ) 2>&1 | grep real
python -c "
import json;d=json.load(open('gpurun_out/r02_bench_final_n8_b.json'))
print(round(d['value']), round(d['e2e']['value']), d['ms_each_step']); print(round(d['also']['greedy']['value']), d['also']['greedy']['ms_each_step'])
print({k:(v['seconds'], round(v['chars_per_s'])) for k,v in d['strong'].items() if isinstance(v,dict)})"
tail -3 gpurun_out/r02_bench_final_n8_b.err

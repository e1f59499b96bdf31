# round 2, call 35: the head state end to end -- whole GPU suite, smoke, default bench, reference arm, training workload,
# and the competitive bracket's launch list at head (host loop: ncu cannot replay conditional graphs per kernel)
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 ) 2>&1 | tee gpurun_out/r02_head_gpu_tests.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2 | tee gpurun_out/r02_head_smoke.log
( time timeout 1500 python bench.py > gpurun_out/r02_bench_head_n1.json 2> gpurun_out/r02_bench_head_n1.err ) 2>&1 | grep real
python -c "
import json;d=json.load(open('gpurun_out/r02_bench_head_n1.json'))
print('beam', round(d['value']), round(d['e2e']['value']), d['ms_per_step'], d['clocks'], d['roofline'])
g=d['also']['greedy']; print('greedy', round(g['value']), round(g['e2e']['value']), g['ms_per_step'], g['clocks'])
print('train', d['also']['train_c5']['ms_per_step'], 'page', d['page']['ms_per_call'], d['page']['beam']['ms_per_call'], 'predict', round(d['predict']['chars_per_s']))
print('co_primary', json.dumps(d.get('co_primary'))[:600])
print({k: (round(v['value']) if isinstance(v, dict) and 'value' in v else None) for k, v in d['also'].items()})
"
( time timeout 900 python bench.py --impl reference > gpurun_out/r02_bench_head_ref.json 2> gpurun_out/r02_bench_head_ref.err ) 2>&1 | grep real
cut -c1-300 gpurun_out/r02_bench_head_ref.json
timeout 900 python bench.py --workload train --steps 10 --warmup 3 > gpurun_out/r02_bench_head_train_n1.json 2> gpurun_out/r02_bench_head_train_n1.err; cut -c1-260 gpurun_out/r02_bench_head_train_n1.json
COR_NO_GRAPH=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file gpurun_out/r02_launches_comp256_head.csv python scripts/beam_probe.py beam_competitive 256 1 > gpurun_out/ncu_comp256_head.log 2>&1
tail -1 gpurun_out/ncu_comp256_head.log | cut -c1-200

# round 2, call 48: FINAL head state (three issuers + second TMA producer in the page GEMM): whole GPU suite, smoke, default bench, reference arm, training workload
# with the graph's K-segments at step 0: whole GPU suite, smoke, default bench, reference arm, training workload
mkdir -p gpurun_out
( time timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 ) 2>&1 | tee gpurun_out/r02_head_gpu_tests.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2 | tee gpurun_out/r02_head_smoke.log
( time timeout 1500 python bench.py > gpurun_out/r02_bench_head_n1.json 2> gpurun_out/r02_bench_head_n1.err ) 2>&1 | grep real
python -c "
import json;d=json.load(open('gpurun_out/r02_bench_head_n1.json'))
print('beam', round(d['value']), round(d['e2e']['value']), d['ms_per_step'], d['clocks'], d['roofline']['frac'], d['parity'])
g=d['also']['greedy']; print('greedy', round(g['value']), round(g['e2e']['value']), g['ms_per_step'], g['clocks'])
print('train', d['also']['train_c5']['ms_per_step'], 'page', d['page']['ms_per_call'], d['page']['beam']['ms_per_call'], 'predict', round(d['predict']['chars_per_s']))
c=d['co_primary']; print('co_primary', round(c['value']), round(c['e2e']), c['parity']['rows_identical'], c['parity']['rows_differing_at_logged_near_tie'], c['parity']['rows_differing_otherwise'])
print({k: (round(v['value']) if isinstance(v, dict) and 'value' in v else None) for k, v in d['also'].items()})
"
( time timeout 900 python bench.py --impl reference > gpurun_out/r02_bench_head_ref.json 2> gpurun_out/r02_bench_head_ref.err ) 2>&1 | grep real
cut -c1-200 gpurun_out/r02_bench_head_ref.json
timeout 900 python bench.py --workload train --steps 10 --warmup 3 > gpurun_out/r02_bench_head_train_n1.json 2> gpurun_out/r02_bench_head_train_n1.err; cut -c1-200 gpurun_out/r02_bench_head_train_n1.json

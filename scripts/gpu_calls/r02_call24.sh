# round 2, call 24: default bench again (clock sampler started before the warm-up), twice
mkdir -p gpurun_out
for k in a b; do
( time timeout 1500 python bench.py > gpurun_out/r02_bench_final_n1_$k.json 2> gpurun_out/r02_bench_final_n1_$k.err ) 2>&1 | grep real
python -c "
import json;d=json.load(open('gpurun_out/r02_bench_final_n1_$k.json'))
print('beam', round(d['value']), round(d['e2e']['value']), d['ms_per_step'], d['clocks'])
g=d['also']['greedy']; print('greedy', round(g['value']), round(g['e2e']['value']), g['ms_per_step'], g['clocks'])
print('train', d['also']['train_c5']['ms_per_step'], 'page', d['page']['ms_per_call'], d['page']['beam']['ms_per_call'], 'predict', round(d['predict']['chars_per_s']))
"
done

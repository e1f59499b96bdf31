# round 2, call 9: variant-K diagnostics; the full default bench (every leg) and the reference arm, as the driver runs them
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_core.py -m gpu -x -q -s -k "variant_k" 2>&1 | tail -12 ) | tee gpurun_out/r02_c09_variant_k.log
( time timeout 1500 python bench.py > gpurun_out/r02_c09_bench.json 2> gpurun_out/r02_c09_bench.err ) 2>&1 | tail -4
cut -c1-600 gpurun_out/r02_c09_bench.json; tail -5 gpurun_out/r02_c09_bench.err
( time timeout 900 python bench.py --impl reference > gpurun_out/r02_c09_bench_ref.json 2> gpurun_out/r02_c09_bench_ref.err ) 2>&1 | tail -4
cut -c1-600 gpurun_out/r02_c09_bench_ref.json

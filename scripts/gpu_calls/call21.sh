set -x
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_p21.json 2> gpurun_out/bench_p21.err; cut -c1-160 gpurun_out/bench_p21.json; tail -2 gpurun_out/bench_p21.err

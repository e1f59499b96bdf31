set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_core.py tests/test_gpu_seq2seq.py -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/gpu_tests_b.log
rm -f gpurun_out/pd_timing3.jsonl
for n in 8 32 64; do COR_PD_TIMING=gpurun_out/pd_timing3.jsonl timeout 300 python bench.py --page-only --page-lines $n > gpurun_out/page11_$n.json 2>> gpurun_out/page11.err; cut -c1-120 gpurun_out/page11_$n.json; done
tail -1 gpurun_out/pd_timing3.jsonl

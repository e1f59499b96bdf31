set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_core.py tests/test_gpu_seq2seq.py tests/test_gpu_beam.py -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/gpu_tests_b.log
for n in 32 48 64; do timeout 300 python bench.py --page-only --page-lines $n > gpurun_out/page10_$n.json 2>> gpurun_out/page10.err; cut -c1-330 gpurun_out/page10_$n.json; done
COR_NO_PAGEDEC=1 COR_NO_RECUR=1 timeout 300 python bench.py --page-only --page-lines 64 > gpurun_out/page10_64_old.json 2>> gpurun_out/page10.err; cut -c1-200 gpurun_out/page10_64_old.json

# round 2, call 28: cor_gemm_noalloc -- its test alone in a fresh process, then after the other GEMM tests in one process;
# then the competitive bracket's launch list at head (256 lines, host loop: ncu cannot replay conditional graphs per kernel)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q -s 2>&1 | tail -12
timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_alignpath.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -4
timeout 300 python scripts/beam_probe.py beam_competitive 256 2 2>&1 | tail -1
COR_NO_GRAPH=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r02_launches_comp256.csv python scripts/beam_probe.py beam_competitive 256 1 > gpurun_out/ncu_comp256.log 2>&1
tail -2 gpurun_out/ncu_comp256.log

# round 2, call 39: the page-sized GEMM (128x32 tiles, 32 rows, N = 2048, K = 1536 and 512) under ncu --set full, warm caches
mkdir -p gpurun_out
timeout 600 ncu --set full --cache-control none --clock-control none --import-source on -k regex:gemm_tc_kernel -s 2 -c 2 -o gpurun_out/r02_ncu_gemm_narrow_k1536 -f python scripts/gemm_probe.py 3 32 2048 1536 > gpurun_out/ncu_gemm_narrow.log 2>&1; tail -1 gpurun_out/ncu_gemm_narrow.log
timeout 600 ncu --metrics gpu__time_duration.sum --cache-control none --clock-control none -k regex:gemm_tc_kernel --csv python scripts/gemm_probe.py 3 32 2048 512 2>&1 | grep gemm_tc | cut -d, -f5,12- | tail -3
timeout 600 ncu --metrics gpu__time_duration.sum --cache-control none --clock-control none -k regex:gemm_tc_kernel --csv python scripts/gemm_probe.py 3 32 2560 512 2>&1 | grep gemm_tc | cut -d, -f5,12- | tail -2
timeout 600 ncu --metrics gpu__time_duration.sum --cache-control none --clock-control none -k regex:gemm_tc_kernel --csv python scripts/gemm_probe.py 3 32 2048 1536 2>&1 | grep gemm_tc | cut -d, -f5,12- | tail -3

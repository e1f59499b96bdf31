# round 2, call 49: ncu --set full of the 128x32-tile GEMM at head (three issuers, two producers), warm caches: the "after" of r02_ncu_full_gemm_narrow_before_raw.csv
mkdir -p gpurun_out
timeout 300 ncu --set full --cache-control none --clock-control none --import-source on -k regex:gemm_tc_kernel -s 2 -c 2 -o gpurun_out/r02_ncu_gemm_narrow_after -f python scripts/gemm_probe.py 3 32 2048 1536 > gpurun_out/ncu_gemm_narrow_after.log 2>&1; tail -1 gpurun_out/ncu_gemm_narrow_after.log

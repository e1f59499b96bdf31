# round 2, call 12: training (parallel directions, one-GEMM dR) re-verified; profile captures of the head state
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_train.py -m gpu -x -q 2>&1 | tail -4 ) | tee gpurun_out/r02_c12_train_tests.log
timeout 900 python bench.py --workload train --steps 10 --warmup 3 > gpurun_out/r02_c12_train.json 2> gpurun_out/r02_c12_train.err; cut -c1-330 gpurun_out/r02_c12_train.json; tail -3 gpurun_out/r02_c12_train.err
export COR_NO_GRAPH=1      # ncu replays kernels one by one: the host-driven loops issue the same kernels as the graphs
# launch lists (cold-cache, serialised: shares only)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r02_launches_beam_9472lines.csv python scripts/beam_probe.py beam 9472 3 40 > gpurun_out/r02_ncu_a.log 2>&1; tail -1 gpurun_out/r02_ncu_a.log | cut -c1-200
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r02_launches_page32_beam.csv python scripts/page_beam_probe.py 32 1 > gpurun_out/r02_ncu_b.log 2>&1; tail -1 gpurun_out/r02_ncu_b.log | cut -c1-200
# full captures: the LSTM GEMM on CTA pairs in the beam step (K = 512 and the K >= 1024 launches), the heavy push, the page select
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:LstmEpi -s 200 -c 6 -o gpurun_out/r02_full_gemm_lstm python scripts/beam_probe.py beam 9472 1 40 > gpurun_out/r02_ncu_c.log 2>&1; tail -1 gpurun_out/r02_ncu_c.log | cut -c1-200
timeout 900 ncu --set full --clock-control none --import-source on -k regex:beam_push_kernel -s 300 -c 3 -o gpurun_out/r02_full_push python scripts/beam_probe.py beam_competitive 256 1 > gpurun_out/r02_ncu_d.log 2>&1; tail -1 gpurun_out/r02_ncu_d.log | cut -c1-200
timeout 900 ncu --set full --clock-control none --import-source on -k regex:beam_pop_kernel -s 100 -c 3 -o gpurun_out/r02_full_pop python scripts/beam_probe.py beam_competitive 256 1 > gpurun_out/r02_ncu_e.log 2>&1; tail -1 gpurun_out/r02_ncu_e.log | cut -c1-200
timeout 900 ncu --set full --clock-control none --import-source on -k regex:beam_expand_warp_kernel -s 100 -c 3 -o gpurun_out/r02_full_expand python scripts/beam_probe.py beam 9472 1 40 > gpurun_out/r02_ncu_f.log 2>&1; tail -1 gpurun_out/r02_ncu_f.log | cut -c1-200
ls -la gpurun_out/*.ncu-rep

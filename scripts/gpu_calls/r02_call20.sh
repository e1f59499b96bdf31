# round 2, call 20: captures of the head state: K = 512 LSTM launches with the TMA-store epilogue (and with direct stores),
# launch lists of the greedy step and of one training step
mkdir -p gpurun_out
export COR_NO_GRAPH=1
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:LstmEpi -s 20 -c 4 -o gpurun_out/r02_full_gemm_lstm_k512_tma python scripts/beam_probe.py beam 9472 1 40 > gpurun_out/r02_ncu_g.log 2>&1; tail -1 gpurun_out/r02_ncu_g.log | cut -c1-200
COR_NO_EPI_TMA=1 timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:LstmEpi -s 20 -c 4 -o gpurun_out/r02_full_gemm_lstm_k512_direct python scripts/beam_probe.py beam 9472 1 40 > gpurun_out/r02_ncu_h.log 2>&1; tail -1 gpurun_out/r02_ncu_h.log | cut -c1-200
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/r02_launches_greedy_16384lines.csv python bench.py --only --workload greedy --steps 1 --warmup 3 --no-cpu --no-parity --strong-greedy-lines 0 --strong-beam-lines 0 > gpurun_out/r02_ncu_i.log 2>&1; tail -1 gpurun_out/r02_ncu_i.log | cut -c1-200
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 9000 -c 5000 --csv --log-file gpurun_out/r02_launches_train_step.csv python bench.py --workload train --steps 1 --warmup 3 --no-cpu > gpurun_out/r02_ncu_j.log 2>&1; tail -1 gpurun_out/r02_ncu_j.log | cut -c1-200
ls -la gpurun_out/*.csv gpurun_out/*k512*

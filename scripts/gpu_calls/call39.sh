set -x
mkdir -p gpurun_out
F="python bench.py --workload greedy --only --lines 16384 --steps 1 --warmup 3 --no-cpu --page-lines 0 --predict-lines 0"
B="python bench.py --workload beam --only --steps 1 --warmup 3 --no-cpu --page-lines 0 --predict-lines 0"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 3300 -c 1400 --csv --log-file gpurun_out/launches_p39_greedy_16384.csv $F > gpurun_out/ncu_g.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 3400 -c 1400 --csv --log-file gpurun_out/launches_p39_beam_9472.csv $B > gpurun_out/ncu_b.log 2>&1
P="python bench.py --page-only --page-lines 32"
COR_PD_NOCOOP=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_p39_page32.csv $P > gpurun_out/ncu_pg.log 2>&1
ls gpurun_out | grep p39

set -x
mkdir -p gpurun_out
B="python bench.py --workload beam --only --beam-lines 8192 --steps 1 --warmup 3 --no-cpu"
$B > gpurun_out/plain_b.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 3400 -c 1200 --csv --log-file gpurun_out/launches_p8_beam_8192.csv $B > gpurun_out/ncu_b.log 2>&1

# round 2, call 22: the default bench and the reference arm as the driver runs them (N = 1), the training row, a launch list of the training step
mkdir -p gpurun_out
( time timeout 1500 python bench.py > gpurun_out/r02_bench_final_n1.json 2> gpurun_out/r02_bench_final_n1.err ) 2>&1 | tail -3
cut -c1-300 gpurun_out/r02_bench_final_n1.json; tail -3 gpurun_out/r02_bench_final_n1.err
( time timeout 900 python bench.py --impl reference > gpurun_out/r02_bench_final_ref.json 2> gpurun_out/r02_bench_final_ref.err ) 2>&1 | tail -3
cut -c1-200 gpurun_out/r02_bench_final_ref.json
timeout 900 python bench.py --workload train --steps 20 --warmup 3 > gpurun_out/r02_bench_final_train_n1.json 2> gpurun_out/r02_c22.err; cut -c1-300 gpurun_out/r02_bench_final_train_n1.json
timeout 600 python bench.py --impl reference --workload train --steps 2 --warmup 1 > gpurun_out/r02_bench_final_train_ref.json 2>> gpurun_out/r02_c22.err; cut -c1-300 gpurun_out/r02_bench_final_train_ref.json
COR_NO_GRAPH=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 9000 -c 5000 --csv --log-file gpurun_out/r02_launches_train_step.csv python bench.py --workload train --steps 1 --warmup 3 --no-cpu > gpurun_out/r02_ncu_k.log 2>&1; tail -1 gpurun_out/r02_ncu_k.log | cut -c1-200

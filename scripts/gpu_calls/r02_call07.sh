# round 2, call 7: where the competitive / stress beam spends its time (launch lists), graph vs host loop
mkdir -p gpurun_out
python scripts/beam_probe.py beam_competitive 512 2 | tee gpurun_out/r02_c07_comp512.json
COR_NO_GRAPH=1 python scripts/beam_probe.py beam_competitive 512 2 | tee gpurun_out/r02_c07_comp512_hostloop.json
python scripts/beam_probe.py beam_stress 64 2 | tee gpurun_out/r02_c07_stress64.json
COR_NO_GRAPH=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 30000 --csv --log-file gpurun_out/r02_c07_launches_comp512.csv python scripts/beam_probe.py beam_competitive 512 1 > gpurun_out/r02_c07_ncu_comp.log 2>&1
tail -1 gpurun_out/r02_c07_ncu_comp.log | cut -c1-300
COR_NO_GRAPH=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 30000 --csv --log-file gpurun_out/r02_c07_launches_stress64.csv python scripts/beam_probe.py beam_stress 64 1 > gpurun_out/r02_c07_ncu_stress.log 2>&1
tail -1 gpurun_out/r02_c07_ncu_stress.log | cut -c1-300

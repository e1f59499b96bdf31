# round 2, call 38: WARM per-kernel durations of one beamed page (ncu --cache-control none: caches as the previous kernels
# left them; host loop with the page's tile shapes forced by COR_NO_GRAPH=1 ... which uses pick_shape from M: narrow for 32 rows)
mkdir -p gpurun_out
COR_NO_GRAPH=1 timeout 300 python scripts/page_beam_probe.py 32 3 2>&1 | tail -1
timeout 300 python scripts/page_beam_probe.py 32 3 2>&1 | tail -1
COR_NO_GRAPH=1 timeout 600 ncu --metrics gpu__time_duration.sum --cache-control none --clock-control none -c 4000 --csv --log-file gpurun_out/r02_launches_page32_beam_warm.csv python scripts/page_beam_probe.py 32 1 > gpurun_out/ncu_page_warm.log 2>&1
tail -1 gpurun_out/ncu_page_warm.log

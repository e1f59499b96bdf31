# round 2, call 5: launch list of one beamed page (host loop: ncu sees every kernel), and the graph beside it
mkdir -p gpurun_out
python scripts/page_beam_probe.py 32 10 | tee gpurun_out/r02_c05_page_beam_graph.json
COR_NO_GRAPH=1 python scripts/page_beam_probe.py 32 10 | tee gpurun_out/r02_c05_page_beam_hostloop.json
COR_NO_GRAPH=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r02_c05_launches_page32_beam.csv python scripts/page_beam_probe.py 32 1 > gpurun_out/r02_c05_ncu_page.log 2>&1
tail -2 gpurun_out/r02_c05_ncu_page.log | cut -c1-300

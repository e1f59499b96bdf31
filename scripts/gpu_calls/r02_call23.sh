# round 2, call 23: greedy loop graph alone / after beam calls (72 GB and 8 GB workspaces)
mkdir -p gpurun_out
python scripts/greedy_after_beam_probe.py 0
python scripts/greedy_after_beam_probe.py 1 72
python scripts/greedy_after_beam_probe.py 1 8

# round 2, call 33: full-quota hint (a call whose lines nearly all needed the full record quota makes the next call start
# with full-quota waves): parity tests, then the stress bracket call by call (first call unhinted), the others as a check
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_beam.py tests/test_gpu_loop_graph.py tests/test_gpu_timed_shape.py tests/test_gpu_reference_pin.py -m gpu -x -q 2>&1 | tail -4
COR_BEAM_TRACE=1 timeout 300 python scripts/beam_probe.py beam_stress 256 3 72 2>&1 | tail -12
timeout 300 python scripts/beam_probe.py beam_stress 64 3 2>&1 | tail -1
timeout 300 python scripts/beam_probe.py beam_competitive 512 3 2>&1 | tail -1
timeout 300 python scripts/beam_probe.py beam 9472 4 72 2>&1 | tail -1

# round 2, call 29: reach truncation of the beam fringe (beam.cu reach_keep): parity tests, then A/B on the brackets whose fringe saturates
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_beam.py tests/test_gpu_loop_graph.py tests/test_gpu_timed_shape.py tests/test_gpu_reference_pin.py tests/test_gpu_seq2seq.py -m gpu -x -q 2>&1 | tail -6
for w in "beam_competitive 256" "beam_competitive 512" "beam_stress 64" "beam 9472"; do
  echo "== $w: full merge copy, then reach truncation"
  COR_NO_REACH_TRUNC=1 timeout 300 python scripts/beam_probe.py $w 2 2>&1 | tail -1
  timeout 300 python scripts/beam_probe.py $w 2 2>&1 | tail -1
done

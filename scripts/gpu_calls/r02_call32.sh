# round 2, call 32: wave structure of the competitive / stress brackets (COR_BEAM_TRACE=1), one call each
mkdir -p gpurun_out
COR_BEAM_TRACE=1 timeout 300 python scripts/beam_probe.py beam_competitive 256 1 2>&1 | tail -12
COR_BEAM_TRACE=1 timeout 300 python scripts/beam_probe.py beam_competitive 2048 1 72 2>&1 | tail -40
COR_BEAM_TRACE=1 timeout 300 python scripts/beam_probe.py beam_stress 256 1 72 2>&1 | tail -20

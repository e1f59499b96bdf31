# round 2, call 34: safety factor of the retry estimate (COR_RETRY_FACTOR), competitive bracket
mkdir -p gpurun_out
for f in 1.25 1.5 2.0 3.0; do
  echo "== factor $f"
  COR_RETRY_FACTOR=$f timeout 300 python scripts/beam_probe.py beam_competitive 2048 3 72 2>&1 | tail -1 | cut -c1-60,230-400
  COR_RETRY_FACTOR=$f timeout 300 python scripts/beam_probe.py beam_competitive 512 3 2>&1 | tail -1 | cut -c1-60,230-400
done
COR_RETRY_FACTOR=2.0 COR_BEAM_TRACE=1 timeout 300 python scripts/beam_probe.py beam_competitive 2048 1 72 2>&1 | tail -14 | cut -c1-200

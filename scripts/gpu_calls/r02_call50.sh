# round 2, call 50: the allocation-free entry point's test alone in a fresh process at the final head
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q -s 2>&1 | tail -6

python scripts/_dbg_hits.py 2>&1 | tail -40

"""print the seed+extend numbers of one bench.py JSON line read from stdin (profiling aid)"""
import json, sys
line = [l for l in sys.stdin.read().splitlines() if l.startswith("{")][-1]
d = json.loads(line).get("extra", {})
for k in ("map_c4", "pairs_c4", "map_c4_strong", "pairs_c4_strong"):
    if k in d:
        v = d[k]
        print(k, {kk: v[kk] for kk in v if isinstance(v[kk], (int, float, str)) and ("per_s" in kk or kk in ("value", "ms", "e2e", "mismatches", "unit"))})

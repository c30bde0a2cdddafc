"""print the full-matrix traceback leg of one bench.py JSON line read from stdin (profiling aid)"""
import json, sys
line = [l for l in sys.stdin.read().splitlines() if l.startswith("{")][-1]
for k, v in json.loads(line)["extra"]["full_traceback"].items():
    if isinstance(v, dict):
        print(k, round(v["pairs_per_s"] / 1e6, 2), "M pairs/s", round(v["ms"], 3), "ms", "paths", v["paths"], v.get("parity", ""))

#!/bin/bash
# round-end style check on the GPU box: GPU tests, smoke, both bench arms; prints a digest (profiling aid)
tag=${1:-final}
python -m pytest tests -m gpu -q -x 2>&1 | tail -1
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; tail -2 gpurun_out/${tag}_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2>&1
python - <<PY
import json
d=json.loads(open("gpurun_out/${tag}_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"]["value"], d["roofline"]["frac"], d["cpu_baseline"]["value"], d["parity"], d["clocks"])
e=d["extra"]
print("sw", e["sw"]["gcups"], e["sw"]["parity"], {k:round(v.get("gcups",0)) for k,v in e["banded"].items() if isinstance(v,dict)})
for k in ("map","map_c4"):
    m=e[k]; print(k, m.get("reads_per_s"), m.get("e2e",{}).get("reads_per_s"), m.get("parity"))
for k in ("pairs","pairs_c4"): print(k, e.get(k,{}).get("pairs_per_s"))
r=json.loads(open("gpurun_out/${tag}_bench_ref.json").read().strip().splitlines()[-1]); print("reference arm", r["value"], r["cpu_baseline"])
PY

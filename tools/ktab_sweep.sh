for k in 13 14 15; do KTAB_K=$k python bench.py --only c4 --steps 3 --reads 4000000 --pairs 0 --no-cpu 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); m=d[\"extra\"][\"map_c4\"]; print(\"k\", m.get(\"error\"), round(m.get(\"reads_per_s\",0)/1e6,1), round(m.get(\"e2e\",{}).get(\"reads_per_s\",0)/1e6,1), m.get(\"params\",\"\")[:90])"; done

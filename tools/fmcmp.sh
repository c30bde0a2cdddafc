for g in 1000000000 100000000; do python bench.py --only fm --genome $g --steps 5 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('genome $g', round(d['value']/1e9,3), 'ktab', round(d['config']['value_with_kmer_jump_table']['value']/1e9,3), 'locate', round(d['config']['locate']['value']/1e9,3), 'e2e', round(d['e2e']['value']/1e9,3))"; done

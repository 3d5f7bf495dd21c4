for s in 2 4 8 16; do
  GASMA_SLICES=$s python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-full-path 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('slices=$s', 'value', d['value'], 'e2e', d['e2e']['value'], d['e2e'].get('h2d_gb_per_s'))
"
done

set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for v in base v_unroll2 unroll2c lb6 lb4 k3step k3unroll; do
  echo "== $v"
  GASMA_B200_LIB=$PWD/approximate-string-matching_b200/variants/$v.so python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-full-path 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print('$v', d['value'], d['ms_per_step'], d['roofline'].get('achieved'), d.get('clocks',{}).get('sm_mhz'))
"
done

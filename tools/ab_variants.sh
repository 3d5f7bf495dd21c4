# A/B of prebuilt library variants (approximate-string-matching_b200/variants/*.so) on the greedy bench line
for v in "$@"; do
  for rep in 1 2; do
  GASMA_B200_LIB=$PWD/approximate-string-matching_b200/variants/$v.so python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-full-path --no-configs --no-parity 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print('$v', '%.4g' % d['value'], 'step %.4f ms' % d['ms_per_step'], 'kernel %.4f ms' % d['roofline']['kernel_ms'])
"
  done
done

# A/B of prebuilt library variants on the full-path leg (greedy + NW verifier + counters)
for v in "$@"; do
  for rep in 1 2; do
  GASMA_B200_LIB=$PWD/approximate-string-matching_b200/variants/$v.so python bench.py --steps 6 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l)['full_path']; print('$v', '%.4g' % d['value'], 'step %.4f ms' % d['ms_per_step'], 'nw %.4f ms' % d['nw_kernel_ms'], d['counters']['greedy_coverage'])
"
  done
done

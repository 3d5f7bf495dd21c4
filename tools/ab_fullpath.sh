# A/B of prebuilt library variants on the full-path leg (greedy + NW verifier + counters)
#   tools/ab_fullpath.sh _build/ab/a.so _build/ab/b.so ...
for rep in 1 2; do
for lib in "$@"; do
  GASMA_B200_LIB=$PWD/$lib python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-configs --no-parity 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l)['full_path']; print('$lib', '%.4g' % d['value'], 'step %.4f ms' % d['ms_per_step'], 'greedy %.4f ms' % d['greedy_kernel_ms'], 'nw %.4f ms' % d['nw_kernel_ms'], d['counters']['greedy_coverage'])
"
done
done

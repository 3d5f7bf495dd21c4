# kernel time of library builds on the greedy bench line, 3 runs x 30 steps each, interleaved
for rep in 1 2 3; do
for lib in "$@"; do
  GASMA_B200_LIB=$PWD/$lib python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-full-path --no-configs --no-parity 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$lib', 'value %.4g' % d['value'], 'step %.4f ms' % d['ms_per_step'], 'kernel %.4f ms' % d['roofline']['kernel_ms'])
"
done
done

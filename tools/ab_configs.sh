# A/B of library builds on the headline leg and the short config legs (kernel ms per leg)
for rep in 1 2; do
for lib in "$@"; do
  GASMA_B200_LIB=$PWD/$lib python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-full-path --no-parity 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$lib', 'head %.4f/%.4f' % (d['ms_per_step'], d['roofline']['kernel_ms']), ' '.join('%s %.4f/%.4f' % (k[:10], v['ms_per_step'], v['roofline']['kernel_ms']) for k, v in d['configs'].items()))
"
done
done

for rep in ${REPS:-1 2}; do
for lib in "$@"; do
  GASMA_B200_LIB=$PWD/$lib python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-parity 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); f=d['full_path']
        print('$lib', 'head %.4f/%.4f' % (d['ms_per_step'], d['roofline']['kernel_ms']), 'full %.4f greedy %.4f nw %.4f |' % (f['ms_per_step'], f['greedy_kernel_ms'], f['nw_kernel_ms']), ' '.join('%s %.4f/%.4f' % (k[:10], v['ms_per_step'], v['roofline']['kernel_ms']) for k, v in d['configs'].items()))
"
done
done

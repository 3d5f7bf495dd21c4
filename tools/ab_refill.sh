# sweep of the refill threshold (idle threads per warp that trigger a hand-out) on every bench leg
for rep in 1 2; do
for th in ${THRESHOLDS:-4 6 8 10 12}; do
  GASMA_REFILL_THRESHOLD=$th python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-parity --in-flight 1 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); f=d['full_path']
        print('threshold=$th', 'head %.4f/%.4f' % (d['ms_per_step'], d['roofline']['kernel_ms']), 'full %.4f greedy %.4f |' % (f['ms_per_step'], f['greedy_kernel_ms']), ' '.join('%s %.4f' % (k[:10], v['roofline']['kernel_ms']) for k, v in list(d['configs'].items())[:2]))
"
done
done

# A/B of prebuilt library variants: headline greedy kernel (L=150, e=0.10, W=3) and the full-path leg (L=100, e=0.20, W=2) of one bench run
#   tools/ab_both.sh _build/ab/a.so _build/ab/b.so ...      (2 interleaved repetitions)
for rep in 1 2; do
for lib in "$@"; do
  GASMA_B200_LIB=$PWD/$lib python bench.py --steps 24 --warmup 4 --no-cpu-baseline --no-configs --no-parity 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); f=d['full_path']
        print('$lib', 'headline kernel %.4f ms step %.4f ms |' % (d['roofline']['kernel_ms'], d['ms_per_step']), 'full path %.4g/s step %.4f ms greedy %.4f nw %.4f' % (f['value'], f['ms_per_step'], f['greedy_kernel_ms'], f['nw_kernel_ms']))
"
done
done

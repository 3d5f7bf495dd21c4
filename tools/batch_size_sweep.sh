for P in 500000 1000000 2000000 4000000; do
python bench.py --pairs $P --steps 12 --warmup 3 --no-full-path --no-configs --no-cpu-baseline --no-parity --in-flight 3 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); f=d.get('in_flight',{})
        print('pairs=$P value %.4g step/1M %.4f kernel/1M %.4f | in flight %.4g' % (d['value'], d['ms_per_step']*1e6/$P, d['roofline']['kernel_ms']*1e6/$P, f.get('value',0)))
"
done

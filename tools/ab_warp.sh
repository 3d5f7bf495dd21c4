# A/B of prebuilt libraries on the long-read legs (configs[4]: 1 kbp, e = 0.15, k = 24 and k = 50; one alignment per warp)
#   tools/ab_warp.sh _build/ab/a.so _build/ab/b.so ...      kernel ms per 100 k pairs, parity against the oracle on a prefix
for lib in "$@"; do
  export GASMA_B200_LIB=$PWD/$lib
  for K in 24 50; do
    python bench.py --length 1000 --error 0.15 --k $K --pairs 100000 --steps 10 --warmup 3 --no-full-path --no-configs --no-cpu-baseline \
        2> gpurun_out/ab_warp.err | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1])
print('$lib k=$K: %.1f M/s, kernel %.4f ms, step %.4f ms, parity %s' % (d['value']/1e6, d['roofline']['kernel_ms'], d['ms_per_step'], d['parity']['ok']))"
  done
done

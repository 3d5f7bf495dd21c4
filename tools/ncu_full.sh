# one --set full capture of one kernel per library build, on the bench command
#   tools/ncu_full.sh <kernel regex> <skip launches> <bench flags...> -- path/to/a.so ...   -> gpurun_out/full_<name>.ncu-rep
K=$1; SKIP=$2; shift 2
FLAGS=""
while [ "$1" != "--" ]; do FLAGS="$FLAGS $1"; shift; done
shift
CMD="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-configs --no-parity $FLAGS"
for lib in "$@"; do
  export GASMA_B200_LIB=$PWD/$lib
  $CMD > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:$K -s $SKIP -c 1 -f -o gpurun_out/full_$(basename $lib .so) $CMD > /dev/null 2>&1
  echo "$lib done rc=$?"
done

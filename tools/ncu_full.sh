# one --set full capture of the greedy kernel per library build (bench command, greedy-only leg)
#   tools/ncu_full.sh path/to/a.so ...   -> gpurun_out/full_<name>.ncu-rep
CMD="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-full-path --no-configs --no-parity"
for lib in "$@"; do
  export GASMA_B200_LIB=$PWD/$lib
  $CMD > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:greedy_kernel -s 3 -c 1 -f -o gpurun_out/full_$(basename $lib .so) $CMD > /dev/null 2>&1
  echo "$lib done rc=$?"
done

# A/B of library builds on the greedy bench line: kernel time from a plain run, executed warp instructions and average
# active threads of the greedy kernel from a metrics-only ncu pass of the same command.
#   tools/ab_instr.sh path/to/a.so path/to/b.so ...
CMD="python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-full-path --no-configs --no-parity"
for lib in "$@"; do
  export GASMA_B200_LIB=$PWD/$lib
  $CMD 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$lib', 'value %.4g' % d['value'], 'kernel %.4f ms' % d['roofline']['kernel_ms'])
" && ncu --metrics smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:greedy_kernel -s 3 -c 1 --csv --log-file gpurun_out/ab_$(basename $lib .so).csv $CMD > /dev/null 2>&1
  python - <<PY
import csv
rows=[r for r in csv.reader(open('gpurun_out/ab_$(basename $lib .so).csv')) if len(r)>10]
for r in rows[1:]:
    print('   ', r[-3], r[-1])
PY
done

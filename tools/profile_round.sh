# Evidence of one library build, in one gpurun call (run from the repo root):  tools/profile_round.sh <tag>
#   1. plain bench run (must exit 0)                        -> gpurun_out/<tag>_bN1.json
#   2. launch list of the same command (cold, serialised)   -> gpurun_out/<tag>_launches.csv
#   3. metrics-only passes (DRAM bytes, instructions, pipes) of the greedy and NW kernels -> profiles/traffic.json
#      (keyed by the library's SHA-1; bench.py only uses figures of the build it has loaded)
#   4. one --set full capture of each dominant kernel       -> gpurun_out/<tag>_{greedy,nw}_full.ncu-rep
TAG=$1
SHORT="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-parity --no-configs"
M=$(python tools/update_traffic.py --metrics)
python bench.py > gpurun_out/${TAG}_bN1.json 2> gpurun_out/${TAG}_bN1.err || { echo "plain bench failed"; exit 1; }
$SHORT > /dev/null 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $SHORT > /dev/null 2>&1
ncu --metrics $M --clock-control none -k regex:greedy_kernel -s 3 -c 4 --csv --log-file gpurun_out/${TAG}_traffic_headline.csv $SHORT --no-full-path > /dev/null 2>&1
python tools/update_traffic.py gpurun_out/${TAG}_traffic_headline.csv headline
ncu --metrics $M --clock-control none -k regex:nw_band_kernel -s 1 -c 4 --csv --log-file gpurun_out/${TAG}_traffic_full.csv $SHORT > /dev/null 2>&1
python tools/update_traffic.py gpurun_out/${TAG}_traffic_full.csv full_path
cp profiles/traffic.json gpurun_out/${TAG}_traffic.json
ncu --set full --clock-control none --import-source on -k regex:greedy_kernel -s 3 -c 1 -f -o gpurun_out/${TAG}_greedy_full $SHORT --no-full-path > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:nw_band_kernel -s 2 -c 1 -f -o gpurun_out/${TAG}_nw_full $SHORT > /dev/null 2>&1
python bench.py --no-cpu-baseline --no-configs --no-parity 2>/dev/null | python -c "
import sys, json
d = json.loads([l for l in sys.stdin if l.startswith('{')][0])
print('with figures:', d['roofline'].get('traffic'), d['roofline'].get('issue', {}).get('frac'), d['full_path']['roofline'].get('traffic'))"
echo "profile_round done"

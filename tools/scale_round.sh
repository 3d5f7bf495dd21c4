# Multi-GPU evidence of one build, in one gpurun --gpus 8 call:  tools/scale_round.sh <tag>
#   link table (H2D / D2H per GPU with 1, 2, 4, 8 GPUs copying at once) and the bench line at N = 2, 4, 8
TAG=$1
python tools/link_probe_multi.py > gpurun_out/${TAG}_link_table.txt 2>&1
for N in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + N)) \
      bench.py --gpus $N --no-cpu-baseline > gpurun_out/${TAG}_bN$N.json 2> gpurun_out/${TAG}_bN$N.err
  echo "N=$N rc=$?"
done
python bench.py > gpurun_out/${TAG}_bN1.json 2> gpurun_out/${TAG}_bN1.err; echo "N=1 rc=$?"

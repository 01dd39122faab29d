# BASELINE.json configs[4]: cover 1200x675 at 4096 spp, sample-split over N = 8, 4, 2, 1 GPUs of one box.
# usage (under gpurun --gpus 8): bash tools/c5_sweep.sh
O=gpurun_out; mkdir -p $O
for N in 8 4 2; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) \
    bench.py --gpus $N --spp 4096 --steps 3 --warmup 1 --no-cpu > $O/c5_n$N.json 2> $O/c5_n$N.err; echo "N=$N rc=$?"; cut -c1-200 $O/c5_n$N.json
done
python bench.py --spp 4096 --steps 3 --warmup 1 --no-cpu > $O/c5_n1.json 2> $O/c5_n1.err; echo "N=1 rc=$?"; cut -c1-200 $O/c5_n1.json

# The driver's scaling run, as one box visit: bench.py at N = 8, 4, 2, 1 back to back + the multi-GPU tests.
# usage (under gpurun --gpus 8): bash tools/scaling_round.sh
O=gpurun_out; mkdir -p $O
python -m pytest tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -2
for N in 8 4 2; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500+N)) \
    bench.py --gpus $N --no-cpu > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "N=$N rc=$?"; cut -c1-260 $O/bench_n$N.json
done
python bench.py --no-cpu > $O/bench_n1.json 2> $O/bench_n1.err; echo "N=1 rc=$?"; cut -c1-260 $O/bench_n1.json

# One GPU-box visit: parity tests, bench, ncu launch list, one full ncu capture of the path kernel.
# usage (under gpurun): bash tools/gpu_round.sh <tag>
TAG=${1:-r01}
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest_gpu.log; tail -3 $O/${TAG}_pytest_gpu.log
python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"; cut -c1-400 $O/${TAG}_bench.json
python bench.py --impl reference --steps 1 --warmup 1 > $O/${TAG}_bench_ref.json 2>> $O/${TAG}_bench.err; cut -c1-300 $O/${TAG}_bench_ref.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu > $O/${TAG}_ncu_bench.log 2>&1; echo "ncu list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:path_kernel -c 1 -f -o $O/${TAG}_path_full \
  python tools/run_once.py --spp 16 > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"

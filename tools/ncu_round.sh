# ncu evidence of one kernel build: launch list of the bench command + one full capture of the path
# kernel on the bench frame.   usage (under gpurun): bash tools/ncu_round.sh <tag> [kernel regex]
TAG=${1:-r02}
KERNEL=${2:-fused_tc_kernel}
O=gpurun_out; mkdir -p $O
python bench.py --steps 2 --warmup 3 --no-cpu --no-other > $O/${TAG}_bench_short.json 2> $O/${TAG}_bench_short.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu --no-other > $O/${TAG}_ncu_bench.log 2>&1; echo "ncu list rc=$?"
python tools/run_once.py --spp 500 > $O/${TAG}_run_once.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$KERNEL -c 1 -f -o $O/${TAG}_fused_full \
  python tools/run_once.py --spp 500 > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -1 $O/${TAG}_run_once.log

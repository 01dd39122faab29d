# One GPU-box visit: parity tests (all render modes), bench (both arms), ncu launch list, one full
# ncu capture of the dominant kernel at the bench workload.   usage (under gpurun): bash tools/gpu_round.sh <tag>
TAG=${1:-r01}
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest_gpu.log; tail -3 $O/${TAG}_pytest_gpu.log
for M in wave; do RT_B200_MODE=$M python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -1 | sed "s/^/mode $M: /"; done | tee $O/${TAG}_pytest_modes.log
python bench.py --impl reference --steps 1 --warmup 1 > $O/${TAG}_bench_ref.json 2> $O/${TAG}_bench.err; cut -c1-300 $O/${TAG}_bench_ref.json
python bench.py > $O/${TAG}_bench.json 2>> $O/${TAG}_bench.err; echo "bench rc=$?"; cut -c1-400 $O/${TAG}_bench.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu > $O/${TAG}_ncu_bench.log 2>&1; echo "ncu list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:fused_kernel -c 1 -f -o $O/${TAG}_fused_full \
  python tools/run_once.py --spp 500 > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
for cfg in "three 400 225 100" "cover 3840 2160 500" "cover 1200 675 4096"; do set -- $cfg
  echo "== $cfg $(python tools/run_once.py --scene $1 --w $2 --h $3 --spp $4 --reps 2 | tail -1)"; done | tee $O/${TAG}_configs.log

# One GPU-box visit: parity tests, bench (both arms), ncu launch list of the bench command, one full
# ncu capture of the dominant kernel at the bench workload (each ncu pass only after the same command
# has exited 0 without it).   usage (under gpurun): bash tools/gpu_round.sh <tag> [kernel regex]
TAG=${1:-r02}
KERNEL=${2:-fused_tc_kernel}
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest_gpu.log; tail -3 $O/${TAG}_pytest_gpu.log
python bench.py --impl reference --steps 1 --warmup 1 > $O/${TAG}_bench_ref.json 2> $O/${TAG}_bench.err; cut -c1-300 $O/${TAG}_bench_ref.json
python bench.py > $O/${TAG}_bench.json 2>> $O/${TAG}_bench.err; echo "bench rc=$?"; cut -c1-400 $O/${TAG}_bench.json
bash tools/ncu_round.sh $TAG $KERNEL
for cfg in "three 400 225 100" "cover 3840 2160 500" "cover 1200 675 4096"; do set -- $cfg
  echo "== $cfg $(python tools/run_once.py --scene $1 --w $2 --h $3 --spp $4 --reps 2 | tail -1)"; done | tee $O/${TAG}_configs.log

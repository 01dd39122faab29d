# A/B of library builds in build_variants/ at the bench workload (run under gpurun).
# usage: bash tools/ab_libs.sh [spp]   -> gpurun_out/ab_libs.log (image hash of every variant)
SPP=${1:-500}
mkdir -p gpurun_out; : > gpurun_out/ab_libs.log
for L in "" $(ls build_variants/*.so 2>/dev/null); do
  N=$(basename "${L:-default}" .so)
  echo "== $N" >> gpurun_out/ab_libs.log
  env ${L:+RT_B200_LIB=$PWD/$L} python tools/run_once.py --spp $SPP --reps 3 2>&1 | tail -2 | python -c "import sys,json; [print(round(json.loads(l)['render_ms'],2), json.loads(l)['image_sha256']) for l in sys.stdin]" >> gpurun_out/ab_libs.log
done
cat gpurun_out/ab_libs.log

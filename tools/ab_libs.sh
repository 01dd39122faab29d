# A/B of library builds in build_variants/ at the bench workload (run under gpurun).
# usage: [RS="2 4"] bash tools/ab_libs.sh [spp]   -> gpurun_out/ab_libs.log (+ md5 of every variant's image)
SPP=${1:-500}
mkdir -p gpurun_out; : > gpurun_out/ab_libs.log
for L in "" $(ls build_variants/*.so 2>/dev/null); do
  N=$(basename "${L:-default}" .so)
  for R in ${RS:-2}; do
    echo "== $N R=$R" >> gpurun_out/ab_libs.log
    env RT_B200_RAYS_PER_LANE=$R ${L:+RT_B200_LIB=$PWD/$L} python tools/run_once.py --spp $SPP --reps 3 --ppm /tmp/ab_$N.ppm 2>&1 | tail -2 | cut -c1-140 >> gpurun_out/ab_libs.log
    md5sum /tmp/ab_$N.ppm | cut -c1-32 >> gpurun_out/ab_libs.log
  done
done
cat gpurun_out/ab_libs.log

# A/B of render modes (run under gpurun).  usage: bash tools/ab_wave.sh [test]
mkdir -p gpurun_out; : > gpurun_out/abw.log
if [ "$1" = "test" ]; then python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -15 gpurun_out/pytest_gpu.log; fi
run() { echo "== $*" >> gpurun_out/abw.log; env "$@" python tools/run_once.py --spp ${SPP:-128} --reps 3 2>&1 | tail -1 | cut -c1-230 >> gpurun_out/abw.log; }
for L in "" $(ls build_variants/*.so 2>/dev/null); do
for M in ${MODES:-fused}; do for R in ${RS:-2 4}; do
  run RT_B200_MODE=$M RT_B200_RAYS_PER_LANE=$R RT_B200_LIB=${L:+$PWD/$L}
done; done; done
cat gpurun_out/abw.log

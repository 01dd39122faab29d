# A/B harness for kernel variants: RT_B200_RAYS_PER_LANE x library builds (run under gpurun)
mkdir -p gpurun_out; : > gpurun_out/ab.log
if [ "$1" = "test" ]; then python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log; fi
for R in 2 4; do for L in "" $(ls build_variants/*.so 2>/dev/null); do
  echo "== R=$R lib=$L" >> gpurun_out/ab.log
  RT_B200_RAYS_PER_LANE=$R RT_B200_LIB=${L:+$PWD/$L} python tools/run_once.py --spp 128 --reps 3 2>&1 | tail -1 | cut -c1-200 >> gpurun_out/ab.log
done; done
cat gpurun_out/ab.log

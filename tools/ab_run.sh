# A/B harness for kernel variants (run under gpurun): env combos x library builds
mkdir -p gpurun_out; : > gpurun_out/ab.log
if [ "$1" = "test" ]; then python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log; fi
for R in ${RS:-2 4}; do for NC in "" 1; do for L in "" $(ls build_variants/*.so 2>/dev/null); do
  echo "== R=$R noconst=$NC lib=$L" >> gpurun_out/ab.log
  RT_B200_RAYS_PER_LANE=$R RT_B200_NO_CONST_ROWS=$NC RT_B200_LIB=${L:+$PWD/$L} python tools/run_once.py --spp 128 --reps 3 2>&1 | tail -1 | cut -c1-200 >> gpurun_out/ab.log
done; done; done
cat gpurun_out/ab.log

# Build an experimental copy of the library with extra -D flags into build_variants/ (A/B runs).
# usage: bash tools/build_variant.sh <name> <nvcc -D flags...>
set -e
NAME=$1; shift
make -s -C ray-tracing-in-one-weekend_b200/csrc clean >/dev/null
make -s -C ray-tracing-in-one-weekend_b200/csrc EXTRA="$*" OUT=../../build_variants/librt_$NAME.so >/dev/null
make -s -C ray-tracing-in-one-weekend_b200/csrc clean >/dev/null
echo "$NAME: $(grep -A2 'fused_tc_kernelILb1ELb0' ray-tracing-in-one-weekend_b200/csrc/rt_wave.ptxas.log 2>/dev/null | grep -E 'registers' | head -1)"

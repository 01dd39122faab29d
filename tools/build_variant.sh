# Build an experimental copy of the library with extra -D flags into build_variants/ (A/B runs).
# usage: bash tools/build_variant.sh <name> <nvcc -D flags...>
set -e
NAME=$1; shift
C=ray-tracing-in-one-weekend_b200/csrc
mkdir -p build_variants /tmp/rtv_$NAME
FL="-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC,-O2,-ffp-contract=off -Xptxas -v"
for f in rt_wave rt_bvh rt_kernels rt_capi; do nvcc "$@" $FL $( [ $f = rt_wave ] && echo -fmad=false -prec-div=false -prec-sqrt=false ) -c -o /tmp/rtv_$NAME/$f.o $C/$f.cu 2> /tmp/rtv_$NAME/$f.log; done
nvcc "$@" $FL -fmad=false -c -o /tmp/rtv_$NAME/rt_primary_f64.o $C/rt_primary_f64.cu 2>/dev/null
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o build_variants/librt_$NAME.so /tmp/rtv_$NAME/*.o -lcudart_static -lpthread -ldl -lrt
echo "$NAME: $(grep -A2 'fused_kernelILi2ELi4ELb1ELb1' /tmp/rtv_$NAME/rt_wave.log | grep -E 'registers|spill' | tr '\n' ' ')"

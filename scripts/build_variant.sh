#!/bin/bash
# dev: build/lib_<tag>.so = the library with extra -D flags in ONE robot's kernels (default panda).
# usage: scripts/build_variant.sh <tag> "<-Dflags>" [robot] [generated-trace dir relative to mr_vamp_b200/, default _gen]
# (alternative traces: VB_GEN_OUT=mr_vamp_b200/_genv VB_...=... python -m mr_vamp_b200.codegen.emit_cuda <robot>)
set -e
cd "$(dirname "$0")/.."
tag=$1; flags=$2; robot=${3:-panda}; gen=${4:-_gen}
cap=$(echo $robot | sed 's/./\U&/')
mkdir -p build/var
nvcc -std=c++17 -O3 -diag-suppress 177 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Xptxas -v $flags \
  -DVB_TRACE_HEADER="\"../${gen}/${robot}_trace.cuh\"" -DVB_TRACE=${cap}Trace -DVB_OPS=robot_ops_${robot} \
  -c mr_vamp_b200/csrc/vb_robot.cu -o build/var/vb_robot_${robot}_${tag}.o 2> build/var/${robot}_${tag}.ptxas.log || (cat build/var/${robot}_${tag}.ptxas.log; exit 1)
objs=""
for r in sphere ur5 panda fetch baxter; do
  if [ "$r" = "$robot" ]; then objs="$objs build/var/vb_robot_${robot}_${tag}.o"; else objs="$objs build/obj/vb_robot_$r.o"; fi
done
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o build/lib_${tag}.so build/obj/vb_capi.o build/obj/vb_capt_build.o $objs -ldl
grep -A2 "fkcc_random_kernel" build/var/${robot}_${tag}.ptxas.log | grep -E "registers|spill" | head -3

#!/bin/bash
# dev: build/libvamp_clk.so = the library with -DVB_PHASE_CLOCKS in the Panda kernels (scripts/phase_clocks.py)
set -e
cd "$(dirname "$0")/.."
make -j8 2>&1 | grep -i "error" -A5 || true
cp build/obj/vb_robot_panda.o /tmp/vb_robot_panda.keep.o
touch mr_vamp_b200/csrc/vb_kernels.cuh
make -j8 NVFLAGS="-std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Xptxas -v -DVB_PHASE_CLOCKS" build/obj/vb_robot_panda.o 2>&1 | grep -i "error" -A5 || true
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o build/libvamp_clk.so build/obj/vb_capi.o build/obj/vb_robot_*.o -ldl
touch mr_vamp_b200/csrc/vb_kernels.cuh
make -j8 2>&1 | grep -i "error" -A5 || true

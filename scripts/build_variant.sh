#!/bin/bash
# A/B builds of the fast fixed kernels with other compile-time knobs:  scripts/build_variant.sh <name> <size> "<-D flags>"
# -> rsp_chains_b200/librsp_b200_<name>.so (all other objects are those of the regular build).  Select it with RSP_B200_LIB=<path>.
set -e
NAME=$1; SIZE=$2; DEFS=$3
cd "$(dirname "$0")/../rsp_chains_b200/csrc"
make -j8 > /dev/null
mkdir -p build_$NAME
nvcc -std=c++17 -O3 --expt-relaxed-constexpr -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Xptxas -v $DEFS -DRSP_FAST_S=$SIZE -c -o build_$NAME/rsp_fast_$SIZE.o rsp_fast.cu 2> build_$NAME/rsp_fast_$SIZE.log || (grep error build_$NAME/rsp_fast_$SIZE.log; exit 1)
OBJS=$(ls build/*.o | grep -v "rsp_fast_$SIZE.o")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../librsp_b200_$NAME.so $OBJS build_$NAME/rsp_fast_$SIZE.o -ldl
grep -E "registers|spill" build_$NAME/rsp_fast_$SIZE.log | sort | uniq -c | sort -rn | head -6

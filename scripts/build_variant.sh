#!/bin/bash
# build_variant.sh <name> <source-without-.cu> "<extra nvcc flags>": one translation unit rebuilt with extra -D flags and linked with the
# library's other objects into cl-waffe2_b200/lib/variants/libwaffe2cuda_<name>.so — for A/B runs inside ONE gpurun call
# (scripts/gpu.sh task ab:<script.py>), because box-to-box variance (3-7 %) is larger than most of the effects being compared.
set -e
cd "$(dirname "$0")/.."
name=$1; src=$2; flags=$3
P=cl-waffe2_b200
mkdir -p $P/_obj/variants $P/lib/variants
python $P/build.py > /dev/null
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xptxas -v $flags \
  -c $P/csrc/$src.cu -o $P/_obj/variants/${name}_$src.o > $P/_obj/variants/${name}_$src.log 2>&1 || { tail -30 $P/_obj/variants/${name}_$src.log; exit 1; }
objs=""
for s in runtime elementwise reduce softmax misc fused comm p2p conv gemm gemm_tc attention; do
  if [ "$s" = "$src" ]; then objs="$objs $P/_obj/variants/${name}_$src.o"; else objs="$objs $P/_obj/$s.o"; fi
done
/usr/local/cuda/bin/nvcc -shared -o $P/lib/variants/libwaffe2cuda_$name.so $objs -ldl
echo "built $P/lib/variants/libwaffe2cuda_$name.so"

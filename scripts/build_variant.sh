#!/bin/bash
# Builds a variant of the CUDA core HERE (no GPU needed) with extra -D flags for wavefront.cu only:
#   scripts/build_variant.sh NAME "-DGSB_X=1 -DGSB_Y=2"     ->  gpurun_variants/NAME/libgpuspectral_b200.so
# The variants travel to the GPU box with the snapshot; scripts/gpu_variants.sh runs the bench scenes with each of them, so an
# A/B of compile-time settings costs box time only for the measurements, not for nvcc.
set -e
NAME=$1; DEFS=$2
cd "$(dirname "$0")/../gpuspectral_b200"
make -s build/api.o build/bvh_build.o > /dev/null
mkdir -p build/var ../gpurun_variants/$NAME
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a $DEFS -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-fvisibility=hidden -ccbin /usr/bin/g++ \
    --expt-relaxed-constexpr -Xptxas -v -prec-div=false -prec-sqrt=false -c csrc/wavefront.cu -o build/var/wavefront_$NAME.o 2> build/var/wavefront_$NAME.ptxas.log
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o ../gpurun_variants/$NAME/libgpuspectral_b200.so build/api.o build/var/wavefront_$NAME.o build/bvh_build.o
echo "$DEFS" > ../gpurun_variants/$NAME/defs.txt
ls -la ../gpurun_variants/$NAME

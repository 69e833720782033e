#!/bin/bash
# Builds kernel variants into my-raytracer_b200/variants/<name>.so: tools/build_variants.sh name "-DFLAG=1 ..." [name flags]...
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
CS=$ROOT/my-raytracer_b200/csrc
mkdir -p $ROOT/my-raytracer_b200/variants
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  /usr/local/cuda/bin/nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xptxas=-v \
    -Xcompiler -fPIC,-ffp-contract=off -shared -ccbin /usr/bin/g++ $flags \
    $CS/rt_api.cu $CS/rt_sah_gpu.cu $CS/rt_flatten_gpu.cu $CS/rt_bvh_gpu.cu $CS/rt_normals_gpu.cu $CS/rt_flatten.cpp $CS/rt_sah.cpp $CS/rt_prep.cpp $CS/rt_scene_io.cpp $CS/rt_image_io.cpp -o $ROOT/my-raytracer_b200/variants/$name.so 2>&1 \
    | grep -A2 -E "render_kernelILi[04]E" | grep -E "registers|spill" | sed "s/^/$name: /"
done

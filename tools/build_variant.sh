#!/bin/bash
# developer helper: rebuild the library with extra -D flags (experiments), e.g. tools/build_variant.sh -DNNRT_EXPERIMENT_FLOAT_ATOMICS
cd "$(dirname "$0")/../dynamicfuion_python_b200" || exit 1
for f in c_api geometry warp_field_build mesh_normals tsdf raster fitter_kernels assemble_kernel solver_kernels linalg_api linalg_blocks assembly_stages fitter; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -fmad=false -std=c++17 --extended-lambda -Xcompiler -fPIC -Xcompiler -fvisibility=hidden -DNNRT_B200_BUILD "$@" -c csrc/$f.cu -o build/$f.o &
done
wait
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o lib/libnnrt_b200.so build/*.o -lcudart

#!/bin/bash
# Builds libsba_b200.so (sm_100a) in-tree.  nvcc cross-compiles without a GPU.
set -e
cd "$(dirname "$0")/steviapp_b200/csrc"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
CXX_HOST=$(test -x /usr/bin/g++ && echo /usr/bin/g++ || echo g++)
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -ccbin $CXX_HOST -Xcompiler -fPIC -I../../include"
OUT=../libsba_b200.so
mkdir -p ../build
pids=()
for f in sba_kernels sba_kernels_schur sba_kernels_op sba_kernels_mv sba_kernels_pcg sba_kernels_pcgsolve sba_kernels_misc sba_kernels_points sba_kernels_explicit sba_kernels_shared sba_kernels_pairs sba_kernels_dense sba_noalloc sba_solver; do
  $NVCC $FLAGS -c $f.cu -o ../build/$f.o &
  pids+=($!)
done
$NVCC $FLAGS -x cu -c sba_comm.cpp -o ../build/sba_comm.o & pids+=($!)
$NVCC $FLAGS -x cu -c sba_capi.cpp -o ../build/sba_capi.o & pids+=($!)
$NVCC $FLAGS -x cu -c sba_bal.cpp -o ../build/sba_bal.o & pids+=($!)
$NVCC $FLAGS -x cu -c sba_projectors.cpp -o ../build/sba_projectors.o & pids+=($!)
$NVCC $FLAGS -x cu -c sba_project.cpp -o ../build/sba_project.o & pids+=($!)
for p in "${pids[@]}"; do wait $p; done
$NVCC -shared -o $OUT ../build/*.o -ccbin $CXX_HOST -lcudart_static -ldl -lpthread -lrt
echo "built $(readlink -f $OUT)"

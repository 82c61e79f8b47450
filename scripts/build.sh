#!/bin/bash
# Build libcib200.so in-tree for sm_100a (cross-compiles without a GPU).
set -e
cd "$(dirname "$0")/.."
mkdir -p cori_intml_examples_b200/lib
SAN=""
[ -n "$SANITIZE" ] && SAN="-Xcompiler -fsanitize=$SANITIZE -Xcompiler -fno-omit-frame-pointer"   # host-side ASAN / TSAN build
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared $SAN \
     -I csrc -I include csrc/cib_api.cu -o cori_intml_examples_b200/lib/libcib200.so "$@"
g++ -O2 -std=c++17 -fPIC -shared $([ -n "$SANITIZE" ] && echo "-fsanitize=$SANITIZE -fno-omit-frame-pointer") \
     csrc/h5writer.cpp csrc/h5reader.cpp -o cori_intml_examples_b200/lib/libcibh5.so
echo "built cori_intml_examples_b200/lib/libcib200.so and libcibh5.so"

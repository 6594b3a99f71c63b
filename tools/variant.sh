#!/bin/bash
# Build a variant of ONE kernel file with extra -D flags into vk-renderer_b200/libvkr_b200_<name>.so (other objects reused):
#   tools/variant.sh <name> <file.cu> "<flags>"      then run with VKR_B200_LIB=$PWD/vk-renderer_b200/libvkr_b200_<name>.so
set -e
cd "$(dirname "$0")/../vk-renderer_b200"
name=$1; src=$2; flags=$3
mkdir -p build/var_$name
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC $flags -c csrc/$src -o build/var_$name/${src%.cu}.o
objs=$(ls build/*.o | grep -v "/host_" | grep -v "/${src%.cu}.o")
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o libvkr_b200_$name.so $objs build/var_$name/${src%.cu}.o -Xcompiler -fPIC
cuobjdump -res-usage build/var_$name/${src%.cu}.o | grep -E "REG" | sort | uniq -c | head -4

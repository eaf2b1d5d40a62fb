#!/bin/sh
# builds the microbenchmark library next to its source (git-ignored, travels with gpurun)
cd "$(dirname "$0")" && nvcc -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -shared -Xcompiler -fPIC \
  -o librandom_rows.so random_rows.cu

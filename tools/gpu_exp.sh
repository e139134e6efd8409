#!/bin/bash
# scratch: A/B of the narrow tile shape (build = 8 lanes: 32 x 64 cells, build_exp4 = 4 lanes: 16 x 128 cells)
mkdir -p gpurun_out
{
  for b in build build_exp4; do
    echo "## $b test [jacobi]"; timeout 200 ./$b/bin/test_cuda "[jacobi]" 2>&1 | tail -2
  done
  for rep in 1 2; do
    for sz in "645 645 323" "813 407 407" "512 512 512" "333 300 200"; do
      for b in build build_exp4; do
        echo "## $b $sz"; timeout 100 ./$b/bin/jacobi3d $sz --exact-size -n 40 --device-time 2>&1 | grep "jacobi3d-device"
      done
    done
  done
  echo "## verify exp4"; timeout 200 ./build_exp4/bin/jacobi3d 645 645 323 --exact-size --verify 20 2>&1 | tail -2
} > gpurun_out/exp.txt 2>&1
cat gpurun_out/exp.txt

#!/bin/bash
# Builds compile-time variants of libvxorb.so for A/B runs on the GPU box:
#   scratch/ab.sh name1 "-DFLAG=1" name2 "-DOTHER=2 -DX=3" ...
# -> visionx_slam_b200/variants/libvxorb_<name>.so (git-ignored *.so; travels with gpurun)
set -e
cd "$(dirname "$0")/../visionx_slam_b200/csrc"
mkdir -p ../variants
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  make -s BUILD=build_$name OUT=../variants/libvxorb_$name.so EXTRA="$flags" > /dev/null
  echo "built variant $name ($flags)"
done

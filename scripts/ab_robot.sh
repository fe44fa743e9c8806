#!/bin/bash
# dev: A/B of one robot's kernel variants (scripts/build_variant.sh): throughput on the MBM / cage scenes + the robot's parity tests
# usage (under gpurun): bash scripts/ab_robot.sh <robot> <tag> [<tag> ...]     (tag "base" = the committed library)
robot=$1; shift
for tag in "$@"; do
  lib=build/lib_$tag.so; [ "$tag" = base ] && lib=mr_vamp_b200/lib/libvamp_b200.so
  for scene in mbm cage; do
    VAMP_B200_LIB=$lib timeout 60 python scripts/probe_throughput.py $robot $scene 2>&1 | tail -1 | sed "s/^/$tag: /"
  done
  if [ "$tag" != base ]; then
    VAMP_B200_LIB=$lib timeout 300 python -m pytest tests/test_fkcc_gpu.py -m gpu -x -q -k "$robot" 2>&1 | tail -2 | sed "s/^/$tag: /"
  fi
done

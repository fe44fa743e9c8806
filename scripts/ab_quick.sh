#!/bin/bash
# dev: throughput only of one robot's variants. usage (under gpurun): bash scripts/ab_quick.sh <robot> <tag> ...
robot=$1; shift
for tag in "$@"; do
  lib=build/lib_$tag.so; [ "$tag" = base ] && lib=mr_vamp_b200/lib/libvamp_b200.so
  for scene in mbm cage clutter; do
    VAMP_B200_LIB=$lib timeout 60 python scripts/probe_throughput.py $robot $scene 2>&1 | tail -1 | sed "s/^/$tag: /"
  done
done

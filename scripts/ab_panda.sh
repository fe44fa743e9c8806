#!/bin/bash
# dev: A/B of Panda kernel variants built by scripts/build_variant.sh: build/lib_<tag>.so ...
# usage (under gpurun): bash scripts/ab_panda.sh tag1 tag2 ...
for tag in "$@"; do
  for scene in mbm clutter cage; do
    VAMP_B200_LIB=build/lib_$tag.so timeout 60 python scripts/probe_throughput.py panda $scene 2>&1 | tail -1 | sed "s/^/$tag: /"
  done
done

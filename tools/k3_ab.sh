#!/bin/bash
# developer A/B of K3 compile-time switches on the GPU box: tools/k3_ab.sh "<defs A>" "<defs B>" ...
# (each variant: developer build; persistent kernel with and without helpers, one-warp-per-pair kernel)
# K3AB_PROFILE=1: also a build with the per-phase cycle counters (SEPB_K3_PROFILE) and their split
# K3AB_SCHEDS: comma list for k3_time.py (default persistent,nohelp,per_pair)
for defs in "$@"; do
  SEPB_DEVELOPER=1 SEPB_K3_DEFS="$defs" python -m multi_robot_slam_separators_b200.build --force > /dev/null 2>&1 || echo "build failed: $defs"
  SEPB_SHORT=1 SEPB_K3_DEFS="$defs" timeout 300 python tools/k3_time.py 10000 ${K3AB_SCHEDS:-persistent,nohelp,per_pair} 2>&1 | grep K3_DEFS
  if [ -n "$K3AB_PROFILE" ]; then
    SEPB_K3_PROFILE=1 SEPB_DEVELOPER=1 SEPB_K3_DEFS="$defs" python -m multi_robot_slam_separators_b200.build --force > /dev/null 2>&1 || echo "profile build failed: $defs"
    SEPB_K3_PROFILE=1 SEPB_K3_DEFS="$defs" timeout 300 python tools/k3_time.py 10000 persistent 2>&1 | grep -E "K3_DEFS|chunks"
  fi
done

#!/bin/bash
# developer A/B of K3 compile-time switches on the GPU box: tools/k3_ab.sh "<defs A>" "<defs B>" ...
for defs in "$@"; do
  SEPB_K3_DEFS="$defs" python -m multi_robot_slam_separators_b200.build --force > /dev/null 2>&1 || echo "build failed: $defs"
  SEPB_K3_DEFS="$defs" python tools/k3_time.py 10000 2>&1 | tail -3
done

#!/bin/bash
# developer A/B: K2T accumulator tile width (run-time switch SEPB_K2T_TN = 128 | 256; default = widest that fits)
for tn in 128 256; do
  echo "== TN=$tn"
  SEPB_K2T_TN=$tn timeout 120 python tools/k2_bench.py 256 20000 tensor 2>&1 | tail -2
done

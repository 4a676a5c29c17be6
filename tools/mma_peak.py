"""Tensor-pipe issue rates of the MMA shapes K1 / K2T use (bare tcgen05.mma loops, one CTA per SM)."""
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from multi_robot_slam_separators_b200 import capi
import torch

ctx = capi.Context(0)
sm = torch.cuda.get_device_properties(0).multi_processor_count
out = {}
for kind in ("mma_i8_n128", "mma_i8_n256", "mma_bf16_n256", "mma_i8_n128_sts"):
    v = ctx.measure_peak(kind)
    out[kind] = {"ops_per_s": v, "ops_per_s_per_sm": v / sm}
print(json.dumps(out))

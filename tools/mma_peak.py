"""Developer tool: tensor-pipe issue rates of the MMA shapes the matchers use (bare tcgen05.mma loops, peaks.cu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from multi_robot_slam_separators_b200 import capi
ctx = capi.Context(0)
for kind in ("mma_i8_n128", "mma_i8_n256", "mma_bf16_n256", "mma_i8_n128_sts", "mma2_i8_m256n256", "mma2_bf16_m256n256"):
    try:
        print("%-22s %.1f Tops/s" % (kind, ctx.measure_peak(kind) / 1e12), flush=True)
    except Exception as e:
        print(kind, "FAILED", e, flush=True)
ctx.close()

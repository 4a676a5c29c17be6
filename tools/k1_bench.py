"""Developer tool: time K1 (exact vs tensor filter) on resident data.  python tools/k1_bench.py [Q N D k]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from multi_robot_slam_separators_b200 import capi, synth  # noqa: E402

Q, N, D, k = [int(x) for x in (sys.argv[1:5] if len(sys.argv) >= 5 else (10000, 10000, 128, 5))]
places = synth.place_assignment(2, max(Q, N), 0.3, np.random.default_rng([2000, 1]))
nv_a, nv_b = synth.make_netvlad(places[0], D, 2000)[:N], synth.make_netvlad(places[1], D, 2000)[:Q]
ctx = capi.Context(0)
db = capi.NetvladDB(ctx, D, N)
db.append(nv_a)
dq = torch.from_numpy(nv_b).cuda()
idx = torch.empty((Q, k), dtype=torch.int32, device="cuda")
dist = torch.empty((Q, k), dtype=torch.float64, device="cuda")
stream = torch.cuda.ExternalStream(ctx.stream)
L = capi.lib()
res = {}
for name, mode in (("exact", 0), ("tensor", 1)):
    def run():
        ctx.check(L.sepb_netvlad_topk_dev(ctx._h, db._h, C.c_void_p(dq.data_ptr()), Q, k, mode, None, None, None,
                                          C.c_void_p(idx.data_ptr()), C.c_void_p(dist.data_ptr())))
    for _ in range(3):
        run()
    ctx.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5 if mode == 0 else 20
    with torch.cuda.stream(stream):
        e0.record(stream)
        for _ in range(reps):
            run()
        e1.record(stream)
    e1.synchronize()
    ms = e0.elapsed_time(e1) / reps
    res[name] = (idx.cpu().numpy().copy(), dist.cpu().numpy().copy())
    print(f"{name}: {ms:.3f} ms  {Q / ms * 1e3:.3e} queries/s  {2.0 * Q * N * D / ms / 1e9:.1f} TFLOP/s algorithmic")
print("identical:", np.array_equal(res["exact"][0], res["tensor"][0]) and np.array_equal(res["exact"][1], res["tensor"][1]))
v, i2, d2, stats = db.filter_debug(nv_b[: min(Q, 2000)], k=k, want_v=False)
print("filter stats (first rows): candidates", stats[0], "proven", stats[1], "fallback", stats[2], "inserts", stats[3])

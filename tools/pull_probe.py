"""Diagnostic for the end-to-end leg: what do the key-point pulls of host-bound stores (K2T tail, mapped pinned memory)
cost the verification when a bulk H2D copy (the next step's descriptors) saturates PCIe at the same time?
    python tools/pull_probe.py [--frames 5000]
Times verify_dev (K2 + K3 + covariance) with CUDA events on the context's stream in four situations:
resident stores / host-bound stores  x  quiet bus / 1 GB cudaMemcpyAsync H2D in flight on another stream."""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=5000)
    ap.add_argument("--reps", type=int, default=4)
    ap.add_argument("--cov", action="store_true", help="with the covariance (engine.verify_dev); default: what the bench's e2e leg asks for")
    a = ap.parse_args()
    import torch
    from multi_robot_slam_separators_b200 import capi, synth
    from multi_robot_slam_separators_b200.engine import Camera, SeparatorEngine
    t0 = time.time()
    A, B = synth.make_robots(2, a.frames, n_feat=1000, dim=128, seed=2000)
    print("world %.1f s" % (time.time() - t0), flush=True)
    eng = SeparatorEngine(Camera(synth.FX, synth.FY, synth.CX, synth.CY), 1000, 128, a.frames, a.frames, device=0)
    pin = lambda x: torch.from_numpy(np.ascontiguousarray(x)).pin_memory().numpy()  # noqa: E731
    hA = {k: pin(getattr(A, k)) for k in ("desc", "kp2d", "kp3d", "count", "octave")}
    hB = {k: pin(getattr(B, k)) for k in ("desc", "kp2d", "kp3d", "count", "octave")}

    def upload(lazy):
        eng.n_local = eng.n_peer = 0
        eng.add_peer_keyframes(hA["desc"], hA["kp2d"], hA["kp3d"], hA["count"], hA["octave"], async_=True, lazy=lazy)
        eng.add_local_keyframes(hB["desc"], hB["kp2d"], hB["kp3d"], hB["count"], hB["octave"], async_=True, lazy=lazy)
        eng.local.wait()
        eng.peer.wait()

    upload(False)
    eng.receive_descriptors(A.netvlad)
    d_q = torch.from_numpy(B.netvlad).cuda()
    idx, _ = eng.candidates_dev(d_q, 5, capi.NN_TENSOR)
    pairs = eng.pairs_from_candidates(idx)
    eng.sync()
    P = pairs.shape[0]
    bulk_src = torch.empty(1 << 30, dtype=torch.uint8).pin_memory()
    bulk_dst = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")
    side = torch.cuda.Stream()

    import ctypes as C

    def verify(pairs):
        if a.cov:
            return eng.verify_dev(pairs)        # + covariance: its To 3-D points are a fetch phase after K3
        o = {k: eng._tensor(k, shp, dt) for k, shp, dt in (
            ("status", (P,), torch.int32), ("n_matches", (P,), torch.int32), ("n_inliers", (P,), torch.int32),
            ("pose", (P, 6), torch.float64), ("transform", (P, 12), torch.float32), ("match_from", (P, 1000), torch.int16),
            ("match_to", (P, 1000), torch.int16))}
        out = capi.VerifyOut(**{k: v.data_ptr() for k, v in o.items()})
        eng.ctx.check(capi.lib().sepb_verify_pairs_dev(eng.ctx._h, eng.peer._h, eng.local._h, C.c_void_p(pairs.data_ptr()), P,
                                                       C.byref(eng.mp), C.byref(eng.pnp), None, 0, C.byref(out)))
        return o

    def run(lazy, busy):
        ms = []
        for rep in range(a.reps + 1):
            upload(lazy)
            torch.cuda.synchronize()
            if busy:
                with torch.cuda.stream(side):
                    bulk_dst.copy_(bulk_src, non_blocking=True)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(eng.stream):
                e0.record(eng.stream)
                res = verify(pairs)
                e1.record(eng.stream)
            torch.cuda.synchronize()
            if rep:
                ms.append(e0.elapsed_time(e1))
        ok = int((res["status"] == 1).sum().item())
        pulled = (eng.peer.fetched_bytes + eng.local.fetched_bytes) if lazy else 0
        print("%-10s %-6s verify %.2f ms (min %.2f) for %d pairs, %d ok, pulled %.1f MB"
              % ("host-bound" if lazy else "resident", "busy" if busy else "quiet", float(np.median(ms)), min(ms), P, ok, pulled / 1e6),
              flush=True)

    for lazy in (False, True):
        for busy in (False, True):
            run(lazy, busy)


if __name__ == "__main__":
    main()

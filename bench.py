#!/usr/bin/env python
"""bench.py -- verified keyframe pairs/s (Hamming match + PnP-RANSAC) on BASELINE.json configs[1]:
2 robots x 10k synthetic stereo keyframes (1000 256-bit features + 128-d NetVLAD each), all-pairs
NetVLAD NN, top-5 candidates per query verified.  One process per GPU; at N > 1 every rank owns an
independent robot pair (candidate pairs partition with no data-path collective: weak scaling).

  python bench.py --gpus N --steps K --warmup W            our arm (CUDA, libsepb200.so)
  python bench.py --impl reference ...                      the reference's CPU library calls (oracle/)
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "verified keyframe pairs/sec (match+PnP-RANSAC)"
UNIT = "pairs/s"


def env_int(name, default):
    return int(os.environ.get(name, default))


def synth_kwargs(args):
    return {"hard_negatives": 0.5} if args.mix == "hard" else {}


def workload_config(args):
    return {"workload": "configs[%d]: 2 robots x %d keyframes, %d ORB-like 256-bit features + %d-d NetVLAD per frame, "
                        "all-pairs NetVLAD NN + top-%d Hamming/PnP-RANSAC verification%s"
                        % (args.config, args.frames, args.feat, args.dim, args.topk,
                           "; hard-negative mix: half of the clutter descriptors come from a shared 2000-word vocabulary, so "
                           "unrelated frames send ~30 all-outlier correspondences into RANSAC" if args.mix == "hard" else ""),
            "keyframes_per_robot": args.frames, "features_per_frame": args.feat, "netvlad_dim": args.dim,
            "top_k": args.topk, "pairs_per_step_per_gpu": args.frames * args.topk, "ransac_iterations": args.iterations,
            "mix": args.mix, "pnp_flags": {0: "ITERATIVE (5-point EPnP sets)", 2: "P3P (4-point sets, EPnP refit)"}[args.pnp_flags], "nndr": 0.8, "netvlad_distance": "inf (every top-k candidate is verified)",
            "l2_policy": "inputs larger than L2 (feature stores 1.1 GB per GPU, 126 MB L2)",
            "partition": "independent robot pair per rank, no collective on the data path"}


# ---------------------------------------------------------------------------------------------
def clocks_sampler_start(gpu_index):
    f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
    q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    try:
        p = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + q, "--format=csv,noheader,nounits",
                              "-lms", "50"], stdout=f, stderr=subprocess.DEVNULL)
    except Exception:
        return None, f.name
    return p, f.name


def clocks_sampler_stop(p, path):
    out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
    if p is not None:
        p.terminate()
        try:
            p.wait(timeout=5)
        except Exception:
            p.kill()
    try:
        rows = [l.strip().split(", ") for l in open(path) if l.strip()]
        sm = [float(r[0]) for r in rows if len(r) >= 7]
        if sm:
            out["sm_mhz"] = float(np.median(sm))
            out["sm_max_mhz"] = float(rows[0][1])
            out["samples"] = len(sm)
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            out["reasons"] = [n for i, n in enumerate(names) if any(r[3 + i].strip() == "Active" for r in rows if len(r) >= 7)]
    except Exception:
        pass
    finally:
        try:
            os.unlink(path)
        except OSError:
            pass
    return out


# ---------------------------------------------------------------------------------------------
def cpu_reference_step(sample, threads, pool=None):
    """The reference's own library calls on a bounded sample of the workload: scipy cdist + top-k,
    then per pair cv2.BFMatcher(HAMMING).knnMatch(k=2)+NNDR+uniqueness and cv2.solvePnPRansac.
    pool: a process pool created by the caller OUTSIDE the timed loop (fork + teardown are not the reference's work).
    Returns (pairs_done, seconds, nn_queries, nn_seconds, verified)."""
    from oracle import find_matches as fm
    t0 = time.perf_counter()
    if pool is not None:      # the NN rows are independent too: one chunk of query rows per worker
        parts = pool.map(_cpu_nn_rows, [(i, threads) for i in range(threads)])
        idx = np.empty((sample["q"].shape[0], sample["k"]), np.int32)
        for i, p_ in enumerate(parts):
            idx[i::threads] = p_
    else:
        idx, _ = fm.topk(sample["q"], sample["db"], sample["k"])
    t_nn = time.perf_counter() - t0
    pairs = [(int(idx[r, j]), r) for r in range(idx.shape[0]) for j in range(idx.shape[1]) if idx[r, j] >= 0]
    t1 = time.perf_counter()
    if pool is not None:
        res = pool.map(_cpu_verify_pair, pairs, chunksize=max(1, len(pairs) // (threads * 8)))
    else:
        _cpu_worker_init(sample)
        res = [_cpu_verify_pair(p) for p in pairs]
    t_ver = time.perf_counter() - t1
    return len(pairs), t_nn + t_ver, idx.shape[0], t_nn, int(sum(res))


def _cpu_nn_rows(a):
    from oracle import find_matches as fm
    i, n = a
    return fm.topk(_S["q"][i::n], _S["db"], _S["k"])[0]


_S = None


def _cpu_worker_init(sample):
    global _S
    import cv2
    cv2.setNumThreads(1)
    _S = sample


def _cpu_verify_pair(p):
    import cv2
    from oracle import matching, pnp
    f, t = p
    S = _S
    A, B = S["A"], S["B"]
    fa, tb = S["a_slot"][f], S["b_slot"][t]
    pairs = matching.global_match(B.desc[tb], A.desc[fa], 0.8)
    obj, img, _, _ = matching.pnp_correspondences(pairs, A.kp3d[fa], B.kp2d[tb])
    if len(obj) < 5:
        return 0
    ok, rv, tv, inl = pnp.cv_solve_pnp_ransac(obj, img, S["K"], S["rv0"], S["tv0"], iterations=S["iterations"],
                                              flags=S.get("pnp_flags", 0))
    return int(ok and len(inl) >= 5)


def make_cpu_sample(args, seed, n_queries):
    """NetVLAD for the full config (cheap) + features only for the frames the sampled queries touch."""
    import cv2
    from multi_robot_slam_separators_b200 import synth
    from oracle import find_matches as fm, pnp
    rng = np.random.default_rng([seed, 1])
    places = synth.place_assignment(2, args.frames, 0.3, rng)
    nv_a = synth.make_netvlad(places[0], args.dim, seed)
    nv_b = synth.make_netvlad(places[1], args.dim, seed)
    rows = np.arange(n_queries)                      # first rows of robot B (the computing robot)
    idx, _ = fm.topk(nv_b[rows], nv_a, args.topk)
    need_a = np.unique(idx[idx >= 0])
    A = synth.make_frames(places[0][need_a], 0, seed, n_feat=args.feat, dim=8, **synth_kwargs(args))
    B = synth.make_frames(places[1][rows], 1, seed, n_feat=args.feat, dim=8, **synth_kwargs(args))
    K = pnp.K_matrix(synth.FX, synth.FY, synth.CX, synth.CY)
    rv0, _ = cv2.Rodrigues(synth.R_LOCAL.T.copy())
    return dict(q=nv_b[rows], db=nv_a, k=args.topk, A=A, B=B, a_slot={int(f): i for i, f in enumerate(need_a)},
                b_slot={int(r): i for i, r in enumerate(rows)}, K=K, rv0=rv0.ravel(), tv0=np.zeros(3),
                iterations=args.iterations, pnp_flags=args.pnp_flags)


def run_reference(args):
    rank = env_int("RANK", 0)
    if rank != 0:
        return
    import multiprocessing as mp
    threads = os.cpu_count() or 1
    # sample sized to the pool: every worker gets a few hundred pairs per step, so start-up and scheduling noise amortise
    n_queries = min(args.frames, max(args.cpu_queries, 16 * threads))
    sample = make_cpu_sample(args, 2000, n_queries)
    times, t_nns = [], []
    done = 0
    with mp.get_context("fork").Pool(threads, initializer=_cpu_worker_init, initargs=(sample,)) as pool:
        for s in range(args.warmup + args.steps):
            n, sec, nq, t_nn, n_ok = cpu_reference_step(sample, threads, pool)
            if s >= args.warmup:
                times.append(sec)
                t_nns.append(t_nn)
                done += n
    total = sum(times)
    val = done / total
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / max(args.steps, 1), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u32+f64", "data": "synthetic", "config": workload_config(args),
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": "%d query rows x %d-row database NN (scipy cdist) + their top-%d = %d pairs through "
                                       "cv2.BFMatcher/solvePnPRansac per step; one pool of %d processes created outside the timed "
                                       "loop; NN %.0f ms + verification %.0f ms per step"
                                       % (n_queries, args.frames, args.topk, n, threads, 1e3 * sum(t_nns) / max(len(t_nns), 1),
                                          1e3 * (total - sum(t_nns)) / max(len(times), 1))},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    from multi_robot_slam_separators_b200 import capi, synth
    from multi_robot_slam_separators_b200.engine import Camera, SeparatorEngine

    rank, world, local_rank = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: CUDA device required (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)

    # ---- synthetic robots of this rank (seed = 1000*config + rank)
    t0 = time.time()
    A, B = synth.make_robots(2, args.frames, n_feat=args.feat, dim=args.dim, seed=2000 + rank, **synth_kwargs(args))
    t_gen = time.time() - t0
    eng = SeparatorEngine(Camera(synth.FX, synth.FY, synth.CX, synth.CY), args.feat, args.dim, args.frames, args.frames,
                          device=local_rank, iterations=args.iterations, pnp_flags=args.pnp_flags)
    eng.ctx.set_match_kernel(args.match)
    match_kernel = "popc" if args.match == "popc" else "tensor"
    P = args.frames * args.topk

    def pin(a):
        return torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()

    hostA = dict(desc=pin(A.desc), kp2d=pin(A.kp2d), kp3d=pin(A.kp3d), octave=pin(A.octave), count=pin(A.count))
    hostB = dict(desc=pin(B.desc), kp2d=pin(B.kp2d), kp3d=pin(B.kp3d), octave=pin(B.octave), count=pin(B.count))
    nvA, nvB = pin(A.netvlad), pin(B.netvlad)
    h2d_bytes = sum(v.nbytes for v in hostA.values()) + sum(v.nbytes for v in hostB.values()) + nvA.nbytes + nvB.nbytes + P * 8

    def upload():
        eng.add_peer_keyframes(hostA["desc"], hostA["kp2d"], hostA["kp3d"], hostA["count"], hostA["octave"])
        eng.add_local_keyframes(hostB["desc"], hostB["kp2d"], hostB["kp3d"], hostB["count"], hostB["octave"])
        eng.receive_descriptors(nvA)

    upload()
    d_q = torch.from_numpy(B.netvlad).to(dev)
    mode = capi.NN_TENSOR if args.nn == "tensor" else capi.NN_EXACT
    stream = eng.stream

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing -------------------------------------------------------------
    for _ in range(args.warmup):
        eng.step_dev(d_q, args.topk, mode)
    eng.sync()
    launches0 = eng.ctx.launch_count
    sampler, spath = clocks_sampler_start(local_rank)
    barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    with torch.cuda.stream(stream):
        ev[0].record(stream)
        for _ in range(args.steps):
            res = eng.step_dev(d_q, args.topk, mode)
        ev[1].record(stream)
    ev[1].synchronize()
    barrier()
    clocks = clocks_sampler_stop(sampler, spath)
    ms_total = ev[0].elapsed_time(ev[1])
    launches = eng.ctx.launch_count - launches0
    n_ok = int((res["status"] == 1).sum().item())
    n_ransac = int((res["n_matches"] >= 5).sum().item())
    iters_mean = float(res["iters_run"][res["n_matches"] >= 5].float().mean().item()) if n_ransac else 0.0
    m_mean = float(res["n_matches"][res["n_matches"] >= 5].float().mean().item()) if n_ransac else 0.0

    # ---- per-kernel durations (CUDA events on the launching stream) ---------------------------
    def timed(fn, reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            fn()
            e0.record(stream)
            for _ in range(reps):
                fn()
            e1.record(stream)
        e1.synchronize()
        return e0.elapsed_time(e1) / reps

    idx, _ = eng.candidates_dev(d_q, args.topk, mode)
    pairs = eng.pairs_from_candidates(idx)
    ms_k1 = timed(lambda: eng.candidates_dev(d_q, args.topk, mode), 10)
    import ctypes as C
    L = capi.lib()
    nm = torch.zeros(P, dtype=torch.int32, device=dev)
    nu = torch.zeros(P, dtype=torch.int32, device=dev)
    mf = torch.zeros((P, args.feat), dtype=torch.int16, device=dev)
    mt = torch.zeros((P, args.feat), dtype=torch.int16, device=dev)

    def k2():
        eng.ctx.check(L.sepb_match_pairs_dev(eng.ctx._h, eng.peer._h, eng.local._h, C.c_void_p(pairs.data_ptr()), P,
                                             C.byref(eng.mp), None, None, C.c_void_p(nm.data_ptr()),
                                             C.c_void_p(mf.data_ptr()), C.c_void_p(mt.data_ptr()), C.c_void_p(nu.data_ptr())))
    ms_k2 = timed(k2, 10)                 # (ten repetitions: K2 and K3 are within a few % of each other)
    ms_verify = timed(lambda: eng.verify_dev(pairs), 10)
    ms_two_pass = timed(lambda: eng.verify_dev(pairs, two_pass=True), 3)   # + second call of estimate_transformation
    ms_k3 = max(ms_verify - ms_k2, 0.0)
    # Vis/BundleAdjustment = 1 (off by default, outside the timed step): the same verification with the
    # 2-view bundle adjustment after PnP
    eng.ctx.set_bundle_adjustment(capi.ba_params(synth.BASELINE * synth.FY / synth.FX))
    ms_verify_ba = timed(lambda: eng.verify_dev(pairs), 3)
    n_ok_ba = int((eng.verify_dev(pairs)["status"] == 1).sum().item())
    eng.ctx.set_bundle_adjustment(None)

    # ---- end to end through the host-buffer C ABI -------------------------------------------
    # Every step copies BOTH robots' keyframes + descriptors from pinned host memory (H2D), searches, verifies and
    # reads the results back (D2H).  Two engines on one context act as a double buffer: the next step's keyframes
    # travel on the copy stream while the current step is verified (sepb_frames_upload_async).
    e2e_steps = max(1, args.steps if args.e2e_steps <= 0 else min(args.steps, args.e2e_steps))
    # the place-recognition databases of the two slots move to a second context whose stream has the device's highest
    # priority: the search of step s+1 (K1 + its small copies) then runs WHILE step s is verified -- its thread blocks
    # are dispatched ahead of the pending blocks of the K2 grid -- instead of behind it
    sctx = capi.Context(local_rank, high_priority=True) if args.e2e_search == "overlap" else eng.ctx
    eng_b = SeparatorEngine(Camera(synth.FX, synth.FY, synth.CX, synth.CY), args.feat, args.dim, args.frames, args.frames,
                            device=local_rank, ctx=eng.ctx, iterations=args.iterations, pnp_flags=args.pnp_flags,
                            search_ctx=sctx)
    if sctx is not eng.ctx:
        eng.db.close()
        eng.db = capi.NetvladDB(sctx, args.dim, args.frames)
        eng.search_ctx = sctx
    slots = [eng, eng_b]

    def pinned(shape, dtype):
        return torch.zeros(shape, dtype=getattr(torch, np.dtype(dtype).name)).pin_memory().numpy()

    outs = [capi.verify_buffers(P, args.feat, want_lists=False, alloc=pinned) for _ in slots]
    pairs_pin = [pinned((P, 2), np.int32) for _ in slots]
    rows = np.repeat(np.arange(args.frames, dtype=np.int32), args.topk)

    def descriptors(e):                # small H2D (10 MB): the peer's NetVLAD rows of this step replace the database
        e.db.clear()
        e.receive_descriptors(nvA)

    lazy = args.e2e_upload == "lazy"

    def keyframes(e):                  # copy stream, returns at once.  full: 1.13 GB (52 B per feature); lazy: descriptors +
        # counts only (32 B per feature, 0.64 GB), the device pulls key points of the frames whose pairs reach PnP
        e.n_local = e.n_peer = 0       # this step's keyframes replace the stores' content
        e.add_peer_keyframes(hostA["desc"], hostA["kp2d"], hostA["kp3d"], hostA["count"], hostA["octave"], async_=True, lazy=lazy)
        e.add_local_keyframes(hostB["desc"], hostB["kp2d"], hostB["kp3d"], hostB["count"], hostB["octave"], async_=True, lazy=lazy)

    def search(e, slot):               # blocking host-buffer call: H2D queries, K1, D2H candidates; then the pair list
        i5, d5 = e.db.topk(nvB, args.topk, mode)
        pairs_pin[slot][:, 0] = i5.reshape(-1)
        pairs_pin[slot][:, 1] = rows

    def verify(e, slot):               # enqueued: H2D pairs, K2 (+ key-point pulls), K3, D2H results to pinned buffers
        capi.verify_pairs(e.ctx, e.peer, e.local, pairs_pin[slot], e.pnp, want_lists=False, out=outs[slot], async_=True)

    # Software pipeline over two slots (engines on one context = double-buffered stores and databases).  Step s:
    #   verify(s) enqueued -> keyframes(s+1) on the copy stream -> descriptors(s+1) + search(s+1) on the search context
    #   (own high-priority stream: concurrent with verify(s); with --e2e-search serial: same stream, behind it) -> wait for
    #   step s's results.  The GPU only idles for the host time between the end of verify(s) and the enqueue of verify(s+1).
    for sl, e in enumerate(slots):
        descriptors(e)
        keyframes(e)
    search(slots[0], 0)
    verify(slots[0], 0)                # warm-up round
    eng.ctx.sync()
    keyframes(slots[0])                # a fresh binding for the first timed step (the warm-up pulled its key points)
    search(slots[0], 0)
    slots[0].local.wait()
    slots[0].peer.wait()
    barrier()
    t0 = time.perf_counter()
    ph = np.zeros(3)
    done = [torch.cuda.Event(), torch.cuda.Event()]
    deep = args.e2e_pipeline == "deep"
    if deep:
        # verify(s+1) is enqueued BEHIND verify(s) before the host waits for step s's results: the device goes from one
        # step's K3 straight into the next step's K2 (no host round trip between them)
        verify(slots[0], 0)
        done[0].record(eng.stream)
    for s_ in range(e2e_steps):
        c, n_ = s_ & 1, (s_ + 1) & 1
        p0 = time.perf_counter()
        if not deep:
            verify(slots[c], c)        # step s
        p1 = time.perf_counter()
        keyframes(slots[n_])           # step s+1's keyframes travel while step s is verified
        p2 = time.perf_counter()
        if s_ + 1 < e2e_steps:
            descriptors(slots[n_])     # step s+1: peer descriptors and candidate search
            search(slots[n_], n_)
            if deep:
                verify(slots[n_], n_)
                done[n_].record(eng.stream)
        if deep:
            done[c].synchronize()      # step s's results are in the caller's buffers
        else:
            eng.ctx.sync()
        p3 = time.perf_counter()
        ph += (p1 - p0, p2 - p1, p3 - p2)
        r_e2e = {k: v.copy() for k, v in outs[c].items()} if deep and s_ == e2e_steps - 1 else outs[c]
    slots[n_].local.wait()             # the last submitted upload completes inside the timed region
    slots[n_].peer.wait()
    torch.cuda.synchronize()
    t_e2e = time.perf_counter() - t0
    e2e_phases_ms = dict(zip(("verify_enqueue", "keyframes_enqueue", "results_and_next_search"), (ph / e2e_steps * 1e3).tolist()))
    d2h_bytes = P * (4 + 4 + 4 + 48 + 48) + P * 12
    if lazy:   # what really crossed the bus in the last step: descriptors + counts + search inputs + the pulled key points
        pulled = slots[c].peer.fetched_bytes + slots[c].local.fetched_bytes
        h2d_bytes = hostA["desc"].nbytes + hostB["desc"].nbytes + hostA["count"].nbytes + hostB["count"].nbytes + \
            nvA.nbytes + nvB.nbytes + P * 8 + pulled
    assert int((r_e2e["status"] == 1).sum()) == n_ok, "e2e and resident paths disagree"

    # ---- aggregate over ranks -----------------------------------------------------------------
    t = torch.tensor([ms_total, t_e2e * 1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total_max, ms_e2e_max = t.tolist()
    value = world * P * args.steps / (ms_total_max * 1e-3)
    e2e_value = world * P * e2e_steps / (ms_e2e_max * 1e-3)

    # ---- N > 1: the one path with a data-path collective (configs[2] shape: the peer database row-sharded over the
    # ranks, ncclAllGather of the top-k lists + merge inside its timed region), measured in the same job so that the
    # scaling record of the driver holds it; it runs after the headline's timed regions, on its own context
    sharded_line = None
    if world > 1 and not args.no_sharded_leg:
        import copy
        a2 = copy.copy(args)
        a2.steps, a2.warmup = max(3, min(args.steps, 10)), 3
        sharded_line = run_config2(a2, embed=True)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        popc_peak = eng.ctx.measure_peak("popc")
        dfma_peak = eng.ctx.measure_peak("dfma")
        bf16_burst = float(peaks.get("bf16_tflops", 1590.0))
        bf16_sust = float(peaks.get("bf16_tflops_sustained", 1400.0))
        peak_src = "MEASURED_PEAKS.json" if peaks else "fallback"
        # DRAM traffic per launch from the committed ncu --set full captures (only valid for the workload they were taken on)
        traffic = {}
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "r02_traffic.json")))
            w = tj["workload"]
            if (w["keyframes_per_robot"], w["features_per_frame"], w["netvlad_dim"], w["top_k"]) == \
                    (args.frames, args.feat, args.dim, args.topk):
                traffic = {k: v["bytes_per_launch"] for k, v in tj.items() if isinstance(v, dict) and "bytes_per_launch" in v}
        except Exception:
            pass
        k2_bytes = P * ((args.feat + args.feat) * 32 + args.feat * 12)        # algorithmic bytes per launch
        k2_gbs = k2_bytes / (ms_k2 * 1e-3) / 1e9
        k2_popc = P * 8.0 * args.feat * args.feat / (ms_k2 * 1e-3)
        k2_tops = P * 2.0 * 256 * args.feat * args.feat / (ms_k2 * 1e-3) / 1e12  # 256-term dot product per distance
        rooflines = {}
        if match_kernel == "tensor":
            rooflines["K2"] = {"kernel": "hamming_mma_persist_kernel (K2T, tcgen05 kind::i8)", "bound": "tensor", "achieved": k2_tops,
                               "peak": 2 * bf16_burst, "unit": "TOP/s", "frac": k2_tops / (2 * bf16_burst), "traffic": traffic.get("K2"),
                               "algorithmic_bytes": k2_bytes,
                               "peak_source": peak_src + ": 2 x measured bf16 burst (int8 dense issues at twice the bf16 rate; "
                                                         "kernel timed alone)",
                               "hbm_gbs": k2_gbs, "popc_equivalent_per_s": k2_popc,
                               # every s32 accumulator is read back once (tcgen05.ld): F_to x F_from x 4 B per pair against the
                               # measured TMEM read rate -- informational: the accumulator read-back is NOT what bounds K2T
                               "tmem_read": {"achieved_tbs": P * 4.0 * args.feat * args.feat / (ms_k2 * 1e-3) / 1e12,
                                             "peak_tbs": eng.ctx.measure_peak("tmem_pack16") / 1e12,
                                             "peak_source": "sepb_measure_peak: tcgen05.ld 32x32b.x16.pack::16b microbenchmark, same run"}}
        else:
            rooflines["K2"] = {"kernel": "hamming_match_kernel (K2, POPC pipe)", "bound": "hbm", "achieved": k2_gbs,
                               "peak": hbm_peak, "unit": "GB/s", "frac": k2_gbs / hbm_peak, "traffic": None,
                               "peak_source": peak_src, "note": "POPC-issue bound, not HBM bound",
                               "int_pipe": {"achieved": k2_popc, "peak": popc_peak, "unit": "popc32/s", "frac": k2_popc / popc_peak,
                                            "peak_source": "sepb_measure_peak microbenchmark, same run"}}
        if match_kernel == "tensor":
            # what one SM's tensor pipe sustains for K2T's instruction shape (cta_group::1, M128 x N256 x K32 int8, both
            # operands in shared memory): a bare tcgen05.mma issue loop, measured in the same run.  K2T issues 9 K-steps
            # per tile (8 descriptor + 1 train-index step) for 8 K-steps of algorithmic work.
            issue_peak = eng.ctx.measure_peak("mma_i8_n256") / 1e12
            rooflines["K2"]["tensor_issue"] = {"peak_tops": issue_peak, "achieved_tops_issued": k2_tops * 9.0 / 8.0,
                                               "frac_issued": k2_tops * 9.0 / 8.0 / issue_peak, "frac_algorithmic": k2_tops / issue_peak,
                                               "peak_source": "sepb_measure_peak mma_i8_n256 microbenchmark, same run"}
        if "tmem_read" in rooflines["K2"]:
            tr = rooflines["K2"]["tmem_read"]
            tr["frac"] = tr["achieved_tbs"] / tr["peak_tbs"]
        k1_tflops = 2.0 * args.frames * args.frames * args.dim / (ms_k1 * 1e-3) / 1e12
        rooflines["K1"] = {"kernel": "netvlad %s top-%d" % (args.nn, args.topk), "bound": "tensor", "achieved": k1_tflops,
                           "peak": bf16_burst, "unit": "TFLOP/s", "frac": k1_tflops / bf16_burst, "traffic": traffic.get("K1"),
                           "peak_source": peak_src + " (bf16 burst)", "algorithmic_flop": "2*Q*N*D",
                           "frac_of_sustained": k1_tflops / bf16_sust,
                           "note": "the 3-term split-bf16 filter issues 3x the algorithmic FLOP on the tensor pipe"}
        c_solve = 1.1e5 if args.pnp_flags == 0 else 2.5e3          # f64 ops of one 5-point EPnP / one 4-point P3P solve
        k3_ops = n_ransac * iters_mean * (c_solve + 30.0 * m_mean)   # H_eff * (C_solve + 30*M)
        ransac_pairs_per_s = world * n_ransac / max(ms_k3 * 1e-3, 1e-12)
        rooflines["K3"] = {"kernel": "pnp_persist_kernel (K3)", "bound": "fp64 issue", "achieved": k3_ops / max(ms_k3, 1e-9) / 1e-3,
                           "peak": dfma_peak, "unit": "f64 op/s", "frac": k3_ops / max(ms_k3, 1e-9) / 1e-3 / dfma_peak,
                           "peak_source": "sepb_measure_peak FP64 FMA microbenchmark, same run", "traffic": traffic.get("K3"),
                           "launches_per_step": 2, "note": "latency-bound by construction: OpenCV's serial f64 operation order is "
                           "replayed per hypothesis (bit-exact inlier sets); pnp_prepare_kernel + one persistent work-queue kernel"}
        # dominant kernel = the one with the largest share of the step
        dominant = max((("K1", ms_k1), ("K2", ms_k2), ("K3", ms_k3)), key=lambda kv: kv[1])[0]
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_total_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "s8/s32 (Hamming GEMM) + f64 (PnP, NetVLAD)" if match_kernel == "tensor" else "u32+f64",
                "data": "synthetic", "config": workload_config(args), "clocks": clocks,
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d_bytes), "d2h_bytes_per_step": int(d2h_bytes),
                        "steps": e2e_steps, "host_phases_ms": e2e_phases_ms,
                        "pipeline": ("the next step's verification is enqueued behind this one before the host waits for results; "
                                     if args.e2e_pipeline == "deep" else "") +
                                    "double-buffered keyframe stores: step s+1 uploads while step s is verified; its candidate search runs "
                                    + ("on a high-priority second stream, concurrently" if args.e2e_search == "overlap" else "behind it on the same stream"),
                        "upload": ("lazy key points: descriptors + counts of every frame are copied, kp2d / kp3d rows are pulled by the "
                                   "device from pinned host memory for the frames whose pairs reach PnP (sepb_frames_bind_host_async)")
                        if lazy else "full keyframes (52 B per feature)"},
                "gpu_launches": int(launches),
                "roofline": dict(rooflines[dominant], dominant=dominant, dominant_rule="largest share of the step; "
                                 "every kernel of the step is in `rooflines`"),
                "rooflines": rooflines,
                "kernels_ms": {"K1_netvlad_topk": ms_k1, "K2_hamming_match": ms_k2, "K3_pnp_ransac": ms_k3,
                               "step_share_K2": ms_k2 / (ms_total / args.steps)},
                "fp64_fma_peak_per_s": dfma_peak,
                "two_pass": {"ms_verify": ms_two_pass, "pairs_per_s": world * P / ((ms_k1 + ms_two_pass) * 1e-3),
                             "note": "whole estimate_transformation service call: K2 + K3, then K2G + K3 seeded with the first result "
                                     "(R/src/stereoCamGeometricTools.cpp:141-164); rank-0 timing x n_gpus"},
                "bundle_adjustment": {"ms_verify": ms_verify_ba, "ms_ba": max(ms_verify_ba - ms_verify, 0.0), "verified_ok": n_ok_ba,
                                      "note": "not in the timed step: verification with Vis/BundleAdjustment = 1 (5 + 20 LM iterations per "
                                              "verified pair, R/src/myRegistrationVis.cpp:1192-1370); rank-0 timing"},
                "nn_queries_per_s": args.frames / (ms_k1 * 1e-3), "nn_mode": args.nn, "match_kernel": match_kernel,
                "verified_ok_per_step": n_ok, "pairs_with_ransac_per_step": n_ransac, "ransac_iters_mean": iters_mean, "matches_per_ransac_pair_mean": m_mean,
                "ransac_pairs_per_s": ransac_pairs_per_s,
                "synth_seconds": t_gen}
        if not args.no_cpu_baseline and world == 1:
            sample = make_cpu_sample(args, 2000, args.cpu_queries)
            n, sec, nq, t_nn, _ = cpu_reference_step(sample, 1)
            line["cpu_baseline"] = {"value": n / sec, "unit": UNIT, "cores": 1, "kind": "port",
                                    "sample": "%d query rows NN (scipy cdist vs %d rows) + their top-%d = %d pairs through "
                                              "cv2.BFMatcher(k=2)+NNDR+uniqueness and cv2.solvePnPRansac, single thread"
                                              % (args.cpu_queries, args.frames, args.topk, n)}
        if sharded_line is not None:
            keep = ("metric", "value", "unit", "ms_per_step", "steps", "scaling", "e2e", "collective", "roofline",
                    "identical_to_single_database", "checked_queries", "gpu_launches")
            line["sharded_nn"] = dict({k: sharded_line[k] for k in keep if k in sharded_line},
                                      workload=sharded_line["config"]["workload"],
                                      db_rows_per_gpu=sharded_line["config"]["db_rows_per_gpu"])
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------
# configs[2]: 8 robots x 50k keyframes -- one robot's 50k queries against the union of the other seven databases
# (350k rows), row-sharded over the N GPUs, ONE NCCL all-gather of the top-k lists inside the timed region.
METRIC_NN = "NetVLAD NN queries/sec (row-sharded database, top-k, exact)"


def _dist_setup():
    import torch
    import torch.distributed as dist
    rank, world, local_rank = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: CUDA device required (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    return rank, world, local_rank


def _make_group(ctx, rank, world, dev):
    """sepb_group over the ranks of this job: rank 0's ncclGetUniqueId travels through torch.distributed (plumbing)."""
    import torch
    import torch.distributed as dist
    from multi_robot_slam_separators_b200 import capi
    uid = torch.zeros(128, dtype=torch.uint8, device=dev)
    if world > 1:
        if rank == 0:
            uid = torch.frombuffer(bytearray(capi.Group.unique_id()), dtype=torch.uint8).to(dev)
        dist.broadcast(uid, 0)
    return capi.Group(ctx=ctx, rank=rank, nranks=world, unique_id=uid.cpu().numpy().tobytes())


def sharded_nn_setup(rows, queries, dim, seed, rank, world):
    """Seeded NetVLAD rows of this rank's shard + the (replicated) queries: 30 % of the queries revisit a database place."""
    from multi_robot_slam_separators_b200 import sharding, synth
    lo, hi = sharding.shard_rows(rows, rank, world)
    db = synth.make_netvlad(np.arange(lo, hi, dtype=np.int64), dim, seed)
    rng = np.random.default_rng([seed, 1])
    qp = np.arange(rows, rows + queries, dtype=np.int64)
    rev = rng.random(queries) < 0.3
    qp[rev] = rng.integers(0, rows, int(rev.sum()))
    q = synth.make_netvlad(qp, dim, seed, noise=0.35)
    return lo, hi, db, q


def run_config2(args, embed=False):
    """embed=True: called from the configs[1] run at N > 1 (the driver's scaling run) so that its line carries the sharded
    search with the NCCL all-gather measured on the same box; returns the line instead of printing it."""
    import ctypes as C
    import torch
    import torch.distributed as dist
    from multi_robot_slam_separators_b200 import capi
    rank, world, local_rank = _dist_setup()
    dev = torch.device("cuda", local_rank)
    R, Q, D, k = args.db_rows, args.queries, args.dim, args.topk
    t0 = time.time()
    lo, hi, db_rows, q = sharded_nn_setup(R, Q, D, 3000, rank, world)
    t_gen = time.time() - t0
    ctx = capi.Context(local_rank)
    grp = _make_group(ctx, rank, world, dev)
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
    db = capi.NetvladDB(ctx, D, max(hi - lo, 1), base=lo)
    db.append(db_rows)
    q_pin = torch.from_numpy(q).pin_memory()
    d_q = q_pin.to(dev)
    out_i = torch.empty((Q, k), dtype=torch.int32, device=dev)
    out_d = torch.empty((Q, k), dtype=torch.float64, device=dev)
    loc_i, loc_d = torch.empty_like(out_i), torch.empty_like(out_d)
    mode = capi.NN_TENSOR if args.nn == "tensor" else capi.NN_EXACT
    L = capi.lib()

    def step():          # shard search (K1 + exact re-rank) -> ncclAllGather of the [Q,k] lists -> merge, all on ctx.stream
        grp.netvlad_topk_dev([db], [d_q.data_ptr()], Q, k, mode, [out_i.data_ptr()], [out_d.data_ptr()])

    def local_only():    # the same search without the exchange (what one shard costs)
        ctx.check(L.sepb_netvlad_topk_dev(ctx._h, db._h, C.c_void_p(d_q.data_ptr()), Q, k, mode, None, None, None,
                                          C.c_void_p(loc_i.data_ptr()), C.c_void_p(loc_d.data_ptr())))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            e0.record(stream)
            for _ in range(n):
                fn()
            e1.record(stream)
        e1.synchronize()
        return e0.elapsed_time(e1)

    for _ in range(args.warmup):
        step()
    ctx.sync()
    launches0 = ctx.launch_count
    sampler, spath = clocks_sampler_start(local_rank)
    barrier()
    ms_total = timed(step, args.steps)
    barrier()
    clocks = clocks_sampler_stop(sampler, spath)
    launches = ctx.launch_count - launches0
    gather_bytes = grp.last_gather_bytes
    for _ in range(2):
        local_only()
    ctx.sync()
    barrier()
    ms_local = timed(local_only, max(3, args.steps // 2)) / max(3, args.steps // 2)
    # ---- end to end: host queries in (pinned), merged lists out, through the host-buffer group call
    res_i = torch.empty((Q, k), dtype=torch.int32).pin_memory().numpy()
    res_d = torch.empty((Q, k), dtype=torch.float64).pin_memory().numpy()
    hq = q_pin.numpy()
    e2e_steps = max(1, args.steps)

    def e2e_step():
        grp.check(L.sepb_group_netvlad_topk(grp._h, grp._handles([db]), C.c_void_p(hq.ctypes.data), Q, k, mode,
                                            C.c_void_p(res_i.ctypes.data), C.c_void_p(res_d.ctypes.data)))
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    t_e2e = time.perf_counter() - t0
    assert np.array_equal(res_i, out_i.cpu().numpy()) and np.array_equal(res_d, out_d.cpu().numpy()), "e2e and resident paths disagree"
    # ---- identical to ONE database: every shard travels to rank 0 once (outside any timed region) and is searched whole
    ok = True
    if world > 1:
        sizes = [0] * world
        for r in range(world):
            from multi_robot_slam_separators_b200 import sharding
            a, b = sharding.shard_rows(R, r, world)
            sizes[r] = b - a
        mx = max(sizes)
        mine = torch.zeros((mx, D), dtype=torch.float64, device=dev)
        mine[: hi - lo] = torch.from_numpy(db_rows).to(dev)
        allrows = torch.empty((world, mx, D), dtype=torch.float64, device=dev)
        dist.all_gather_into_tensor(allrows.view(-1), mine.view(-1))
        if rank == 0:
            whole_rows = torch.cat([allrows[r, : sizes[r]] for r in range(world)]).cpu().numpy()
    else:
        whole_rows = db_rows
    if rank == 0:
        nchk = min(args.check_queries, Q)
        if world > 1:
            whole = capi.NetvladDB(ctx, D, R)
            whole.append(whole_rows)
            wi, wd = whole.topk(q[:nchk], k, capi.NN_EXACT)
            ok = bool(np.array_equal(res_i[:nchk], wi) and np.array_equal(res_d[:nchk], wd))
        from oracle import find_matches as fm      # checker only (a few hundred rows of the CPU oracle, untimed)
        oi, od = fm.topk(q[:256], whole_rows, k)
        ok = ok and bool(np.array_equal(res_i[:256], oi) and np.array_equal(res_d[:256], od))
    t = torch.tensor([ms_total, t_e2e * 1e3, ms_local], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total_max, ms_e2e_max, ms_local_max = t.tolist()
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        bf16 = float(peaks.get("bf16_tflops", 1590.0))
        ms_step = ms_total_max / args.steps
        flop = 2.0 * Q * (hi - lo) * D
        line = {"metric": METRIC_NN, "value": Q / (ms_step * 1e-3), "unit": "queries/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "bf16 split (tcgen05 filter) + f64 (exact re-rank)" if mode == capi.NN_TENSOR else "f64",
                "data": "synthetic",
                "config": {"workload": "configs[2]: one robot's %d queries against the other robots' %d keyframes (NetVLAD %d-d), "
                                       "database row-sharded over %d GPU(s), top-%d, NCCL all-gather of the candidate lists + merge"
                                       % (Q, R, D, world, k),
                           "db_rows_total": R, "db_rows_per_gpu": hi - lo, "queries": Q, "netvlad_dim": D, "top_k": k,
                           "l2_policy": "database shard + queries (%.0f MB f64 + split-bf16 copy) exceed L2 at N <= 2; at N >= 4 the "
                                        "shard fits L2 (stated, not flushed: the search is tensor-bound)" % ((hi - lo + Q) * D * 8 / 1e6),
                           "partition": "database rows by rank (sepb_db_set_base), queries replicated"},
                "clocks": clocks,
                "e2e": {"value": Q / (ms_e2e_max / e2e_steps * 1e-3), "unit": "queries/s", "h2d_bytes_per_step": int(Q * D * 8),
                        "d2h_bytes_per_step": int(Q * k * 12), "steps": e2e_steps},
                "gpu_launches": int(launches),
                "collective": {"op": "ncclAllGather (i32 idx + f64 dist, one fused group)", "bytes_received_per_rank_per_step": int(gather_bytes),
                               "ms_exchange_plus_merge": max(ms_step - ms_local_max, 0.0), "ms_shard_search": ms_local_max,
                               "nccl_version": capi.Group.nccl_version()},
                "roofline": {"kernel": "netvlad_gemm_kernel + split + re-rank (K1) on one shard", "bound": "tensor",
                             "achieved": flop / (ms_local_max * 1e-3) / 1e12, "peak": bf16, "unit": "TFLOP/s",
                             "frac": flop / (ms_local_max * 1e-3) / 1e12 / bf16, "traffic": None,
                             "peak_source": ("MEASURED_PEAKS.json" if peaks else "fallback") + " (bf16 burst); algorithmic 2*Q*N_shard*D, "
                                            "the 3-term split costs 3x that on the pipe"},
                "identical_to_single_database": ok, "checked_queries": int(min(args.check_queries, Q)), "synth_seconds": t_gen}
        if not args.no_cpu_baseline and world == 1 and not embed:
            from oracle import find_matches as fm
            n = min(args.cpu_queries, Q)
            t0 = time.perf_counter()
            fm.topk(q[:n], whole_rows, k)
            sec = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": n / sec, "unit": "queries/s", "cores": 1, "kind": "port",
                                    "sample": "%d queries x %d rows: scipy cdist (f64) + stable argsort top-%d, single thread" % (n, R, k)}
        if not embed:
            print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
    grp.close()
    ctx.close()
    if embed:
        if rank == 0:
            line["identical_to_single_database"] = ok
            return line
        return None
    if world > 1:
        dist.destroy_process_group()
    if not ok:
        raise SystemExit("sharded search differs from the single-database search")


def _cpu_nn_chunk(a):          # the database reaches the workers through fork (module global), not through a pipe
    from oracle import find_matches as fm
    i, n = a
    return fm.topk(_S["q"][i::n], _S["db"], _S["k"])[0].shape[0]


def run_reference_config2(args):
    if env_int("RANK", 0) != 0:
        return
    import multiprocessing as mp
    from multi_robot_slam_separators_b200 import synth
    threads = os.cpu_count() or 1
    R, D, k = args.db_rows, args.dim, args.topk
    n = min(args.cpu_queries * threads // 4 + threads, args.queries)
    global _S
    _, _, db, q = sharded_nn_setup(R, n, D, 3000, 0, 1)
    _S = dict(q=q, db=db, k=k)
    chunks = [(i, threads) for i in range(threads)]
    times = []
    with mp.get_context("fork").Pool(threads) as pool:          # created once, outside the timed loop
        for s in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            pool.map(_cpu_nn_chunk, chunks)
            if s >= args.warmup:
                times.append(time.perf_counter() - t0)
    val = n * len(times) / sum(times)
    print(json.dumps({"impl": "reference", "metric": METRIC_NN, "value": val, "unit": "queries/s", "n_gpus": args.gpus,
                      "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(times) / max(len(times), 1),
                      "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                      "config": {"workload": "configs[2]: %d-row database, top-%d" % (R, k)},
                      "cpu_baseline": {"value": val, "unit": "queries/s", "cores": threads, "kind": "port",
                                       "sample": "%d queries x %d rows per step: scipy cdist + argsort, %d processes" % (n, R, threads)},
                      "e2e": {"value": val, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}), flush=True)


# ---------------------------------------------------------------------------------------------
# configs[4]: batched geometric verification sweep, 1k - 1M candidate pairs from a resident frame store, pair list
# partitioned across ranks (no collective on the data path; strong scaling per batch size).  One JSON line per size,
# the LAST line (largest batch) is the headline of the run.
def run_config4(args):
    import torch
    import torch.distributed as dist
    from multi_robot_slam_separators_b200 import capi, sharding, synth
    from multi_robot_slam_separators_b200.engine import Camera, SeparatorEngine
    rank, world, local_rank = _dist_setup()
    dev = torch.device("cuda", local_rank)
    A, B = synth.make_robots(2, args.frames, n_feat=args.feat, dim=args.dim, seed=5000)      # same store on every rank
    eng = SeparatorEngine(Camera(synth.FX, synth.FY, synth.CX, synth.CY), args.feat, args.dim, args.frames, args.frames,
                          device=local_rank)
    eng.add_peer_keyframes(A.desc, A.kp2d, A.kp3d, A.count, A.octave)
    eng.add_local_keyframes(B.desc, B.kp2d, B.kp3d, B.count, B.octave)
    eng.receive_descriptors(A.netvlad)
    d_q = torch.from_numpy(B.netvlad).to(dev)
    idx, _ = eng.candidates_dev(d_q, args.topk, capi.NN_TENSOR)
    cand = eng.pairs_from_candidates(idx)            # top-k candidates per keyframe: the pool the sweep draws from
    eng.sync()
    g = torch.Generator(device="cpu").manual_seed(5)
    stream = eng.stream
    for P in [int(s) for s in args.sizes.split(",")]:
        pick = torch.randint(cand.shape[0], (P,), generator=g)          # same list on every rank
        lo, hi = sharding.pair_block(P, rank, world)
        mine = cand[pick[lo:hi].to(dev)].contiguous()
        for _ in range(max(args.warmup, 1)):
            res = eng.verify_dev(mine)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        launches0 = eng.ctx.launch_count
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            e0.record(stream)
            for _ in range(args.steps):
                res = eng.verify_dev(mine)
            e1.record(stream)
        e1.synchronize()
        run = res["n_matches"] >= 5
        t = torch.tensor([e0.elapsed_time(e1) / args.steps], dtype=torch.float64, device=dev)
        cnt = torch.tensor([float((res["status"] == 1).sum().item()), float(run.sum().item())], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
        if rank == 0:
            print(json.dumps({"metric": METRIC, "value": P / (t[0].item() * 1e-3), "unit": UNIT, "n_gpus": world,
                              "steps": args.steps, "warmup": args.warmup, "ms_per_step": t[0].item(), "higher_is_better": True,
                              "scaling": "strong", "vs_baseline": None, "dtype": "s8/s32 (Hamming GEMM) + f64 (PnP)",
                              "data": "synthetic", "gpu_launches": int(eng.ctx.launch_count - launches0),
                              "pairs": P, "pairs_per_gpu": hi - lo, "verified_ok": int(cnt[0].item()),
                              "pairs_with_ransac": int(cnt[1].item()),
                              "config": {"workload": "configs[4]: %d candidate pairs drawn from the top-%d candidates of 2 x %d keyframes "
                                                     "(F=%d), store replicated per GPU, pair list partitioned across ranks"
                                                     % (P, args.topk, args.frames, args.feat)}}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()



def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", type=int, default=1, choices=[1, 2, 3, 4],
                    help="BASELINE.json configs[N]: 1 = 2 x 10k keyframes on one GPU (the headline; default), 2 = row-sharded "
                         "NetVLAD database with the NCCL top-k gather, 3 = KITTI-sized frames (F=2000, D=4096, H=1000), "
                         "4 = verification sweep 1k..1M pairs")
    ap.add_argument("--frames", type=int, default=None)
    ap.add_argument("--feat", type=int, default=None)
    ap.add_argument("--dim", type=int, default=None)
    ap.add_argument("--iterations", type=int, default=None, help="RANSAC iterations (Vis/Iterations)")
    ap.add_argument("--topk", type=int, default=5)
    ap.add_argument("--mix", default="clean", choices=["clean", "hard"],
                    help="synthetic clutter: uniform random descriptors (clean: unrelated frames produce ~no matches) or a shared "
                         "vocabulary (hard: unrelated frames reach RANSAC with all-outlier sets and burn every iteration)")
    ap.add_argument("--pnp-flags", type=int, default=0, choices=[0, 2],
                    help="Vis/PnPFlags: 0 = ITERATIVE (the reference's default: 5-point EPnP minimal sets), 2 = P3P (4-point sets)")
    ap.add_argument("--nn", default="tensor", choices=["exact", "tensor"],
                    help="NetVLAD search: tcgen05 split-bf16 filter + exact f64 re-rank (tensor) or full f64 scan (exact); identical results")
    ap.add_argument("--match", default="auto", choices=["auto", "popc", "tensor"],
                    help="Hamming matcher kernel: tcgen05 int8 GEMM (tensor), POPC pipe (popc); identical results")
    ap.add_argument("--e2e-upload", default="lazy", choices=["lazy", "full"],
                    help="end-to-end leg: how keyframes reach the device (identical results)")
    ap.add_argument("--e2e-pipeline", default="simple", choices=["deep", "simple"],
                    help="end-to-end leg: deep = the next step's verification is enqueued before the host waits for this "
                         "step's results; simple = enqueue, wait, enqueue")
    ap.add_argument("--e2e-search", default="overlap", choices=["overlap", "serial"],
                    help="end-to-end leg: the next step's candidate search on a high-priority second context (overlap) or "
                         "queued behind the verification on the same stream (serial)")
    ap.add_argument("--e2e-steps", type=int, default=0, help="steps of the end-to-end leg (0 = --steps)")
    ap.add_argument("--cpu-queries", type=int, default=200)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--db-rows", type=int, default=350000, help="configs[2]: rows of the sharded database (7 peers x 50k)")
    ap.add_argument("--queries", type=int, default=50000, help="configs[2]: query descriptors per step")
    ap.add_argument("--check-queries", type=int, default=4096, help="configs[2]: queries compared with the unsharded search")
    ap.add_argument("--no-sharded-leg", action="store_true",
                    help="configs[1] at N > 1: skip the sharded-database leg (NCCL all-gather) appended to the line")
    ap.add_argument("--sizes", default="1000,10000,100000,1000000", help="configs[4]: batch sizes")
    args = ap.parse_args()
    # stdout carries the JSON line(s) only: libraries that write to fd 1 (NCCL's version banner under NCCL_DEBUG) go to
    # stderr from here on
    sys.stdout.flush()
    out_fd = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(out_fd, "w", buffering=1)
    defaults = {1: (10000, 1000, 128, 300), 2: (10000, 1000, 128, 300), 3: (4000, 2000, 4096, 1000), 4: (10000, 1000, 128, 300)}[args.config]
    for name, v in zip(("frames", "feat", "dim", "iterations"), defaults):
        if getattr(args, name) is None:
            setattr(args, name, v)
    if args.config == 2:
        (run_reference_config2 if args.impl == "reference" else run_config2)(args)
    elif args.config == 4 and args.impl == "ours":
        run_config4(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

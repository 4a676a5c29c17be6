"""Developer tool: K3 duration on the bench workload (verify - match), CUDA events on the library stream.
usage: python tools/k3_time.py [frames] [persistent|per_pair]"""
import ctypes as C
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from multi_robot_slam_separators_b200 import capi, synth
from multi_robot_slam_separators_b200.engine import Camera, SeparatorEngine

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
scheds = sys.argv[2].split(",") if len(sys.argv) > 2 else ["persistent"]     # persistent | nohelp | per_pair
hard = float(os.environ.get("SEPB_HARD_NEG", "0"))
import pickle
cache = "/tmp/k3_synth_%d_%g.pkl" % (frames, hard)      # the 18 s of numpy are paid once per box
if os.path.exists(cache):
    A, B = pickle.load(open(cache, "rb"))
else:
    A, B = synth.make_robots(2, frames, n_feat=1000, dim=128, seed=2000, **({"hard_negatives": hard} if hard else {}))
    pickle.dump((A, B), open(cache, "wb"), protocol=4)
eng = SeparatorEngine(Camera(synth.FX, synth.FY, synth.CX, synth.CY), 1000, 128, frames, frames)
eng.add_peer_keyframes(A.desc, A.kp2d, A.kp3d, A.count, A.octave)
eng.add_local_keyframes(B.desc, B.kp2d, B.kp3d, B.count, B.octave)
eng.receive_descriptors(A.netvlad)
d_q = torch.from_numpy(B.netvlad).cuda()
idx, _ = eng.candidates_dev(d_q, 5, capi.NN_TENSOR)
pairs = eng.pairs_from_candidates(idx)
stream = eng.stream
P = pairs.shape[0]
L = capi.lib()


def timed(fn, reps=3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        fn()
        e0.record(stream)
        for _ in range(reps):
            fn()
        e1.record(stream)
    e1.synchronize()
    return e0.elapsed_time(e1) / reps


nm = torch.zeros(P, dtype=torch.int32, device="cuda")
nu = torch.zeros(P, dtype=torch.int32, device="cuda")
mf = torch.zeros((P, 1000), dtype=torch.int16, device="cuda")
mt = torch.zeros((P, 1000), dtype=torch.int16, device="cuda")


def k2():
    eng.ctx.check(L.sepb_match_pairs_dev(eng.ctx._h, eng.peer._h, eng.local._h, C.c_void_p(pairs.data_ptr()), P,
                                         C.byref(eng.mp), None, None, C.c_void_p(nm.data_ptr()),
                                         C.c_void_p(mf.data_ptr()), C.c_void_p(mt.data_ptr()), C.c_void_p(nu.data_ptr())))


ms_k2 = timed(k2)
for sched in scheds:
    eng.ctx.set_pnp_scheduler("per_pair" if sched == "per_pair" else "persistent")
    os.environ.pop("SEPB_K3_NO_HELP", None)
    if sched == "nohelp":
        os.environ["SEPB_K3_NO_HELP"] = "1"       # read by developer builds (SEPB_DEVELOPER=1) at every launch
    ms_verify = timed(lambda: eng.verify_dev(pairs))
    res = eng.verify_dev(pairs)
    eng.sync()
    run = res["n_matches"] >= 5
    print("K3_DEFS=%r sched=%s: verify %.3f ms, K2 %.3f ms, K3 (+cov) %.3f ms; ok %d, ransac pairs %d, iters mean %.1f, checksum %d"
          % (os.environ.get("SEPB_K3_DEFS", ""), sched, ms_verify, ms_k2, ms_verify - ms_k2, int((res["status"] == 1).sum()),
             int(run.sum()), float(res["iters_run"][run].float().mean()) if int(run.sum()) else 0.0,
             int(res["n_inliers"].sum().item()) * 7 + int(res["iters_run"].sum().item())), flush=True)

if os.environ.get('SEPB_SHORT'):
    sys.exit(0)
L.sepb_debug_k3_prof.argtypes = [C.c_void_p, C.c_int]
out = (C.c_ulonglong * 8)()
L.sepb_debug_k3_prof(out, 1)
eng.verify_dev(pairs); eng.sync()
if L.sepb_debug_k3_prof(out, 1):
    v = np.array(list(out), dtype=np.float64); ch = max(v[4], 1)
    print("  chunks %d; per chunk kcycles: epnp5 %.1f (jacobi %.1f) scoring %.1f; refit per pair %.1f"
          % (ch, v[0] / ch / 1e3, v[5] / ch / 1e3, v[2] / ch / 1e3, v[3] / max(int(run.sum()), 1) / 1e3))

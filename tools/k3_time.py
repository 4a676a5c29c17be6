"""Developer tool: K3 duration on the bench workload (verify - match), CUDA events on the library stream."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from multi_robot_slam_separators_b200 import capi, synth
from multi_robot_slam_separators_b200.engine import Camera, SeparatorEngine

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
A, B = synth.make_robots(2, frames, n_feat=1000, dim=128, seed=2000)
eng = SeparatorEngine(Camera(synth.FX, synth.FY, synth.CX, synth.CY), 1000, 128, frames, frames)
eng.add_peer_keyframes(A.desc, A.kp2d, A.kp3d, A.count, A.octave)
eng.add_local_keyframes(B.desc, B.kp2d, B.kp3d, B.count, B.octave)
eng.receive_descriptors(A.netvlad)
d_q = torch.from_numpy(B.netvlad).cuda()
idx, _ = eng.candidates_dev(d_q, 5, capi.NN_TENSOR)
pairs = eng.pairs_from_candidates(idx)
stream = eng.stream


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


ms_verify = timed(lambda: eng.verify_dev(pairs))
res = eng.verify_dev(pairs)
eng.sync()
print("K3_DEFS=%r verify %.3f ms (K2 + K3 + covariance), ok %d" % (os.environ.get("SEPB_K3_DEFS", ""), ms_verify, int((res["status"] == 1).sum())))

if os.environ.get('SEPB_SHORT'):
    sys.exit(0)
# clocks under a sustained K3 load (a per-warp cycle count that improves while the wall time does not = clock drop)
import subprocess, time
p = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_event_reasons.sw_power_cap", "--format=csv,noheader,nounits",
                      "-lms", "100"], stdout=subprocess.PIPE, text=True)
t0 = time.time()
n = 0
while time.time() - t0 < 2.5:
    eng.verify_dev(pairs)
    eng.sync()
    n += 1
p.terminate()
rows = [l.strip().split(", ") for l in p.stdout.read().splitlines() if l.strip()]
mhz = [float(r[0]) for r in rows[3:]]
print("  sustained: %.3f ms per verify over %d calls; SM MHz median %.0f min %.0f; power max %.0f W; power cap active: %s"
      % (1e3 * (time.time() - t0) / n, n, np.median(mhz), min(mhz), max(float(r[1]) for r in rows), any(r[2] == "Active" for r in rows)))
L = capi.lib()
import ctypes as C
L.sepb_debug_k3_prof.argtypes = [C.c_void_p, C.c_int]
out = (C.c_ulonglong * 8)()
L.sepb_debug_k3_prof(out, 1)
eng.verify_dev(pairs); eng.sync()
if L.sepb_debug_k3_prof(out, 1):
    v = np.array(list(out), dtype=np.float64); ch = max(v[4], 1)
    print("  per chunk kcycles: epnp5 %.1f (jacobi %.1f) scoring %.1f; refit per pair %.1f" % (v[0] / ch / 1e3, v[5] / ch / 1e3, v[2] / ch / 1e3, v[3] / 3000 / 1e3))

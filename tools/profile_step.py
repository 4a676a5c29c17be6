"""Minimal driver for ncu: a few resident steps of the hot path (no e2e / CPU legs)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from multi_robot_slam_separators_b200 import capi, synth
from multi_robot_slam_separators_b200.engine import Camera, SeparatorEngine

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
mode = capi.NN_TENSOR if (len(sys.argv) > 3 and sys.argv[3] == "tensor") else capi.NN_EXACT
A, B = synth.make_robots(2, frames, n_feat=1000, dim=128, seed=2000)
eng = SeparatorEngine(Camera(synth.FX, synth.FY, synth.CX, synth.CY), 1000, 128, frames, frames)
eng.add_peer_keyframes(A.desc, A.kp2d, A.kp3d, A.count, A.octave)
eng.add_local_keyframes(B.desc, B.kp2d, B.kp3d, B.count, B.octave)
eng.receive_descriptors(A.netvlad)
d_q = torch.from_numpy(B.netvlad).cuda()
for _ in range(steps):
    res = eng.step_dev(d_q, 5, mode)
eng.sync()
print("ok", int((res["status"] == 1).sum()))

"""BASELINE.json configs[4]: batched geometric verification sweep, 1k - 1M candidate pairs drawn from a resident
20 000-frame store (2 robots x --frames keyframes), at N GPUs: the pair list is partitioned across ranks
(sharding.pair_block, no collective on the data path), store replicated per GPU (strong scaling per batch size).
One JSON line per batch size: verified pairs/s (K2 + K3 + covariance, CUDA events on the library stream, max over ranks).
  python tools/sweep_pairs.py                       (1 GPU)
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/sweep_pairs.py"""
import argparse, json, os, sys
import numpy as np, torch, torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from multi_robot_slam_separators_b200 import capi, sharding, synth  # noqa: E402
from multi_robot_slam_separators_b200.engine import Camera, SeparatorEngine  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=10000)
ap.add_argument("--sizes", default="1000,10000,100000,1000000")
ap.add_argument("--reps", type=int, default=3)
args = ap.parse_args()
rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
dev = torch.device("cuda", lr)
A, B = synth.make_robots(2, args.frames, n_feat=1000, dim=128, seed=5000)      # same store on every rank
eng = SeparatorEngine(Camera(synth.FX, synth.FY, synth.CX, synth.CY), 1000, 128, args.frames, args.frames, device=lr)
eng.add_peer_keyframes(A.desc, A.kp2d, A.kp3d, A.count, A.octave)
eng.add_local_keyframes(B.desc, B.kp2d, B.kp3d, B.count, B.octave)
eng.receive_descriptors(A.netvlad)
d_q = torch.from_numpy(B.netvlad).to(dev)
idx, _ = eng.candidates_dev(d_q, 5, capi.NN_TENSOR)
cand = eng.pairs_from_candidates(idx)            # 5 candidates per keyframe: the pool the sweep draws from
eng.sync()
g = torch.Generator(device="cpu").manual_seed(5)
for P in [int(s) for s in args.sizes.split(",")]:
    pick = torch.randint(cand.shape[0], (P,), generator=g)          # same list on every rank
    lo, hi = sharding.pair_block(P, rank, world)
    mine = cand[pick[lo:hi].to(dev)].contiguous()
    stream = eng.stream
    with torch.cuda.stream(stream):
        res = eng.verify_dev(mine)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(args.reps):
            res = eng.verify_dev(mine)
        e1.record(stream)
    e1.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / args.reps, float((res["status"] == 1).sum().item())], dtype=torch.float64, device=dev)
    ms, ok = t.clone(), t.clone()
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(ok, op=dist.ReduceOp.SUM)
    if rank == 0:
        print(json.dumps({"metric": "verified keyframe pairs/sec (match+PnP-RANSAC)", "value": P / (ms[0].item() * 1e-3),
                          "unit": "pairs/s", "n_gpus": world, "pairs": P, "pairs_per_gpu": hi - lo, "ms": ms[0].item(),
                          "verified_ok": int(ok[1].item()), "scaling": "strong",
                          "config": "configs[4]: pairs drawn from the top-5 candidates of 2 x %d keyframes, store replicated per GPU" % args.frames}),
              flush=True)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()

"""Row-sharded NetVLAD search over N GPUs (BASELINE.json configs[2] shape, scaled by --rows / --queries):
every rank owns a slice of the peer database, searches it with global ids, one NCCL all-gather of the [Q,k]
lists, device merge.  Checks the merged result against the same search over the whole database on rank 0 and
prints one JSON line (queries/s, max over ranks, CUDA events).
  torchrun --nproc-per-node N --master-addr 127.0.0.1 tools/sharded_nn_check.py --rows 350000 --queries 50000"""
import argparse, json, os, sys, ctypes as C
import numpy as np, torch, torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from multi_robot_slam_separators_b200 import capi, sharding, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=350000)
ap.add_argument("--queries", type=int, default=50000)
ap.add_argument("--dim", type=int, default=128)
ap.add_argument("--k", type=int, default=5)
ap.add_argument("--steps", type=int, default=5)
ap.add_argument("--check-queries", type=int, default=4096)
args = ap.parse_args()
rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
dev = torch.device("cuda", lr)
# the same seeded descriptors on every rank (cheap: NetVLAD vectors only)
rng = np.random.default_rng([3000, 1])
n_places = max(args.rows, args.queries)
places = synth.place_assignment(2, n_places, 0.3, rng)
db_all = synth.make_netvlad(places[0][: args.rows], args.dim, 3000)
q_all = synth.make_netvlad(places[1][: args.queries], args.dim, 3000)
lo, hi = sharding.shard_rows(args.rows, rank, world)
ctx = capi.Context(lr)
stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
db = capi.NetvladDB(ctx, args.dim, max(hi - lo, 1), base=lo)
db.append(db_all[lo:hi])
d_q = torch.from_numpy(q_all).to(dev)
Q, k = args.queries, args.k
idx = torch.empty((Q, k), dtype=torch.int32, device=dev)
dd = torch.empty((Q, k), dtype=torch.float64, device=dev)
merge = sharding.merge_on_device(ctx)
L = capi.lib()


def local():
    ctx.check(L.sepb_netvlad_topk_dev(ctx._h, db._h, C.c_void_p(d_q.data_ptr()), Q, k, capi.NN_TENSOR, None, None, None,
                                      C.c_void_p(idx.data_ptr()), C.c_void_p(dd.data_ptr())))
    return idx, dd


def step():
    if world > 1:
        return sharding.sharded_topk(local, merge)
    return local()


with torch.cuda.stream(stream):          # the collective orders itself against the library's stream
    for _ in range(2):
        out_i, out_d = step()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0.record(stream)
    for _ in range(args.steps):
        out_i, out_d = step()
    e1.record(stream)
e1.synchronize()
ms = torch.tensor([e0.elapsed_time(e1) / args.steps], device=dev, dtype=torch.float64)
if world > 1:
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
ok = True
if rank == 0:
    nchk = min(args.check_queries, Q)
    full = capi.NetvladDB(ctx, args.dim, args.rows)
    full.append(db_all)
    ri, rd = full.topk(q_all[:nchk], k=k, mode=capi.NN_EXACT)
    ok = bool(np.array_equal(out_i[:nchk].cpu().numpy(), ri) and np.array_equal(out_d[:nchk].cpu().numpy(), rd))
    print(json.dumps({"metric": "NetVLAD NN queries/sec (row-sharded database, top-%d)" % k, "value": Q / (ms.item() * 1e-3),
                      "unit": "queries/s", "n_gpus": world, "ms_per_step": ms.item(), "rows_total": args.rows,
                      "rows_per_gpu": hi - lo, "queries": Q, "dim": args.dim,
                      "collective": "all_gather of [Q,k] (i32 idx + f64 dist) = %d bytes per rank" % (Q * k * 12),
                      "identical_to_single_database": ok, "checked_queries": nchk}), flush=True)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
sys.exit(0 if ok else 1)

"""Host topology probe for the e2e leg: which NUMA node each GPU hangs off, and what pinned-memory H2D bandwidth GPU 0 sees
from buffers first-touched on each node (the thread is bound to the node's cores before the allocation)."""
import glob
import os
import subprocess
import sys
import time

import torch


def cpulist(s):
    out = []
    for part in s.strip().split(","):
        if not part:
            continue
        a, _, b = part.partition("-")
        out += list(range(int(a), int(b or a) + 1))
    return out


def main():
    print(subprocess.run("nvidia-smi topo -m; lscpu | grep -i -E 'numa|socket|model name|^CPU\\(s\\)'; nproc", shell=True,
                         capture_output=True, text=True).stdout)
    nodes = {}
    for p in sorted(glob.glob("/sys/devices/system/node/node[0-9]*")):
        nodes[int(p.rsplit("node", 1)[1])] = cpulist(open(p + "/cpulist").read())
    print("nodes:", {k: (v[0], v[-1], len(v)) for k, v in nodes.items() if v})
    allowed = os.sched_getaffinity(0)
    print("allowed cpus:", len(allowed))
    for d in range(torch.cuda.device_count()):
        pr = torch.cuda.get_device_properties(d)
        bus = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        try:
            nn = open("/sys/bus/pci/devices/%s/numa_node" % bus).read().strip()
        except OSError as e:
            nn = "?(%s)" % e
        print("gpu", d, bus, "numa_node", nn)
    torch.cuda.set_device(0)
    dst = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")
    for node, cpus in nodes.items():
        cpus = [c for c in cpus if c in allowed]
        if not cpus:
            continue
        os.sched_setaffinity(0, cpus)
        src = torch.empty(1 << 30, dtype=torch.uint8).pin_memory()
        src.fill_(1)
        for _ in range(2):
            dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            dst.copy_(src, non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        print("node %d -> gpu0: %.1f GB/s" % (node, 5 * (1 << 30) / e0.elapsed_time(e1) / 1e6))
        del src
    os.sched_setaffinity(0, allowed)
    # aggregate host -> device bandwidth with 1, 2, 4, 8 GPUs copying at once (what bounds the end-to-end leg at N = 8)
    ng = torch.cuda.device_count()
    size = 1 << 29
    srcs = [torch.empty(size, dtype=torch.uint8).pin_memory() for _ in range(ng)]
    dsts = [torch.empty(size, dtype=torch.uint8, device="cuda:%d" % d) for d in range(ng)]
    for k in sorted({1, 2, 4, 8, ng}):
        if k > ng:
            continue
        for rep in range(2):
            for d in range(k):
                torch.cuda.synchronize(d)
            t0 = time.perf_counter()
            for _ in range(4):
                for d in range(k):
                    dsts[d].copy_(srcs[d], non_blocking=True)
            for d in range(k):
                torch.cuda.synchronize(d)
            dt = time.perf_counter() - t0
        print("%d GPUs copying at once: %.1f GB/s in aggregate (%.1f per GPU)" % (k, 4 * k * size / dt / 1e9, 4 * size / dt / 1e9))


if __name__ == "__main__":
    sys.exit(main())

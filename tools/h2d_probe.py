#!/usr/bin/env python3
"""Per-rank pinned host<->device copy bandwidth with ALL ranks copying at the same time (the feed of the e2e path):
   python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/h2d_probe.py [--bind]
Rank 0 prints one JSON object: per-rank H2D / D2H GB/s (149 MB / 41 MB pieces = one bench step), the host topology each rank sees
and whether it was bound to its GPU's local CPUs (orbb200_bind_host_to_device)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb-slam2_b200"))
import torch
import torch.distributed as dist

rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
bound = 0
if "--bind" in sys.argv:
    import orbb200
    bound = orbb200.lib().orbb200_bind_host_to_device(lr)
nb_in, nb_out = 320 * 1241 * 376, 41 * 1000 * 1000
h_in = torch.empty(nb_in, dtype=torch.uint8).pin_memory()
h_in.random_(0, 255)
d_in = torch.empty(nb_in, dtype=torch.uint8, device=dev)
d_out = torch.empty(nb_out, dtype=torch.uint8, device=dev)
h_out = torch.empty(nb_out, dtype=torch.uint8).pin_memory()
res = {}
for name, fn, nb in (("h2d", lambda: d_in.copy_(h_in, non_blocking=True), nb_in), ("d2h", lambda: h_out.copy_(d_out, non_blocking=True), nb_out),
                     ("both", lambda: (d_in.copy_(h_in, non_blocking=True), None), nb_in)):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    s2 = torch.cuda.Stream()
    t0 = time.perf_counter()
    for _ in range(8):
        fn()
        if name == "both":
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    res[name] = 8 * (nb + (nb_out if name == "both" else 0)) / dt / 1e9
try:
    cpus = sorted(os.sched_getaffinity(0))
    aff = "%d cpus [%d..%d]" % (len(cpus), cpus[0], cpus[-1])
except Exception:
    aff = "?"
mine = {"rank": rank, "gbs": res, "affinity": aff, "bound_cpus": bound}
if world > 1:
    allr = [None] * world
    dist.all_gather_object(allr, mine)
else:
    allr = [mine]
if rank == 0:
    numa = ""
    try:
        numa = open("/sys/devices/system/node/online").read().strip()
    except Exception:
        pass
    print(json.dumps({"world": world, "bind": "--bind" in sys.argv, "host_cpus": os.cpu_count(), "numa_nodes_online": numa, "ranks": allr,
                      "sum_h2d_gbs": sum(r["gbs"]["h2d"] for r in allr), "sum_both_gbs": sum(r["gbs"]["both"] for r in allr)}))
if world > 1:
    dist.destroy_process_group()

"""Host<->device copy bandwidth of N GPUs at once (is the end-to-end number limited by the box or by us?).
usage: python -m torch.distributed.run --nproc-per-node N tools/probe_pcie.py
Each rank copies 1 GiB pinned buffers H2D only, D2H only, then both directions on two streams; the phases start at
the same wall-clock instants on every rank (gloo barrier), rank 0 prints per-rank and total GB/s."""
import os
import time

import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("MASTER_PORT", "29533")
dist.init_process_group("gloo", rank=rank, world_size=world)
torch.cuda.set_device(local)
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from quade_b200 import affinity  # noqa: E402
loc = affinity.gpu_locality(local)
if os.environ.get("PROBE_AFFINITY"):            # pin the rank (and its first-touch pinned pages) next to its GPU
    loc = affinity.bind_to_gpu(local, world)
infos = [None] * world
dist.all_gather_object(infos, {k: v for k, v in loc.items() if k != "local_cpus"})
if rank == 0:
    print("binding:", "on" if os.environ.get("PROBE_AFFINITY") else "off", infos, flush=True)
n = 1 << 30
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
h_in.fill_(1)
h_out.fill_(2)
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.ones(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def phase(h2d, d2h, reps=8):
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        if h2d:
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    gbs = torch.tensor([reps * n * (int(h2d) + int(d2h)) / dt / 1e9], dtype=torch.float64)
    alls = [torch.zeros(1, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(alls, gbs)
    return [round(float(a), 1) for a in alls]


for name, a, b in (("h2d", True, False), ("d2h", False, True), ("both", True, True)):
    phase(a, b, 2)
    r = phase(a, b)
    if rank == 0:
        print(name, "per rank GB/s", r, "total", round(sum(r), 1), flush=True)
dist.destroy_process_group()

import sys, time, os
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from quade_b200 import DemuxContext, PinnedBuffer
from quade_b200.synth import Synth
import os
s = Synth.from_config(os.environ.get("QUADE_CFG", "cfg3")); P = 2_000_000
sizes = s.sizes(0, P)
host = [np.empty(n, dtype=np.uint8) for n in sizes]
s.fill(0, P, host, n_threads=16)
dev = [torch.empty(n + 64, dtype=torch.uint8, device="cuda") for n in sizes]
for t, h in zip(dev, host): t[:h.size].copy_(torch.from_numpy(h))
ctx = DemuxContext(s.dual, s.spans(), s.min_qual, s.sample_keys(), device=0, max_stream_bytes=max(sizes) + 4096, max_pairs=P + 1024, n_slots=int(os.environ.get("QUADE_SLOTS", "2")))
ptrs = [t.data_ptr() for t in dev]
for rep in range(3):
    ts = []; tc = []
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for i in range(16):
        if ctx.in_flight == int(os.environ.get("QUADE_SLOTS", "2")):
            a = time.perf_counter(); r = ctx.collect(); tc.append(time.perf_counter() - a)
        a = time.perf_counter(); ctx.submit_device(ptrs[:len(sizes)], sizes, last=True, ticket=i); ts.append(time.perf_counter() - a)
    while ctx.in_flight:
        a = time.perf_counter(); r = ctx.collect(); tc.append(time.perf_counter() - a)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print("rep", rep, "total ms/step %.3f" % (dt / 16 * 1e3), "submit ms", [round(x * 1e3, 2) for x in ts], "collect ms", [round(x * 1e3, 2) for x in tc], "kernels", round(r.ms_kernels, 3), {k: round(v, 3) for k, v in r.ms_kernel.items()})
print("reframed_tiles", r.reframed_tiles)

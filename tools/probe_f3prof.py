"""Debug: per-phase cycle shares of k_frame3 (library built with QUADE_BUILD_DEFS=-DQUADE_F3_PROF)."""
import ctypes, os, sys
sys.path.insert(0, "/root/repo")
os.environ.setdefault("QUADE_FRAME", "3")
import numpy as np, torch
from quade_b200 import DemuxContext, _lib
from quade_b200.synth import Synth
s = Synth.from_config("cfg3"); P = 2_000_000
sizes = s.sizes(0, P)
host = [np.empty(n, dtype=np.uint8) for n in sizes]
s.fill(0, P, host, n_threads=16)
dev = [torch.empty(n + 64, dtype=torch.uint8, device="cuda") for n in sizes]
for t, h in zip(dev, host): t[:h.size].copy_(torch.from_numpy(h))
ctx = DemuxContext(s.dual, s.spans(), s.min_qual, s.sample_keys(), device=0, max_stream_bytes=max(sizes) + 4096, max_pairs=P + 1024, n_slots=2)
ptrs = [t.data_ptr() for t in dev]
L = _lib.lib()
out = (ctypes.c_ulonglong * 16)()
for rep in range(3):
    ctx.submit_device(ptrs, sizes, last=True, ticket=rep); r = ctx.collect()
    L.quade_debug_f3_prof(out)
    v = list(out)
    names = (["tma wait", "masks+B1", "scan+B2", "list+B3", "lookback|parse", "B4 wait", "store+B5", "take+B6"] if os.environ["QUADE_FRAME"] == "3"
             else ["tma wait", "masks+bar", "scan+bar", "publish+list", "k0 wait", "parse+store", "bar", "issue+bar"])
    tile_bytes = 8192 if os.environ["QUADE_FRAME"] == "5" else 16384
    if int(os.environ["QUADE_FRAME"]) >= 7:
        tiles = sum(sizes) / 16384
        nm = ["scan:full wait", "scan:work", "lb:go wait", "lb:lookback", "parse:listed wait", "parse:k0 wait", "parse:work", "parse:take"]
        print("rep", rep, "frame ms %.3f" % r.ms_kernel["frame"], " | ".join("%s %.0f" % (n, x / tiles) for n, x in zip(nm, v[:8])), "(cycles per tile)")
        continue
    for w in range(2):
        tot = sum(v[w * 8:w * 8 + 8])
        print("rep", rep, "warp", w, "frame ms %.3f" % r.ms_kernel["frame"], " ".join("%s %.1f%%" % (n, 100.0 * x / tot) for n, x in zip(names, v[w * 8:w * 8 + 8])), "cycles/tile %.0f" % (tot / (sum(sizes) / tile_bytes)))

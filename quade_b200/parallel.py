"""Multi-GPU plumbing: one process per GPU (torchrun), chunks sharded round-robin, one all-reduce of the counters.

Chunks (the n-th entries of seq_R1 / seq_R2 / index_R1[/ index_R2], src/Quade.py:198,229) are independent in
the reference: generators are re-created per chunk (:203-206) and the skipped-record desynchronisation never
crosses a chunk.  The only cross-chunk state is (i) the additive counters (src/Sample.py:32,144) and (ii) the
append order inside each bin file = chunk order (src/FastqWriter.py:68-69).  So rank r demultiplexes chunks
r, r+W, r+2W, ... on its own GPU, every bin file is written as one `.partNNNNNN` per chunk, the counters are
summed with ONE all-reduce (NCCL on the GPUs; gloo in the CPU tests) and the parts are concatenated in chunk
order on the host (concatenated gzip members are a gzip stream).  No read data ever crosses GPUs.
"""
import glob
import os
import shutil

import numpy as np


def env_world():
    """(rank, world_size, local_rank) from the torchrun environment; (0, 1, 0) when launched plainly."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def init_process_group(backend=None):
    """Join the job's process group if WORLD_SIZE > 1 (idempotent).  Returns (rank, world, local_rank)."""
    rank, world, local_rank = env_world()
    if world > 1:
        import torch
        import torch.distributed as dist
        if not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29533")
            if backend is None:
                backend = "nccl" if torch.cuda.is_available() else "gloo"
            kw = {}
            if backend == "nccl":
                torch.cuda.set_device(local_rank)
                kw["device_id"] = torch.device("cuda", local_rank)
            dist.init_process_group(backend, rank=rank, world_size=world, **kw)
    return rank, world, local_rank


def _dist():
    try:
        import torch.distributed as dist
        return dist if dist.is_available() and dist.is_initialized() else None
    except Exception:
        return None


def chunks_of_rank(n_chunks, rank, world):
    """Round-robin shard: chunk c runs on rank c mod world."""
    return list(range(rank, n_chunks, world))


def allreduce_counters(counters, device=None):
    """Sum the per-bin pair counters over all ranks: the path's only collective (int64[1+2S])."""
    c = np.ascontiguousarray(counters, dtype=np.int64)
    dist = _dist()
    if dist is None or dist.get_world_size() == 1:
        return c.copy()
    import torch
    t = torch.from_numpy(c.copy())
    if dist.get_backend() == "nccl":
        t = t.to(torch.device("cuda", device if device is not None else torch.cuda.current_device()))
    dist.all_reduce(t)
    return t.cpu().numpy()


def any_rank_failed(failed, device=None):
    """True on every rank if some rank reports a failure (one tiny all-reduce: the ranks agree before they meet in
    the counters' all-reduce or a barrier, so that one rank's exception does not hang the others)."""
    return bool(allreduce_counters(np.array([1 if failed else 0], dtype=np.int64), device)[0])


def barrier():
    dist = _dist()
    if dist is not None and dist.get_world_size() > 1:
        dist.barrier()


def merge_parts(final_names, rank=0, world=1, directory="."):
    """Restore output order by chunk index: `<name>.part000000`, `.part000001`, ... -> `<name>`.

    A bin file exists in the end iff some chunk created a part for it (lazy creation, FastqWriter.py:55-57).
    The list of names is split over the ranks so that the concatenation runs in parallel.  Returns the names
    this rank produced.
    """
    made = []
    for i, name in enumerate(sorted(final_names)):
        if i % world != rank:
            continue
        parts = sorted(glob.glob(os.path.join(directory, glob.escape(name) + ".part*")))
        if not parts:
            continue
        with open(os.path.join(directory, name), "wb") as out:
            for p in parts:
                with open(p, "rb") as fh:
                    shutil.copyfileobj(fh, out, 4 << 20)
        for p in parts:
            os.remove(p)
        made.append(name)
    return made

#!/usr/bin/env python
"""bench.py -- pairs/s of the demultiplexing hot path on N B200s (one process per GPU) + roofline + CPU baseline.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config cfg3] [--pairs P] [--impl ours|reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A step = one batch of P synthetic read pairs per GPU through parse -> key -> lookup -> min-phred -> compaction.
`value`: inputs resident in HBM, outputs left in HBM.  `e2e`: the same batches from pinned HOST buffers through
quade_submit/quade_collect, H2D and D2H inside the timed region, double-buffered.  Weak scaling: every rank
demultiplexes its own chunks; the only collective is one NCCL all-reduce of the report counters.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "paired reads/sec demuxed (dual index)"
UNIT = "pairs/s"
WORKLOADS = {
    "cfg2": "synthetic paired-end 2x150, single 8-bp index, 96 samples, min phred 25",
    "cfg3": "synthetic paired-end 2x150, dual index 8+8 with 6-bp molecular index, 384 samples, min phred 25",
    "cfg4": "synthetic paired-end 2x150, dual 10+10 index, 384 samples, min phred 25",
    "cfg5": "synthetic paired-end 2x150, dual 8+8 + 6-bp molecular, 384 samples, 30% undetermined, min phred 25",
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="cfg3", choices=sorted(WORKLOADS))
    ap.add_argument("--pairs", type=int, default=2_000_000, help="read pairs per step per GPU")
    ap.add_argument("--min-qual", type=int, default=None)
    ap.add_argument("--cpu-pairs", type=int, default=1_000_000, help="sample size of the CPU baseline leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--slots", type=int, default=6,
                    help="batch slots in flight (submit/collect ring): the device queue rides out host hiccups of (slots-1) steps")
    ap.add_argument("--spinup-ms", type=float, default=400.0,
                    help="untimed extra warm-up before the W warm-up steps: the GPU sits idle while the inputs are generated "
                         "on the host and needs tens of ms of load to be back at full clocks")
    return ap.parse_args()


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel`, from the committed ncu --set full summary
    of this same command (profiles/r01_traffic.json, written by tools/ncu_traffic.py); None if absent."""
    try:
        with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as fh:
            return json.load(fh)["kernels"][kernel]["dram_bytes_per_launch"]
    except Exception:
        return None


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region.  ONE long-running `nvidia-smi -lms` process,
    started before the big pinned allocations: forking this process once it maps gigabytes of pinned memory takes
    milliseconds with the GIL held, which stalled the submit loop inside the timed region (2.0 -> 2.3-3.6 ms per
    step when the sampler forked per sample).  Period 100 ms: at 25 ms the driver queries themselves delayed kernel
    launches (2 of 6 runs +18 %); a region shorter than the period reports the nearest sample."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.rows, self.t0, self.t1 = [], None, None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", os.environ.get("QUADE_BENCH_SMI_MS", "100")],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
        self.start()

    def run(self):
        if self.proc is None:
            return
        for line in self.proc.stdout:                      # blocking read: releases the GIL
            parts = [x.strip() for x in line.strip().split(",")]
            if len(parts) >= 6:
                self.rows.append((time.time(), parts))

    def begin(self):
        self.t0 = time.time()

    def stop(self):
        self.t1 = time.time()
        time.sleep(0.06)                                   # one more sample after the region
        if self.proc is not None:
            self.proc.terminate()
        self.join(timeout=3)
        rows = [p for (t, p) in self.rows if self.t0 is not None and self.t0 <= t <= self.t1 + 0.05]
        if not rows and self.rows:                         # region shorter than the sampling period: nearest sample
            rows = [min(self.rows, key=lambda tp: abs(tp[0] - self.t1))[1]]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(float(r[0]) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][1]), "reasons": reasons, "samples": len(rows)}


class CpuLeg:
    """The CPU restatement of the reference path (oracle/, C, one chunk per thread) on a bounded sample."""

    def __init__(self, synth, min_qual, n_pairs, threads):
        from oracle import oracle as O
        self.O = O
        self.threads = max(1, threads)
        self.n_chunks = self.threads * 2
        self.per = max(1000, n_pairs // self.n_chunks)
        self.chunks = [synth.generate(c * self.per, self.per) for c in range(self.n_chunks)]
        self.params = O.OracleParams(synth.dual, synth.spans(), min_qual, synth.sample_keys())

    def run(self):
        pairs, secs, out_bytes = self.O.demux_chunks_mt(self.params, self.chunks, self.threads)
        return pairs, secs

    def describe(self):
        return ("%d chunks x %d pairs, one chunk per thread, C restatement of the reference loop "
                "(oracle/quade_oracle.c), inputs and outputs in RAM" % (self.n_chunks, self.per))


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU algorithm on the host cores (oracle port; the reference itself is
    Python 2.7 and /root/reference does not exist on the GPU box).  Rank 0 only."""
    if rank != 0:
        return
    from quade_b200.synth import Synth
    synth = Synth.from_config(args.config)
    min_qual = synth.min_qual if args.min_qual is None else args.min_qual
    cores = os.cpu_count() or 1
    sample = min(args.pairs, max(100_000, args.cpu_pairs))
    times, pairs_done = [], 0
    leg = CpuLeg(synth, min_qual, sample, cores)
    for it in range(args.warmup + args.steps):
        pairs, secs = leg.run()
        if it >= args.warmup:
            times.append(secs)
            pairs_done += pairs
    total = sum(times)
    value = pairs_done / total
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * total / max(1, args.steps), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic", "impl": "reference",
            "config": {"workload": WORKLOADS[args.config], "config": args.config, "pairs_per_step": leg.n_chunks * leg.per},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": leg.describe() + " per step"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from quade_b200 import DemuxContext, PinnedBuffer, _lib
    from quade_b200.synth import Synth
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    sampler = ClockSampler(local_rank)                 # forked now, while this process is still small
    synth = Synth.from_config(args.config)
    min_qual = synth.min_qual if args.min_qual is None else args.min_qual
    P, ns = args.pairs, synth.n_streams

    # this rank's chunk: pairs [rank*P, rank*P + P) of the data set, generated into pinned host buffers
    sizes = synth.sizes(rank * P, P)
    pinned = [PinnedBuffer(n) for n in sizes]
    threads = max(1, (os.cpu_count() or 1) // world)
    synth.fill(rank * P, P, [b.array for b in pinned], n_threads=min(64, threads))
    dev_in = [torch.empty(n + 64, dtype=torch.uint8, device=dev) for n in sizes]
    for t, b in zip(dev_in, pinned):
        t[:b.nbytes].copy_(torch.from_numpy(b.array), non_blocking=False)
    ctx = DemuxContext(synth.dual, synth.spans(), min_qual, synth.sample_keys(), device=local_rank,
                       max_stream_bytes=max(sizes) + 4096, max_pairs=P + 1024, max_records=P + P // 8, n_slots=args.slots)
    stream = torch.cuda.ExternalStream(ctx.compute_stream(), device=dev)
    ptrs = [t.data_ptr() for t in dev_in]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    diag = []                                              # (seconds, what) of every host call, QUADE_BENCH_DIAG=1

    def device_steps(n):
        """n batches back to back, args.slots in flight; returns the collected results."""
        res = []
        for i in range(n):
            if ctx.in_flight == args.slots:
                t = time.perf_counter()
                res.append(ctx.collect())
                diag.append((time.perf_counter() - t, "collect"))
            t = time.perf_counter()
            ctx.submit_device(ptrs, sizes, last=True, ticket=i)
            diag.append((time.perf_counter() - t, "submit"))
        while ctx.in_flight:
            res.append(ctx.collect())
        return res

    # ---- value: HBM-resident inputs, kernels only --------------------------------------------------------
    def timed_block(n):
        """What the timed region runs: n steps, then the report counters summed over the ranks."""
        res = device_steps(n)
        cnt = torch.from_numpy(ctx.counters()).to(dev)
        if world > 1:
            dist.all_reduce(cnt)           # the path's only collective: the report counters (NCCL)
        return res, cnt

    # untimed: spin-up, then the W warm-up steps through the very statements of the timed region (first use of a
    # code path on a fresh box pages in library text and loads CUDA modules lazily: tens of ms)
    t_spin = time.perf_counter()
    while (time.perf_counter() - t_spin) * 1e3 < args.spinup_ms:
        device_steps(4)                    # no collective here: the number of rounds is wall-clock bound per rank
    barrier()
    timed_block(args.warmup)
    barrier()
    sampler.begin()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    del diag[:]
    results, counters = timed_block(args.steps)
    ev1.record(stream)
    if os.environ.get("QUADE_BENCH_DIAG"):
        print("diag: slowest host calls in the timed region (ms):",
              [(round(a * 1e3, 2), w) for a, w in sorted(diag, reverse=True)[:4]], file=sys.stderr)
    barrier()
    clocks = sampler.stop()
    ms_total = ev0.elapsed_time(ev1)
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    r0 = results[-1]
    assert all(r.n_pairs == P for r in results), "a batch did not demultiplex every pair"
    launches = sum(r.gpu_launches for r in results)
    value = world * P * args.steps / (ms_total * 1e-3)

    # ---- roofline of the dominant kernel (CUDA events on the compute stream, averaged over the timed steps) --
    kms = {k: sum(r.ms_kernel[k] for r in results) / len(results) for k in _lib.K_NAMES}
    in_bytes, out_bytes = r0.in_bytes, r0.out_bytes
    insert_in = sizes[0] + sizes[1]
    alg = {"frame": in_bytes, "compact": insert_in + out_bytes, "classify": sum(sizes[2:])}
    dom = max(("frame", "compact", "classify"), key=lambda k: kms[k])
    kname = {"frame": "k_frame6", "compact": "k_compact3", "classify": "k_classify2"}[dom]
    peak, peak_src = peaks()
    ach = alg[dom] / (kms[dom] * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": kname, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": ncu_traffic(kname), "peak_source": peak_src, "algorithmic_bytes_per_launch": alg[dom],
                "ms_per_launch": kms[dom],
                "per_kernel": {k: {"ms": kms[k], "GB/s": (alg[k] / (kms[k] * 1e-3) / 1e9) if k in alg else None,
                                   "frac": (alg[k] / (kms[k] * 1e-3) / 1e9 / peak) if k in alg else None} for k in kms},
                "pipeline": {"algorithmic_bytes_per_step": in_bytes + out_bytes,
                             "achieved": (in_bytes + out_bytes) * args.steps / (ms_total * 1e-3) / 1e9,
                             "frac": (in_bytes + out_bytes) * args.steps / (ms_total * 1e-3) / 1e9 / peak}}

    # ---- e2e: pinned host buffers -> quade_submit (H2D) -> kernels -> quade_collect (D2H), double-buffered --
    e2e = None
    if not args.no_e2e:
        host = [b.array for b in pinned]

        def host_steps(n):
            done = 0
            for i in range(n):
                if ctx.in_flight == min(args.slots, 3):                # PCIe-bound: a deeper ring only adds pinned memory
                    r = ctx.collect()
                    done += int(r.arena(0)[-1]) * 0 + r.n_pairs      # touches the copied-back result
                ctx.submit(host, last=True, ticket=i)
            while ctx.in_flight:
                r = ctx.collect()
                done += int(r.arena(1)[-1]) * 0 + r.n_pairs
            return done
        host_steps(max(args.slots + 1, args.warmup))           # every slot's pinned output arena exists before timing
        barrier()
        t0 = time.perf_counter()
        done = host_steps(args.steps)
        barrier()
        dt = time.perf_counter() - t0
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        assert done == P * args.steps
        e2e = {"value": world * P * args.steps / float(tt.item()), "unit": UNIT,
               "h2d_bytes_per_step": int(sum(sizes)), "d2h_bytes_per_step": int(out_bytes),
               "timing": "wall clock between device synchronisations, max over ranks",
               "pcie_GBps_per_gpu": (sum(sizes) + out_bytes) * args.steps / float(tt.item()) / 1e9}

    # ---- CPU baseline: rank 0, N=1 only ---------------------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        leg = CpuLeg(synth, min_qual, min(args.cpu_pairs, P), cores)
        leg.run()
        reps, secs, pairs = 0, 0.0, 0
        while secs < 5.0 and reps < 20:
            pr, sc = leg.run()
            pairs, secs, reps = pairs + pr, secs + sc, reps + 1
        cpu = {"value": pairs / secs, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": leg.describe() + ", %d repetitions" % reps}

    if rank == 0:
        und = int(counters[0].item())
        tot = int(counters.sum().item())
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u8", "data": "synthetic",
                "config": {"workload": WORKLOADS[args.config], "config": args.config, "pairs_per_step_per_gpu": P,
                           "min_qual": min_qual, "in_bytes_per_pair": in_bytes / P, "out_bytes_per_pair": out_bytes / P,
                           "l2": "inputs (%.0f MB per step) larger than the 126 MB L2" % (in_bytes / 1e6),
                           "parallelism": "chunks sharded over %d GPU(s), counters all-reduced (NCCL)" % world,
                           "undetermined_frac": und / max(1, tot)},
                "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
                "clocks": clocks}
        emit(line)
    ctx.close()


_JSON_OUT = None


def emit(line):
    """The ONE JSON line goes to the real stdout; everything else a library prints (NCCL's version banner under
    NCCL_DEBUG=VERSION, torchrun notices) has been pointed at stderr by main()."""
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    global _JSON_OUT
    args = parse_args()
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")      # keep the real stdout for the JSON line ...
    os.dup2(2, 1)                              # ... and send fd 1 (C libraries included) to stderr
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world > 1:
        import torch
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local_rank))
    try:
        run_ours(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()

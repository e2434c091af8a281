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
    ap.add_argument("--ref-pairs", type=int, default=8_000, help="pairs per process and step of the reference-Python legs")
    ap.add_argument("--distinct", type=int, default=0,
                    help="distinct device-generated chunks the steps cycle through (0 = one per step incl. warm-up, capped by --distinct-gb)")
    ap.add_argument("--distinct-gb", type=float, default=48.0, help="HBM budget for the resident input chunks")
    ap.add_argument("--no-bind", action="store_true", help="do not bind the rank's threads and pinned pages next to its GPU")
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
    of this same command (profiles/r02_traffic.json, written by tools/ncu_traffic.py); None if absent."""
    try:
        with open(os.path.join(ROOT, "profiles", "r02_traffic.json")) as fh:
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
    """The C restatement of the reference path (oracle/quade_oracle.c, one chunk per thread) on a bounded sample:
    kind "port"."""

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


class RefLeg:
    """The reference's OWN code (a-slide/Quade src/*.py, unmodified on disk under oracle/_ref, run under Python 3 by
    oracle/ref_shim.py) on a bounded sample: `procs` independent single-threaded processes, one chunk of `per`
    pairs each, plain FASTQ in, fastq.gz out, local scratch.  The clock is the reference's own `Done in` line
    (src/Quade.py:172,192); a step takes as long as its slowest process.  kind "reference"."""

    def __init__(self, synth, min_qual, per, procs):
        import tempfile
        from oracle import ref_shim
        if not ref_shim.available():
            raise RuntimeError("reference sources not found (oracle/make_ref.py copies them to oracle/_ref)")
        self.per, self.procs = per, procs
        self.root = tempfile.mkdtemp(prefix="quade_ref_")
        self.dirs = []
        for c in range(procs):
            d = os.path.join(self.root, "p%02d" % c)
            os.makedirs(d)
            arrays = synth.generate(c * per, per)
            names = []
            for k, a in enumerate(arrays):
                fn = os.path.join(d, "in_%d.fastq" % k)
                a.tofile(fn)
                names.append([fn])
            with open(os.path.join(d, "conf.txt"), "w") as fh:
                fh.write(synth.conf_text(names, min_qual))
            self.dirs.append(d)

    def run(self):
        """One step: every process demultiplexes its chunk once.  Returns (pairs, seconds)."""
        env = dict(os.environ, PYTHONPATH=ROOT)
        ps = [subprocess.Popen([sys.executable, "-m", "oracle.ref_shim", os.path.join(d, "conf.txt")], cwd=d, env=env,
                               stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for d in self.dirs]
        worst = 0.0
        for p in ps:
            out, _ = p.communicate()
            secs = None
            for line in out.splitlines():
                if line.startswith("Done in ") and line.endswith("s"):
                    secs = float(line[len("Done in "):-1])
            if p.returncode != 0 or secs is None:
                raise RuntimeError("the reference failed:\n" + out[-2000:])
            worst = max(worst, secs)
        return self.per * self.procs, worst

    def close(self):
        import shutil
        shutil.rmtree(self.root, ignore_errors=True)

    def describe(self):
        return ("%d process(es) x %d pairs, the reference's own Python (oracle/_ref, %s through oracle/ref_shim.py), one "
                "thread each, plain FASTQ in, fastq.gz out, its own 'Done in' clock" % (
                    self.procs, self.per, "python%d.%d" % sys.version_info[:2]))


def host_info():
    model = "?"
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                model = line.split(":", 1)[1].strip()
                break
    except Exception:
        pass
    return {"nproc": os.cpu_count() or 1, "cpu_model": model}


def workload_config(args, synth, min_qual, sizes):
    """The `config` object: the workload only, the same for both arms."""
    return {"workload": WORKLOADS[args.config], "config": args.config, "pairs_per_step_per_gpu": args.pairs,
            "min_qual": min_qual,
            "l2": "inputs (%.0f MB per step) larger than the 126 MB L2" % (sum(sizes) / 1e6),
            "parallelism": "chunks sharded over %d GPU(s), counters all-reduced (NCCL)" % args.gpus}


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation on the host cores.  The reference is pure Python
    and single-threaded, so "all the host threads it can use" is one independent process per core, each on its own
    chunk (src/Quade.py:198-206: chunks are independent).  Falls back to the C port only if oracle/_ref is absent.
    Rank 0 only; nothing of the product library is loaded here."""
    if rank != 0:
        return
    from quade_b200.synth import Synth
    from oracle import ref_shim
    synth = Synth.from_config(args.config)
    min_qual = synth.min_qual if args.min_qual is None else args.min_qual
    cores = os.cpu_count() or 1
    sizes = synth.sizes(0, args.pairs)
    port_leg = CpuLeg(synth, min_qual, min(args.pairs, max(100_000, args.cpu_pairs)), cores)
    port_leg.run()
    pp, ps = port_leg.run()
    port = {"value": pp / ps, "unit": UNIT, "cores": cores, "kind": "port", "sample": port_leg.describe()}
    times, pairs_done = [], 0
    if ref_shim.available():
        leg = RefLeg(synth, min_qual, args.ref_pairs, cores)
        try:
            for it in range(args.warmup + args.steps):
                pairs, secs = leg.run()
                if it >= args.warmup:
                    times.append(secs)
                    pairs_done += pairs
        finally:
            leg.close()
        kind, sample = "reference", leg.describe() + " per step"
    else:
        leg = port_leg
        for it in range(args.warmup + args.steps):
            pairs, secs = leg.run()
            if it >= args.warmup:
                times.append(secs)
                pairs_done += pairs
        kind, sample = "port", leg.describe() + " per step (oracle/_ref absent: reference sources were not materialised)"
    total = sum(times)
    value = pairs_done / total
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * total / max(1, args.steps), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic", "impl": "reference",
            "config": workload_config(args, synth, min_qual, sizes),
            "cpu_baseline": dict({"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample}, **host_info()),
            "cpu_baseline_port": port,
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
    from quade_b200 import affinity
    placement = affinity.bind_to_gpu(local_rank, world) if not args.no_bind else {"mode": "off"}   # before any pinned allocation
    synth = Synth.from_config(args.config)
    min_qual = synth.min_qual if args.min_qual is None else args.min_qual
    P, ns = args.pairs, synth.n_streams

    # ---- inputs of `value`: D DISTINCT chunks of P pairs, generated in HBM by the device-side generator (the bytes
    # are those of the host generator: tests/test_gpu_parity.py).  Rank r owns chunks r*D .. r*D + D - 1 of the data set.
    sizes0 = synth.sizes(rank * P, P)
    D = args.distinct or (args.steps + args.warmup)
    D = max(1, min(D, int(args.distinct_gb * 1e9 // max(1, sum(sizes0)))))
    chunks = []
    for j in range(D):
        first = (rank * D + j) * P
        sz = synth.sizes(first, P)
        bufs = [torch.empty(n + 64, dtype=torch.uint8, device=dev) for n in sz]
        got = synth.fill_device(first, P, [t.data_ptr() for t in bufs])
        assert got == sz, (got, sz)
        chunks.append(([t.data_ptr() for t in bufs], sz, bufs))
    cap = max(max(c[1]) for c in chunks) + 4096
    ctx = DemuxContext(synth.dual, synth.spans(), min_qual, synth.sample_keys(), device=local_rank, max_stream_bytes=cap,
                       max_pairs=P + 1024, max_records=P + P // 8, n_slots=args.slots, device_only=True)
    stream = torch.cuda.ExternalStream(ctx.compute_stream(), device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    diag = []                                              # (seconds, what) of every host call, QUADE_BENCH_DIAG=1

    def device_steps(n, start=0):
        """n batches back to back, args.slots in flight, step i on chunk (start + i) mod D; returns the results."""
        res = []
        for i in range(n):
            if ctx.in_flight == args.slots:
                t = time.perf_counter()
                res.append(ctx.collect())
                diag.append((time.perf_counter() - t, "collect"))
            t = time.perf_counter()
            ptrs, sz, _ = chunks[(start + i) % D]
            ctx.submit_device(ptrs, sz, last=True, ticket=i)
            diag.append((time.perf_counter() - t, "submit"))
        while ctx.in_flight:
            res.append(ctx.collect())
        return res

    # ---- value: HBM-resident inputs, kernels only --------------------------------------------------------
    def timed_block(n, start=0):
        """What the timed region runs: n steps, then the report counters summed over the ranks."""
        res = device_steps(n, start)
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
    ctx.reset_counters()
    barrier()
    sampler.begin()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    del diag[:]
    results, counters = timed_block(args.steps, args.warmup)
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
    assert all(r.n_pairs == P for r in results), "a batch did not demultiplex every pair"
    launches = sum(r.gpu_launches for r in results)
    reframed = sum(r.reframed_tiles for r in results)
    value = world * P * args.steps / (ms_total * 1e-3)

    # ---- roofline of the dominant kernel (CUDA events on the compute stream, averaged over the timed steps) --
    nres = len(results)
    kms = {k: sum(r.ms_kernel[k] for r in results) / nres for k in _lib.K_NAMES}
    in_bytes = sum(r.in_bytes for r in results) / nres
    out_bytes = sum(r.out_bytes for r in results) / nres
    insert_in = sum(chunks[(args.warmup + i) % D][1][0] + chunks[(args.warmup + i) % D][1][1] for i in range(args.steps)) / args.steps
    index_in = in_bytes - insert_in
    alg = {"frame": in_bytes, "compact": insert_in + out_bytes, "classify": index_in}
    dom = max(("frame", "compact", "classify"), key=lambda k: kms[k])
    kname = {"frame": "k_frame", "compact": "k_compact", "classify": "k_classify"}[dom]
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
    und = int(counters[0].item())
    tot = int(counters.sum().item())
    ctx.close()
    del chunks
    torch.cuda.empty_cache()

    # ---- e2e: pinned host buffers -> quade_submit (H2D) -> kernels -> quade_collect (D2H), double-buffered --
    e2e = None
    if not args.no_e2e:
        e_slots = min(args.slots, 3)                           # PCIe-bound: a deeper ring only adds pinned memory
        pinned = [PinnedBuffer(n) for n in sizes0]
        threads = max(1, (os.cpu_count() or 1) // world)
        synth.fill(rank * P, P, [b.array for b in pinned], n_threads=min(64, threads))
        host = [b.array for b in pinned]
        ectx = DemuxContext(synth.dual, synth.spans(), min_qual, synth.sample_keys(), device=local_rank,
                            max_stream_bytes=max(sizes0) + 4096, max_pairs=P + 1024, max_records=P + P // 8, n_slots=e_slots)
        e_out = [0]

        def host_steps(n):
            done = 0
            for i in range(n):
                if ectx.in_flight == e_slots:
                    r = ectx.collect()
                    done += int(r.arena(0)[-1]) * 0 + r.n_pairs      # touches the copied-back result
                    e_out[0] = r.out_bytes
                ectx.submit(host, last=True, ticket=i)
            while ectx.in_flight:
                r = ectx.collect()
                done += int(r.arena(1)[-1]) * 0 + r.n_pairs
                e_out[0] = r.out_bytes
            return done
        host_steps(max(e_slots + 1, args.warmup))
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
               "h2d_bytes_per_step": int(sum(sizes0)), "d2h_bytes_per_step": int(e_out[0]),
               "timing": "wall clock between device synchronisations, max over ranks",
               "pcie_GBps_per_gpu": (sum(sizes0) + e_out[0]) * args.steps / float(tt.item()) / 1e9,
               "placement": placement}
        ectx.close()

    # ---- CPU baseline: rank 0, N=1 only: the reference's own Python on one core; the C port on all cores beside it --
    cpu = port = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import ref_shim
        cores = os.cpu_count() or 1
        leg = CpuLeg(synth, min_qual, min(args.cpu_pairs, P), cores)
        leg.run()
        reps, secs, pairs = 0, 0.0, 0
        while secs < 3.0 and reps < 20:
            pr, sc = leg.run()
            pairs, secs, reps = pairs + pr, secs + sc, reps + 1
        port = {"value": pairs / secs, "unit": UNIT, "cores": cores, "kind": "port",
                "sample": leg.describe() + ", %d repetitions" % reps}
        if ref_shim.available():
            rleg = RefLeg(synth, min_qual, 2 * args.ref_pairs, 1)
            try:
                rleg.run()
                pr, sc = rleg.run()
            finally:
                rleg.close()
            cpu = dict({"value": pr / sc, "unit": UNIT, "cores": 1, "kind": "reference", "sample": rleg.describe()}, **host_info())
        else:
            cpu = dict(port, **host_info())

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u8", "data": "synthetic",
                "config": workload_config(args, synth, min_qual, sizes0),
                "workload_stats": {"in_bytes_per_pair": in_bytes / P, "out_bytes_per_pair": out_bytes / P,
                                   "undetermined_frac": und / max(1, tot), "distinct_chunks_per_gpu": D,
                                   "inputs": "generated in HBM by the device-side generator, step i on chunk i mod %d" % D,
                                   "reframed_tiles_per_step": reframed / nres},
                "roofline": roofline, "cpu_baseline": cpu, "cpu_baseline_port": port, "e2e": e2e, "gpu_launches": int(launches),
                "clocks": clocks}
        emit(line)


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

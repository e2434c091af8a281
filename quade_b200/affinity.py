"""CPU / memory placement of one rank next to its GPU.

The end-to-end path is host-bound (pinned staging buffers, H2D / D2H copies, gzip): a rank whose threads and
page-locked pages sit on the other socket pays the inter-socket link on every copy.  bind_to_gpu() restricts the
calling process (and the threads it starts afterwards) to the CPUs the kernel reports as local to the GPU
(/sys/bus/pci/devices/<bdf>/local_cpulist, numa_node); page-locked memory allocated AFTER the call is first-touched
on that node.  Where the box gives no locality (numa_node -1, one node: the usual VM) the CPUs are split evenly over
the local ranks instead, so that the ranks' reader / writer threads do not pile onto the same cores.
"""
import os


def _parse_cpulist(text):
    cpus = []
    for part in text.strip().split(","):
        if not part:
            continue
        a, _, b = part.partition("-")
        cpus += list(range(int(a), int(b or a) + 1))
    return cpus


def gpu_locality(index):
    """{'bdf', 'numa_node', 'local_cpus'} of CUDA device `index` as the kernel reports it (None fields if unknown)."""
    info = {"bdf": None, "numa_node": None, "local_cpus": None}
    try:
        import pynvml
        pynvml.nvmlInit()
        visible = os.environ.get("CUDA_VISIBLE_DEVICES")
        phys = int(visible.split(",")[index]) if visible and visible.split(",")[index].strip().isdigit() else index
        h = pynvml.nvmlDeviceGetHandleByIndex(phys)
        bdf = pynvml.nvmlDeviceGetPciInfo(h).busId
        bdf = bdf.decode() if isinstance(bdf, bytes) else bdf
        bdf = bdf.lower()
        if len(bdf.split(":")[0]) == 8:              # nvml prints an 8-digit domain, sysfs a 4-digit one
            bdf = bdf[4:]
        info["bdf"] = bdf
        base = "/sys/bus/pci/devices/" + bdf
        with open(base + "/numa_node") as fh:
            info["numa_node"] = int(fh.read().strip())
        with open(base + "/local_cpulist") as fh:
            info["local_cpus"] = _parse_cpulist(fh.read())
    except Exception:
        pass
    return info


def bind_to_gpu(local_rank, local_world=1):
    """Restrict this process to the CPUs next to GPU `local_rank`.  Returns a dict describing what was done."""
    allowed = sorted(os.sched_getaffinity(0))
    loc = gpu_locality(local_rank)
    done = dict(loc, allowed=len(allowed), mode="none")
    cpus = [c for c in (loc["local_cpus"] or []) if c in allowed]
    if loc["numa_node"] is not None and loc["numa_node"] >= 0 and cpus and len(cpus) < len(allowed):
        done["mode"] = "numa-local"
    elif local_world > 1:
        per = max(1, len(allowed) // local_world)
        cpus = allowed[local_rank * per:(local_rank + 1) * per] or allowed
        done["mode"] = "even-split"
    else:
        cpus = allowed
    if done["mode"] != "none":
        try:
            os.sched_setaffinity(0, cpus)
        except OSError:
            done["mode"] = "none"
    done["cpus"] = len(cpus) if done["mode"] != "none" else len(allowed)
    done.pop("local_cpus", None)
    return done

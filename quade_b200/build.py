"""Builds quade_b200/libquade_b200.so in-tree with nvcc for sm_100a (no JIT cache, no torch extension)."""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.environ.get("QUADE_LIB_PATH") or os.path.join(HERE, "libquade_b200.so")   # override: A/B builds of kernel variants
SOURCES = ["quade_kernels.cu", "quade_api.cu", "quade_static.cu"]
DEPS = SOURCES + sorted(f for f in os.listdir(CSRC) if f.endswith((".h", ".cuh", ".inc"))) + [os.path.join("..", "..", "include", "quade_b200.h")]
#: the synthetic-data generator is test / benchmark infrastructure: its own library, never linked into the product
SYNTH_LIB = os.path.join(HERE, "libquade_synth.so")
SYNTH_SOURCES = ["quade_synth.cu"]
SYNTH_DEPS = SYNTH_SOURCES + [os.path.join("..", "..", "include", "quade_synth.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC,-O3,-Wall", "-cudart", "static"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libquade_b200.so cannot be built")


def _stale(lib, deps):
    if not os.path.exists(lib):
        return True
    t = os.path.getmtime(lib)
    return any(os.path.getmtime(os.path.join(CSRC, d)) > t for d in deps)


def is_stale():
    return _stale(LIB, DEPS)


def _compile_and_link(lib, sources, verbose):
    """Objects and the library are written under per-process names and published with os.replace: the ranks of a
    torchrun job may all find the library stale and build at the same time."""
    nvcc = _nvcc()
    objs = []
    procs = []
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    tag = "%s.%d" % (os.path.basename(lib), os.getpid())
    for src in sources:
        obj = os.path.join(HERE, "build", "%s.%s.o" % (os.path.splitext(src)[0], tag))
        objs.append(obj)
        cmd = [nvcc] + NVCC_FLAGS + os.environ.get("QUADE_BUILD_DEFS", "").split() + (["-Xptxas", "-v"] if verbose else []) + ["-c", os.path.join(CSRC, src), "-o", obj]
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    try:
        for cmd, p in procs:
            out, _ = p.communicate()
            if verbose and out:
                print(out)
            if p.returncode != 0:
                raise RuntimeError("nvcc failed: %s\n%s" % (" ".join(cmd), out))
        tmp = "%s.%d.tmp" % (lib, os.getpid())
        subprocess.check_call([nvcc, "-shared", "-cudart", "static", "-Wno-deprecated-gpu-targets", "-o", tmp] + objs + ["-lpthread", "-ldl"])
        os.replace(tmp, lib)
    finally:
        for o in objs:
            if os.path.exists(o):
                os.remove(o)
    return lib


def build(force=False, verbose=False):
    """Compile the CUDA kernels + C-ABI into quade_b200/libquade_b200.so. Returns the path."""
    if not force and not is_stale():
        return LIB
    return _compile_and_link(LIB, SOURCES, verbose)


IO_LIB = os.path.join(HERE, "libquade_io.so")
IO_DEPS = ["quade_io.cpp", os.path.join("..", "..", "include", "quade_io.h")]


def build_io(force=False, verbose=False):
    """Compile the host output sink (gzip-member writer pool; g++, zlib) into quade_b200/libquade_io.so."""
    if not force and not _stale(IO_LIB, IO_DEPS):
        return IO_LIB
    cxx = os.environ.get("CXX") or shutil.which("g++") or "g++"
    tmp = "%s.%d.tmp" % (IO_LIB, os.getpid())
    cmd = [cxx, "-O2", "-g", "-std=c++17", "-fPIC", "-Wall", "-Wextra", "-shared", "-o", tmp, os.path.join(CSRC, "quade_io.cpp"),
           "-lz", "-lpthread"]
    out = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose and out.stdout:
        print(out.stdout)
    if out.returncode != 0:
        raise RuntimeError("g++ failed: %s\n%s" % (" ".join(cmd), out.stdout))
    os.replace(tmp, IO_LIB)
    return IO_LIB


def build_synth(force=False, verbose=False):
    """Compile the synthetic-data generator (host + device) into quade_b200/libquade_synth.so."""
    if not force and not _stale(SYNTH_LIB, SYNTH_DEPS):
        return SYNTH_LIB
    return _compile_and_link(SYNTH_LIB, SYNTH_SOURCES, verbose)


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_synth(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_io(force="--force" in sys.argv, verbose="-v" in sys.argv))

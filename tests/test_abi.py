"""The C-ABI boundary without a GPU: libquade_b200.so loads, exports exactly the entry points that
include/quade_b200.h declares, and the ctypes mirrors in quade_b200/_lib.py have the header's struct sizes
(checked by compiling the header with gcc).  No compute call is made here; without a device the compute entry
points must fail loudly (QUADE_ERR_CUDA), never fall back."""
import ctypes as C
import os
import re
import subprocess

import pytest

from quade_b200 import _lib, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "quade_b200.h")


def _declared():
    text = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    return re.findall(r"^(?:const\s+)?\w+\s+\*?(quade_\w+)\s*\(", text, flags=re.M)


def test_header_declarations_are_the_bound_symbols():
    names = _declared()
    assert len(names) == len(set(names)) >= 13
    assert sorted(names) == sorted(_lib.SYMBOLS)


def test_library_exports_every_declared_symbol():
    L = C.CDLL(_lib.lib_path()) if os.path.exists(_lib.lib_path()) else _lib.lib()
    for name in _declared():
        assert getattr(L, name) is not None, name
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.lib_path()], capture_output=True, text=True, check=True).stdout
    exported = {ln.split()[-1] for ln in out.splitlines() if " T " in ln}
    assert set(_declared()) <= exported
    # nothing but the declared C entry points carries the quade_ prefix: the boundary is the header
    assert {s for s in exported if s.startswith("quade_")} == set(_declared())


def test_ctypes_structs_match_the_header(tmp_path):
    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "quade_b200.h"\n#include "quade_synth.h"\n'
                   'int main(void){printf("%zu %zu %zu %zu %zu %zu %d %d\\n", sizeof(quade_params), sizeof(quade_batch),'
                   ' sizeof(quade_result), sizeof(quade_synth_params), offsetof(quade_result, ms_kernel),'
                   ' offsetof(quade_result, gpu_launches), QUADE_K_COUNT, QUADE_ABI_VERSION);return 0;}\n')
    exe = tmp_path / "sizes"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)],
                   check=True)
    got = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    assert got[:4] == [C.sizeof(_lib.Params), C.sizeof(_lib.Batch), C.sizeof(_lib.Result), C.sizeof(synth.SynthParams)]
    assert got[4] == _lib.Result.ms_kernel.offset and got[5] == _lib.Result.gpu_launches.offset
    assert got[6] == _lib.K_COUNT
    assert got[7] == _lib.lib().quade_version()


def test_no_cpu_fallback_without_a_device():
    if _lib.device_count() > 0:
        pytest.skip("a CUDA device is present: the loud-failure leg is for the CPU container")
    p = _lib.Params()
    p.dual, p.n_samples = 1, 1
    ctx = C.c_void_p()
    rc = _lib.lib().quade_ctx_create(0, C.byref(p), C.byref(ctx))
    assert rc == _lib.ERR_CUDA and not ctx.value
    assert _lib.lib().quade_last_error()
    with pytest.raises(_lib.QuadeError):
        _lib.check(rc)


def test_the_generator_is_not_in_the_product_library():
    """quade_synth_* (test / benchmark infrastructure) lives in libquade_synth.so: the product library exports none of
    it, and the generator library exports nothing of the product."""
    L = synth.lib()
    for name in synth.SYMBOLS:
        assert getattr(L, name) is not None
    from quade_b200 import build
    out = subprocess.run(["nm", "-D", "--defined-only", build.SYNTH_LIB], capture_output=True, text=True, check=True).stdout
    exported = {ln.split()[-1] for ln in out.splitlines() if " T " in ln and ln.split()[-1].startswith("quade_")}
    assert exported == set(synth.SYMBOLS)
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.lib_path()], capture_output=True, text=True, check=True).stdout
    assert "quade_synth" not in out

"""Load-time shim that runs the UNMODIFIED reference sources under Python 3.

TEST INFRASTRUCTURE ONLY.  Nothing in the product path (`quade_b200/`) may import this module.
It is used (a) by `tests/golden/make_golden.py` to generate golden vectors from the reference's own
code, and (b) by tests that cross-check the C oracle when `/root/reference` is present (it never is
on the GPU box, so those tests skip there).

No reference code is committed to this repository: the shim reads `/root/reference/src/*.py` (or the git-ignored
copy under `oracle/_ref/`, made by `oracle/make_ref.py` so that the sources reach the GPU box) as text at run time, applies the textual substitutions listed in SURVEY.md section 8(c) (Python 2 -> 3 spelling
only, no behavioural change inside the ASCII FASTQ domain) and `exec`s the result into synthetic
modules.  Substitutions:

  pyFastq/FastqReader.py:57-58  `raise StopIteration(...)`  -> `return`        (PEP 479)
  pyFastq/FastqReader.py:30     binary open -> latin-1 text mode, newline="\\n" (Py2 str == bytes)
  FastqWriter.py:86,89          `.write(buffer)` -> `.write(buffer.encode("latin-1"))`
  Sample.py:113-126             `*100/` -> `*100//` (6 sites)                  (Py2 int division)
  Conf_file.py:17               'wb' -> 'w'
  Quade.py:211-214,241-243      `_gen.next()` -> `_gen.__next__()`
  Quade.py:187                  report "wb" -> "w"
  Quade.py:92                   RawConfigParser(...) + strict=False, interpolation=None, inline ';'
  Quade.py:290                  `if __name__ == '__main__':` -> `if False:`
Environment: builtins.unichr = chr; sys.modules['ConfigParser'] = configparser; FastqSeq and
FastqWriter pre-registered as top-level modules; stub package `pyFastq`.
"""
import builtins
import configparser
import contextlib
import io
import os
import sys
import types

def _root():
    """QUADE_REFERENCE_ROOT, else /root/reference (build container), else the copy oracle/make_ref.py left under
    oracle/_ref (the GPU box)."""
    env = os.environ.get("QUADE_REFERENCE_ROOT")
    if env:
        return env
    if os.path.isfile("/root/reference/src/Quade.py"):
        return "/root/reference"
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


REFERENCE_ROOT = _root()
_SRC = os.path.join(REFERENCE_ROOT, "src")


def available():
    return os.path.isfile(os.path.join(_SRC, "Quade.py"))


def _must_replace(text, old, new, count=None):
    n = text.count(old)
    if n == 0 or (count is not None and n != count):
        raise RuntimeError("shim substitution %r matched %d times (expected %s)" % (old, n, count))
    return text.replace(old, new)


def _load(name, relpath, subs, modules, extra_globals=None):
    with open(os.path.join(_SRC, relpath), "r", encoding="utf-8") as fh:
        text = fh.read()
    for old, new, count in subs:
        text = _must_replace(text, old, new, count)
    mod = types.ModuleType(name)
    mod.__file__ = os.path.join(_SRC, relpath)
    if extra_globals:
        mod.__dict__.update(extra_globals)
    modules[name] = mod
    saved = {k: sys.modules.get(k) for k in modules}
    sys.modules.update(modules)
    try:
        exec(compile(text, mod.__file__, "exec"), mod.__dict__)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return mod


def load_reference():
    """Return a dict of freshly exec'd reference modules (fresh class-level Sample state)."""
    if not available():
        raise RuntimeError("reference sources not found under %s" % _SRC)
    builtins.unichr = chr  # FastqSeq.py:73
    mods = {"ConfigParser": configparser}
    pkg = types.ModuleType("pyFastq")
    pkg.__path__ = []
    mods["pyFastq"] = pkg
    _load("FastqSeq", "pyFastq/FastqSeq.py", [], mods)
    fr = _load("pyFastq.FastqReader", "pyFastq/FastqReader.py", [
        ('raise StopIteration("\\t{} sequences parsed".format(i))', "return", 1),
        ('gopen(fastq_file, "rb") if fastq_file[-2:] == "gz" else open(fastq_file, "rb")',
         'gopen(fastq_file, "rt", encoding="latin-1", newline="\\n") if fastq_file[-2:] == "gz" '
         'else open(fastq_file, "r", encoding="latin-1", newline="\\n")', 1),
    ], mods)
    pkg.FastqReader = fr
    _load("FastqWriter", "FastqWriter.py", [
        ("fastq_file.write(self.R1_buffer)", 'fastq_file.write(self.R1_buffer.encode("latin-1"))', 1),
        ("fastq_file.write(self.R2_buffer)", 'fastq_file.write(self.R2_buffer.encode("latin-1"))', 1),
    ], mods)
    _load("Sample", "Sample.py", [("*100/", "*100//", 6)], mods)
    _load("Conf_file", "Conf_file.py", [("'wb'", "'w'", 1)], mods)
    _load("Quade", "Quade.py", [
        ("_gen.next()", "_gen.__next__()", 7),
        ('open ("Quade_report.csv", "wb")', 'open ("Quade_report.csv", "w")', 1),
        ("ConfigParser.RawConfigParser(allow_no_value=True)",
         'ConfigParser.RawConfigParser(allow_no_value=True, strict=False, interpolation=None, '
         'inline_comment_prefixes=(";",))', 1),
        ("if __name__ == '__main__':", "if False:", 1),
    ], mods)
    return mods


def run_reference(conf_path, workdir, quiet=True):
    """Run `Quade(conf_path)()` from the reference sources with cwd=workdir.

    Returns (elapsed_seconds_reported_by_reference_or_None, stdout_text).
    """
    mods = load_reference()
    Quade = mods["Quade"].Quade
    old = os.getcwd()
    buf = io.StringIO()
    os.chdir(workdir)
    try:
        with contextlib.redirect_stdout(buf) if quiet else contextlib.nullcontext():
            q = Quade(conf_file=conf_path)
            q()
    finally:
        os.chdir(old)
    out = buf.getvalue()
    secs = None
    for line in out.splitlines():
        if line.startswith("Done in ") and line.endswith("s"):
            secs = float(line[len("Done in "):-1])
    return secs, out


if __name__ == "__main__":
    s, o = run_reference(os.path.abspath(sys.argv[1]), os.getcwd(), quiet=False)

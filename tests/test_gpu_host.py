"""GPU twin of tests/test_host.py: the reference's documented smoke test (README.md:67-72) through
Quade.__call__ with the real CUDA engine, compared with the reference's golden outputs."""
import os
import subprocess
import sys

import pytest

from tests.test_host import ROOT, _check_golden_outputs, _fresh, _stage_dataset

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("batch_bytes", [8 << 20, 4096])
def test_golden_run_on_the_gpu(tmp_path, monkeypatch, batch_bytes):
    from quade_b200.quade import Quade
    _fresh()
    wd = _stage_dataset(tmp_path)
    monkeypatch.chdir(wd)
    monkeypatch.setattr(Quade, "ENGINE_FACTORY", None)
    monkeypatch.setattr(Quade, "BATCH_BYTES", batch_bytes)
    q = Quade(conf_file="Quade_conf_file.txt")
    assert q() == 0
    _check_golden_outputs(wd)


def test_command_line_entry(tmp_path):
    wd = _stage_dataset(tmp_path)
    env = dict(os.environ, PYTHONPATH=ROOT, QUADE_BATCH_BYTES=str(1 << 20))
    out = subprocess.run([sys.executable, os.path.join(ROOT, "Quade.py"), "-c", "Quade_conf_file.txt"], cwd=str(wd),
                         env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "Start parsing files: 3 chunks to be parsed" in out.stdout and "Done in" in out.stdout
    assert out.stdout.count("Skipping the sequence") == 1          # C1_R1 record 0
    _check_golden_outputs(wd)


def test_sharded_run_two_gpus_nccl(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (the gloo twin runs in tests/test_host.py)")
    wd = _stage_dataset(tmp_path)
    env = dict(os.environ, PYTHONPATH=ROOT, QUADE_BATCH_BYTES=str(1 << 20))
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29642", "-m", "quade_b200", "-c",
                          "Quade_conf_file.txt"], cwd=str(wd), env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout + out.stderr
    _check_golden_outputs(wd)

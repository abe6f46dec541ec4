"""The plain-C multi-GPU driver (examples/multi_gpu_gather.c): one process per GPU, CUDA IPC handles exchanged over
pipes (no torch, no NCCL), the label / score gather fused into the kernel; every rank verifies every shard of its
gathered arrays against the owner's plain run.  Runs on min(2, visible GPUs) -- with one GPU the same protocol with a
single rank; the two-rank form is exercised with `gpurun --gpus 2`."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "examples")])
    return os.path.join(ROOT, "examples", "multi_gpu_gather")


def test_example_builds_and_links_no_torch():
    exe = _build()
    out = subprocess.run(["ldd", exe], capture_output=True, text=True).stdout
    assert "libscrubjay_cuda.so" in out and "torch" not in out and "nccl" not in out


@pytest.mark.gpu
def test_example_runs_and_verifies_every_shard():
    import torch
    exe = _build()
    n = min(2, torch.cuda.device_count())
    r = subprocess.run([exe, str(n), "3000"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.count("verified") == n and "OK: %d GPU(s)" % n in r.stdout, r.stdout + r.stderr

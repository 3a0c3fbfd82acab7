"""Collective correctness on real GPUs (SURVEY.md §4 tier 5): fused all-reduce / barrier /
reduce+optimizer vs torch.distributed NCCL, one process per GPU, launched with torchrun."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _run(nproc, symm="auto"):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", "29631", os.path.join(HERE, "mp_fused_check.py")]
    env = dict(os.environ, TONY_B200_COMM_TIMEOUT_MS="20000", TONY_B200_SYMM=symm)
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert res.returncode == 0 and "MP_OK" in res.stdout, res.stdout[-3000:] + res.stderr[-3000:]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("symm", ["ipc", "vmm"])
def test_fused_collectives_2gpu(symm):
    """ipc: cudaMalloc + CUDA IPC, P2P ld/st kernels.  vmm: cuMemCreate + fd exchange + NVLS multicast,
    in-switch reduction (multimem.ld_reduce / multimem.st) for the fused optimizer and all-reduce."""
    _run(2, symm)


@pytest.mark.skipif(torch.cuda.device_count() < 8, reason="needs 8 GPUs")
def test_fused_collectives_8gpu():
    _run(8, "auto")

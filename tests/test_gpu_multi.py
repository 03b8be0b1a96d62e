"""Multi-GPU parity (needs >= 2 GPUs on the box, otherwise skipped): launches tests/mp_check.py under torchrun, one rank
per GPU over NCCL, and requires own- and halo-node tracers to be bit-identical to the oracle on the undivided mesh and
the distributed SSH solve to agree to solver tolerance."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def ngpus():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.parametrize("mesh,nranks", [("pi", 2), ("band", 2), ("farclike", 2), ("soufflet", 4), ("farclike", 4), ("band", 8)])
def test_partitioned_gpu_run_matches_undivided_oracle(mesh, nranks):
    if ngpus() < nranks:
        pytest.skip("needs %d GPUs" % nranks)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nranks), "--master-addr", "127.0.0.1",
           "--master-port", str(29500 + nranks + (10 if mesh == "farclike" else 0)), os.path.join(ROOT, "tests", "mp_check.py"), mesh]
    out = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=180, cwd=ROOT).stdout
    assert "MPCHECK_RESULT PASS" in out, out[-3000:]

"""CPU execution of the MULTI-RANK device path through the CUDA-on-CPU emulation (tests/emu, test infrastructure): one
process per rank under torchrun, the kernels' source (halo pack/unpack kernels included) compiled with g++, NCCL send/recv
carried over files.  tests/mp_check.py --emulate compares every rank's own and halo tracers with the oracle on the undivided
mesh bit for bit, --ale adds the device-side ALE producers (--wsplit: with the explicit/implicit split of Wvel) and the SSH
assembly with their halo exchanges against the undivided host mirror, and the distributed SSH solves against the undivided one."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("mesh,nranks,extra", [("pi", 2, ["--wsplit"]), ("band", 3, [])])
def test_emulated_partitioned_run_matches_undivided(mesh, nranks, extra):
    from tests import emu
    emu.build()                                           # once, before the ranks start: they must not compile concurrently
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nranks), "--master-addr", "127.0.0.1",
           "--master-port", str(29700 + nranks + os.getpid() % 100), os.path.join(ROOT, "tests", "mp_check.py"), mesh, "--emulate", "--ale"] + extra
    out = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900, cwd=ROOT).stdout
    assert "MPCHECK_RESULT PASS" in out, out[-3000:]

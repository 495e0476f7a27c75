"""Multi-rank tests.

* `-m gpu`: launches tests/dist_worker.py with one process per GPU (2, and 4 when the box has
  them) under torch.distributed.run; skipped on a single-GPU box.
* `-m "not gpu"`: world_size-2 (and 4) gloo runs of tests/dist_host_worker.py on the CPU: the
  host-side logic of the distributed path -- per-rank decomposition, map_comms tables, the
  peer-grouped message plan the NCCL exchange uses (dlesm_comm_plan) -- driving a real
  multi-process exchange over gloo with numpy pack/unpack, checked against the oracle's
  exchange_generic.
"""
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _torchrun(script, nproc, timeout, extra_env=None):
    env = dict(os.environ)
    env.update(extra_env or {})
    env.pop("DL_ESM_ALIGNMENT", None)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc),
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), os.path.join(ROOT, "tests", script)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env, cwd=ROOT)
    return res


@pytest.mark.parametrize("nproc", [2, 4])
def test_host_side_distributed_logic_over_gloo(nproc):
    res = _torchrun("dist_host_worker.py", nproc, 300)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "HOST DIST OK" in res.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("nproc", [2, 4])
def test_multi_gpu_parity(nproc):
    import torch
    if torch.cuda.device_count() < nproc:
        pytest.skip("needs %d GPUs, box has %d" % (nproc, torch.cuda.device_count()))
    res = _torchrun("dist_worker.py", nproc, 900)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "DIST OK" in res.stdout

"""tests/c_driver.c: the reference's example program (example/model.f90:54-109) and its halo test
(tests/dist_mem/test_halos.f90:60-265) as a plain-C program over the C ABI -- the call sequence the B200
build of field_mod / grid_mod (fortran/patches/) makes, with no Python or ctypes in between."""
import os
import subprocess
import sys
import tempfile

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402


def run(args, nranks=1, timeout=300):
    exe = entry.build_c_driver()
    if nranks == 1:
        env = dict(os.environ, RANK="0", WORLD_SIZE="1", LOCAL_RANK="0")
        env.pop("DL_ESM_ALIGNMENT", None)
        p = subprocess.run([exe] + args, env=env, capture_output=True, text=True, timeout=timeout)
        return [(p.returncode, p.stdout, p.stderr)]
    with tempfile.TemporaryDirectory() as d:
        procs = []
        for r in range(nranks):
            env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(nranks), LOCAL_RANK=str(r),
                       DLESM_ID_FILE=os.path.join(d, "nccl_id"))
            env.pop("DL_ESM_ALIGNMENT", None)
            procs.append(subprocess.Popen([exe] + args, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                                          text=True))
        out = []
        for p in procs:
            try:
                so, se = p.communicate(timeout=timeout)
            except subprocess.TimeoutExpired:
                for q in procs:
                    q.kill()
                raise
            out.append((p.returncode, so, se))
        return out


def test_c_driver_builds_and_fails_loudly_without_a_gpu():
    """gcc + the header + the shared library are enough to build an application; without a CUDA device
    the first device call ends the program with the library's message (no CPU fallback)"""
    import torch
    exe = entry.build_c_driver()
    assert os.access(exe, os.X_OK)
    if torch.cuda.is_available():
        pytest.skip("a GPU is present: the run itself is the gpu-marked test")
    (rc, so, se), = run(["model"])
    assert rc == 2 and "no CPU fallback" in se, (rc, se)


@pytest.mark.gpu
def test_c_driver_example_model_known_answers():
    (rc, so, se), = run(["model"])
    assert rc == 0, (so, se)
    # 1 rank: four checksums of 40.0, for host-resident and for device-resident fields
    assert so.count("checksum =  4.00000000E+01") == 8, so
    assert "all checks passed on 1 rank(s)" in so


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(10, 4), (4, 10), (37, 23)])
def test_c_driver_halos_one_rank(shape):
    (rc, so, se), = run(["halos", str(shape[0]), str(shape[1])])
    assert rc == 0, (so, se)


@pytest.mark.gpu
@pytest.mark.parametrize("nranks", [2, 4])
def test_c_driver_several_ranks(nranks):
    """one process per GPU, NCCL id through a file: the 60.0 answer of example/model.f90 at two ranks and
    the hill-halo test with real neighbours"""
    import torch
    if torch.cuda.device_count() < nranks:
        pytest.skip("needs %d GPUs" % nranks)
    res = run(["model"], nranks)
    assert all(rc == 0 for rc, _, _ in res), res
    if nranks == 2:
        assert res[0][1].count("checksum =  6.00000000E+01") == 8, res[0][1]
    for shape in ((10, 10), (37, 23), (64, 48)):
        res = run(["halos", str(shape[0]), str(shape[1])], nranks)
        assert all(rc == 0 for rc, _, _ in res), res
        assert all("halo cells checked" in so for _, so, _ in res)

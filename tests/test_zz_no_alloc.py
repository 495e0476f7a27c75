"""dlesm_dev_copy_patch_static -- the patch copy for hosts that cannot allocate device memory.

Some applications that embed this library have no cudaMalloc to spare: their own pool (an OpenACC /
CUDA Fortran runtime, a caching allocator) already owns the whole GPU so a further allocation
fails, or their time-step loop must not take the device-wide synchronisation an allocation costs.
dlesm_dev_copy_patch_static serves them: an overlapping copy is staged through a fixed area that
is part of the library image (DLESM_STATIC_SCRATCH_DOUBLES doubles per GPU) instead of through
scratch memory the library allocates.

The check: through the C ABI, on the same field contents, the new entry point must leave exactly
the bits dlesm_dev_copy_patch leaves (a copy moves doubles, so the bar is bit equality, as in
tests/test_gpu_parity.py) -- on two calls in a row, the second smaller than the first and over
other cells, so that anything the second call took from what the first left in the shared area
would show.  Patches beyond the area's size must be refused with a message that names the limit.

This file sorts last on purpose (pytest -x stops at the first failure; a failure here must not hide
the parity tests), and it passes alone in a fresh process as well as after the other files.
"""
import ctypes as C

import numpy as np
import pytest

import dl_esm_inf_b200 as fd
from dl_esm_inf_b200 import _lib

BCS_PER = (fd.GO_BC_PERIODIC, fd.GO_BC_PERIODIC, fd.GO_BC_NONE)
LIMIT = 262144   # DLESM_STATIC_SCRATCH_DOUBLES of include/dlesm_b200.h


@pytest.fixture
def serial(monkeypatch):
    monkeypatch.delenv("DL_ESM_ALIGNMENT", raising=False)
    fd.gocean_initialise(1, 1, device=0)
    yield


def _field(n_x, n_y, seed):
    g = fd.grid_type(fd.GO_ARAKAWA_C, BCS_PER, fd.GO_OFFSET_SW)
    g.decompose(n_x, n_y)
    fd.grid_init(g, 1.0, 1.0, None)
    f = fd.r2d_field(g, fd.GO_T_POINTS)
    src = np.asfortranarray(np.random.default_rng(seed).uniform(-1.0, 1.0, size=(g.nx, g.ny)))
    return g, f, src


def _existing(f, src, s, d):
    """the library-allocating entry point on a fresh copy of `src`"""
    L = _lib.lib()
    f.set_data(src)
    assert L.dlesm_dev_copy_patch(C.c_void_p(f.device_ptr), *s, *d) == 0, L.dlesm_last_error()
    return f.get_data().copy(order="F")


def _static(f, g, s, d, blocking=1):
    L = _lib.lib()
    assert L.dlesm_dev_sync() == 0          # the library's own stream is not the one used below
    rc = L.dlesm_dev_copy_patch_static(C.c_void_p(f.device_ptr), g.nx, g.ny, *s, *d, None, blocking)
    assert rc == 0, L.dlesm_last_error()


def test_header_states_the_limit_used_here():
    import os
    import re
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "dlesm_b200.h")).read()
    assert int(re.search(r"#define\s+DLESM_STATIC_SCRATCH_DOUBLES\s+(\d+)", hdr).group(1)) == LIMIT


@pytest.mark.gpu
def test_static_scratch_patch_copy_equals_the_allocating_one_twice_in_a_row(serial):
    g, f, src = _field(150, 90, seed=7)
    try:
        # two overlapping copies (boxes are xstart, xstop, ystart, ystop): a large one, then a small one
        # elsewhere, shifted the other way; then a one-cell shift along x and a disjoint copy
        calls = [((2, 120, 3, 70), (9, 127, 11, 78)),
                 ((40, 60, 30, 50), (37, 57, 28, 48)),
                 ((1, g.nx - 1, 5, 5), (2, g.nx, 5, 5)),
                 ((1, 10, 1, 10), (60, 69, 70, 79))]
        for s, d in calls:
            assert (s[1] - s[0], s[3] - s[2]) == (d[1] - d[0], d[3] - d[2])
        # expected: the existing entry point applied call by call
        want, cur = [], src
        for s, d in calls:
            cur = _existing(f, cur, s, d)
            want.append(cur)
        assert not np.array_equal(want[0], src) and not np.array_equal(want[1], want[0])
        # the entry point under test, the calls back to back on one field, no re-upload in between
        f.set_data(src)
        for k, (s, d) in enumerate(calls):
            _static(f, g, s, d)
            got = f.get_data()
            assert np.array_equal(got, want[k]), "call %d differs in %d cells" % (k, int(np.sum(got != want[k])))
        # and once more without a host read-back between the two overlapping calls (non-blocking, stream order)
        f.set_data(src)
        _static(f, g, *calls[0], blocking=0)
        _static(f, g, *calls[1], blocking=1)
        assert np.array_equal(f.get_data(), want[1])
    finally:
        f.free()


@pytest.mark.gpu
def test_static_scratch_refuses_what_does_not_fit_and_bad_boxes(serial):
    L = _lib.lib()
    g, f, src = _field(600, 600, seed=11)
    try:
        f.set_data(src)
        L.dlesm_dev_sync()
        p = C.c_void_p(f.device_ptr)
        # 513 x 512 overlapping cells: one row more than the area holds
        assert 513 * 512 > LIMIT
        rc = L.dlesm_dev_copy_patch_static(p, g.nx, g.ny, 1, 513, 1, 512, 2, 514, 2, 513, None, 1)
        assert rc == -1                                      # DLESM_ERR_ARGS
        msg = L.dlesm_last_error().decode()
        assert str(LIMIT) in msg and "DLESM_STATIC_SCRATCH_DOUBLES" in msg
        assert np.array_equal(f.get_data(), src)             # a refused call changes nothing
        # exactly the limit fits (512 x 512), and agrees with the allocating entry point
        s, d = (1, 512, 1, 512), (2, 513, 2, 513)
        want = _existing(f, src, s, d)
        f.set_data(src)
        _static(f, g, s, d)
        assert np.array_equal(f.get_data(), want)
        # a disjoint copy larger than the area needs no staging and is accepted
        s, d = (1, 600, 1, 290), (1, 600, 301, 590)
        assert 600 * 290 > LIMIT // 2
        want = _existing(f, src, s, d)
        f.set_data(src)
        _static(f, g, s, d)
        assert np.array_equal(f.get_data(), want)
        # argument errors: shape mismatch, box outside the allocation, null pointer, host memory
        assert L.dlesm_dev_copy_patch_static(p, g.nx, g.ny, 1, 4, 1, 4, 2, 6, 2, 5, None, 1) == -1
        assert L.dlesm_dev_copy_patch_static(p, g.nx, g.ny, 1, 4, 1, 4, g.nx - 2, g.nx + 1, 1, 4, None, 1) == -1
        assert L.dlesm_dev_copy_patch_static(None, g.nx, g.ny, 1, 4, 1, 4, 2, 5, 2, 5, None, 1) == -1
        host = np.zeros((g.nx, g.ny), order="F")
        rc = L.dlesm_dev_copy_patch_static(host.ctypes.data_as(C.c_void_p), g.nx, g.ny, 1, 4, 1, 4, 2, 5, 2, 5, None, 1)
        assert rc < 0 and b"not device memory" in L.dlesm_last_error()
        # an empty box is a no-op, like the allocating entry point
        assert L.dlesm_dev_copy_patch_static(p, g.nx, g.ny, 5, 4, 1, 4, 9, 8, 1, 4, None, 1) == 0
    finally:
        f.free()

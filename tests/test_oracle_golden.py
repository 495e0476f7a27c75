"""Pins the CPU oracle against every golden vector the reference's own tests hold for the
hot path.  Each test restates one reference test program; citations are to
/root/reference/finite_difference (not read at run time)."""
import numpy as np
import pytest

from oracle import oracle as O

BCS_EXT = (O.GO_BC_EXTERNAL, O.GO_BC_EXTERNAL, O.GO_BC_NONE)
POINTS = (O.GO_U_POINTS, O.GO_V_POINTS, O.GO_T_POINTS, O.GO_F_POINTS)


def make_ranks(nranks, jpiglo, jpjglo, with_tmask=True, alignment=1):
    """gocean_initialise + grid_type + decompose + grid_init on every rank of an np-rank run."""
    grids = []
    for rank in range(1, nranks + 1):
        g = O.Grid(O.GO_ARAKAWA_C, BCS_EXT, O.GO_OFFSET_NE)
        g.decompose(jpiglo, jpjglo, rank=rank, nranks=nranks)
        tmask = None
        if with_tmask:
            # allocate(tmask(subdomain%global%nx, subdomain%global%ny)); tmask = 1
            tmask = np.ones((g.subdomain.glob.nx, g.subdomain.glob.ny), dtype=np.int32, order="F")
        g.init(1.0, 1.0, tmask=tmask, alignment=alignment, dist_mem_enabled=True)
        grids.append(g)
    return grids


def halo_exchange_all(fields):
    O.exchange_all([f.data for f in fields], [f.grid.comms for f in fields])


# --- example/model.f90:54-109 ---------------------------------------------------------
@pytest.mark.parametrize("nranks,expected", [(1, 40.0), (2, 60.0)])
def test_example_model(nranks, expected):
    grids = make_ranks(nranks, 4, 10)
    if nranks == 2:  # 4x10 on 2 ranks -> 1x2 tiles of 4x5
        assert (grids[0].decomp.nx, grids[0].decomp.ny) == (1, 2)
    for pts in POINTS:
        flds = [O.Field(g, pts) for g in grids]
        for r, f in enumerate(flds):
            f.data[:, :] = float(r + 1)          # init_field_by_rank
        halo_exchange_all(flds)
        total = sum(f.checksum() for f in flds)  # field_checksum + global_sum
        assert total == expected


def test_example_alloc_shape():
    """SURVEY 8 size table: 4x10 example allocates 7x13 (A=1) and 64x13 (A=64)."""
    g = make_ranks(1, 4, 10)[0]
    assert (g.nx, g.ny) == (7, 13)
    f = O.Field(g, O.GO_T_POINTS)
    assert f.internal.box() == (2, 5, 2, 11) and (f.whole.nx, f.whole.ny) == (6, 12)
    g64 = make_ranks(1, 4, 10, alignment=64)[0]
    assert (g64.nx, g64.ny) == (64, 13)


# --- tests/dist_mem/test_gsum.f90:107-154 ----------------------------------------------
@pytest.mark.parametrize("nranks", [1, 4, 6])
def test_gsum(nranks):
    jpiglo, jpjglo = 4, 10
    grids = make_ranks(nranks, jpiglo, jpjglo)
    for pts in POINTS:
        flds = [O.Field(g, pts) for g in grids]
        for f in flds:
            f.data[:, :] = -100.0
            i = f.internal
            f.data[i.xstart - 1:i.xstop, i.ystart - 1:i.ystop] = 1.0
        total = sum(f.checksum() for f in flds)
        assert abs(total - jpiglo * jpjglo) <= 1.0e-8


# --- tests/dist_mem/test_halos.f90:153-265 ---------------------------------------------
def hill(f, ji, jj):
    g = f.grid
    xpos = g.metrics["xt"][ji - 1, jj - 1]
    ypos = g.metrics["yt"][ji - 1, jj - 1]
    assert g.offset == O.GO_OFFSET_NE
    if f.b.defined_on == O.GO_U_POINTS:
        xpos = xpos + 0.5 * g.metrics["dx_u"][ji - 1, jj - 1]
    elif f.b.defined_on == O.GO_F_POINTS:
        xpos = xpos + 0.5 * g.metrics["dx_f"][ji - 1, jj - 1]
        ypos = ypos + 0.5 * g.metrics["dy_f"][ji - 1, jj - 1]
    elif f.b.defined_on == O.GO_V_POINTS:
        ypos = ypos + 0.5 * g.metrics["dy_v"][ji - 1, jj - 1]
    return float(np.float32(10000.0 * xpos + ypos))   # hill = real(...): default real kind


def init_field_hill(f):
    i = f.internal
    f.data[:, :] = 0.0
    for jj in range(i.ystart, i.ystop + 1):
        for ji in range(i.xstart, i.xstop + 1):
            f.data[ji - 1, jj - 1] = hill(f, ji, jj)
    for jj in range(1, f.grid.ny + 1):
        f.data[0:i.xstart - 1, jj - 1] = f.data[i.xstart - 1, jj - 1]
        f.data[i.xstop:, jj - 1] = f.data[i.xstop - 1, jj - 1]
    for ji in range(1, f.grid.nx + 1):
        f.data[ji - 1, 0:i.ystart - 1] = f.data[ji - 1, i.ystart - 1]
        f.data[ji - 1, i.ystop:] = f.data[ji - 1, i.ystop - 1]


def check_hill_halos(f):
    g, i, tol = f.grid, f.internal, 1.0e-8
    sub = g.subdomain.glob
    nchecked = 0
    if sub.xstart > 1:
        for jj in range(i.ystart, i.ystop + 1):
            assert abs(f.data[i.xstart - 2, jj - 1] - hill(f, i.xstart - 1, jj)) <= tol
            nchecked += 1
    if sub.xstop < g.global_nx:
        for jj in range(i.ystart, i.ystop + 1):
            assert abs(f.data[i.xstop, jj - 1] - hill(f, i.xstop + 1, jj)) <= tol
            nchecked += 1
    if sub.ystart > 1:
        for ji in range(i.xstart, i.xstop + 1):
            assert abs(f.data[ji - 1, i.ystart - 2] - hill(f, ji, i.ystart - 1)) <= tol
            nchecked += 1
    if sub.ystop < g.global_ny:
        for ji in range(i.xstart, i.xstop + 1):
            assert abs(f.data[ji - 1, i.ystop] - hill(f, ji, i.ystop + 1)) <= tol
            nchecked += 1
    return nchecked


# tests/dist_mem/Makefile:64-72
@pytest.mark.parametrize("nranks,jpiglo,jpjglo", [(2, 10, 4), (2, 4, 10), (4, 10, 10), (6, 10, 10)])
def test_halos(nranks, jpiglo, jpjglo):
    grids = make_ranks(nranks, jpiglo, jpjglo)
    nchecked = 0
    for pts in (O.GO_U_POINTS, O.GO_V_POINTS, O.GO_T_POINTS, O.GO_F_POINTS):
        flds = [O.Field(g, pts) for g in grids]
        for f in flds:
            init_field_hill(f)
        halo_exchange_all(flds)
        nchecked += sum(check_hill_halos(f) for f in flds)
    assert nchecked > 0


def test_halos_corners_too():
    """Corners are not checked by the reference test; map_comms' 1x1 diagonal messages
    (src/parallel_comms_mod.f90:988-1170) must fill them with the diagonal owner's value."""
    grids = make_ranks(4, 10, 10)
    flds = [O.Field(g, O.GO_T_POINTS) for g in grids]
    for f in flds:
        init_field_hill(f)
    halo_exchange_all(flds)
    f = flds[0]  # SW tile: its NE corner halo comes from rank 4
    i = f.internal
    assert f.data[i.xstop, i.ystop] == hill(f, i.xstop + 1, i.ystop + 1)
    c = grids[0].comms
    assert c.nsend == 3 and sorted(s["dir"] for s in c.sends()) == [O.Iminus, O.Jminus, O.IminusJminus]


# --- tests/dist_mem/test_reduction.f90:128-192 -------------------------------------------
@pytest.mark.parametrize("nranks", [1, 4, 6])
def test_reduction(nranks):
    jpiglo = jpjglo = 10
    glob = np.empty((jpiglo, jpjglo), order="F")
    for j in range(1, jpjglo + 1):
        for i in range(1, jpiglo + 1):
            glob[i - 1, j - 1] = (i - 1) + (j - 1) * jpiglo      # unique_global_value
    grids = make_ranks(nranks, jpiglo, jpjglo, with_tmask=False)   # grid_init(model_grid, dx, dy)
    flds = [O.Field(g, O.GO_T_POINTS, init_global_data=glob) for g in grids]
    for f in flds:                                                 # check_field_distribution
        i, sub = f.internal, f.grid.subdomain.glob
        for jj in range(i.ystart, i.ystop + 1):
            for ji in range(i.xstart, i.xstop + 1):
                want = (sub.xstart + ji - i.xstart - 1) + (sub.ystart + jj - i.ystart - 1) * jpiglo
                assert f.data[ji - 1, jj - 1] == want
        f.data[i.xstart - 1:i.xstop, i.ystart - 1:i.ystop] += 1    # update_field
    out = O.gather_inner_data(grids[0].decomp, flds)
    assert np.array_equal(out, glob + 1)                           # check_gathered_field


# --- tests/device_computation/test_device_io.f90:157-180 ---------------------------------
@pytest.mark.parametrize("alignment", [1, 2, 8, 64])
def test_device_io(alignment):
    g = make_ranks(1, 5, 5, alignment=alignment)[0]
    if alignment == 1:
        assert (g.nx, g.ny) == (8, 8)
    f = O.Field(g, O.GO_U_POINTS)
    device = np.full(g.nx * g.ny, np.nan)                 # init_device_memory
    f.data[:, :] = 0
    O.write_to_device(f.data, device)                     # all device data is 0
    f.data[:, :] = 1
    O.write_to_device(f.data, device, 2, 2, 5, 5)         # 5x5 block starting at 2,2 is 1
    device *= 2                                           # simulate_device_computation
    O.read_from_device(device, f.data, 5, 5, 4, 4)        # read the bottom-right quadrant
    assert f.data[0, 0] == 1.0 and f.data[4, 4] == 2.0 and f.data[6, 6] == 0.0
    if alignment == 1:
        want = np.ones((8, 8))
        want[4:8, 4:8] = 0.0
        want[4:6, 4:6] = 2.0
        assert np.array_equal(f.data, want)


# --- the fixture file itself -----------------------------------------------------------
def test_golden_fixture_matches_what_these_tests_assert():
    """tests/golden/reference_known_answers.json is the committed record of the reference's
    known answers; the numbers asserted above must be the ones in the file."""
    import json
    import os
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_known_answers.json")
    gold = json.load(open(path))
    for nranks, want in gold["example_model"]["checksum_by_nranks"].items():
        test_example_model(int(nranks), want)
    g = make_ranks(1, *gold["example_model"]["domain"])[0]
    assert [g.nx, g.ny] == gold["example_model"]["alloc_A1"]
    assert gold["test_gsum"]["checksum"] == gold["test_gsum"]["domain"][0] * gold["test_gsum"]["domain"][1]
    for nranks, jpiglo, jpjglo in gold["test_halos"]["cases"]:
        test_halos(nranks, jpiglo, jpjglo)
    for a in gold["test_device_io"]["alignments"]:
        test_device_io(a)

"""The PSy layer (metadata-driven kernel dispatch): kernel_mod / argument_mod
metadata + a per-point code(ji, jj, ...) -> a generated sm_100a sweep.

CPU part: NVRTC needs no GPU, so kernel objects are built (and their compile errors reported)
here.  GPU part: the NEMOLite2D-style kernels of examples/nemolite2d.py through the C ABI
against oracle/nemolite2d_oracle.py on the same inputs: bit-exact where no sin() is involved,
1e-12 relative otherwise (tolerance of BASELINE.json's north_star)."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import dl_esm_inf_b200 as fd  # noqa: E402
from dl_esm_inf_b200 import _lib  # noqa: E402
from dl_esm_inf_b200.api import (GO_ALL_PTS, GO_CT, GO_CU, GO_EVERY, GO_GRID_AREA_T, GO_GRID_DX_CONST,  # noqa: E402
                                 GO_GRID_X_MAX_INDEX, GO_GRID_X_MIN_INDEX, GO_I_SCALAR, GO_INTERNAL_PTS,
                                 GO_R_SCALAR, GO_READ, GO_SUM, GO_TIME_STEP, GO_WRITE, field_arg, grid_arg,
                                 kernel_type, scalar_arg)
from examples import nemolite2d as N  # noqa: E402
from oracle import nemolite2d_oracle as NO  # noqa: E402

BCS_EXT = (fd.GO_BC_EXTERNAL, fd.GO_BC_EXTERNAL, fd.GO_BC_NONE)
RTOL = 1.0e-12


# ---------------------------------------------------------------------------------------------
# CPU: building kernels
# ---------------------------------------------------------------------------------------------
def test_reference_copy_kernel_builds_and_generates_the_loop_nest():
    """src/infrastructure_mod.f90:13-41: meta_args = (GO_WRITE, GO_EVERY), (GO_READ, GO_EVERY);
    ITERATES_OVER = GO_ALL_PTS; index_offset = GO_OFFSET_ANY"""
    k = kernel_type("field_copy", N.FIELD_COPY, [field_arg(GO_WRITE, GO_EVERY), field_arg(GO_READ, GO_EVERY)],
                    GO_ALL_PTS, fd.GO_OFFSET_ANY)
    src = k.generated_source
    assert "extern \"C\" __global__" in src and "dlesm_sweep_field_copy(" in src
    assert "dlesm_field a0, dlesm_cfield a1" in src           # write -> lvalue accessor, read -> const
    assert "field_copy_code(ji, jj, a0, a1)" in src
    k.free()


def test_every_nemolite2d_kernel_compiles_for_sm_100a():
    srcs = {"continuity": (N.CONTINUITY, N.CONTINUITY_META), "momentum_u": (N.MOMENTUM_U, N.MOMENTUM_U_META),
            "momentum_v": (N.MOMENTUM_V, N.MOMENTUM_V_META), "bc_ssh": (N.BC_SSH, N.BC_SSH_META)}
    for name, (code, meta) in srcs.items():
        k = kernel_type(name, code, meta, GO_INTERNAL_PTS, fd.GO_OFFSET_NE)
        assert ("%s_code(ji, jj" % name) in k.generated_source
        k.free()


def test_argument_types_follow_the_metadata():
    code = """__device__ void t_code(int ji, int jj, dlesm_field o, double s, int n, dlesm_cfield area, double dx,
                                     int xmin, int xmax) { o(ji, jj) = s * n + area(ji, jj) * dx + xmin + xmax; }"""
    k = kernel_type("t", code, [field_arg(GO_WRITE, GO_CT), scalar_arg(GO_R_SCALAR), scalar_arg(GO_I_SCALAR),
                                grid_arg(GO_GRID_AREA_T), grid_arg(GO_GRID_DX_CONST), grid_arg(GO_GRID_X_MIN_INDEX),
                                grid_arg(GO_GRID_X_MAX_INDEX)])
    assert "dlesm_field a0, double a1, int a2, dlesm_cfield a3, double a4, int a5, int a6" in k.generated_source
    k.free()


def test_compile_errors_are_reported_like_gocean_stop():
    with pytest.raises(fd.GoceanStop) as e:
        kernel_type("bad", "__device__ void bad_code(int ji, int jj, dlesm_field o) { o(ji, jj) = nope; }",
                    [field_arg(GO_WRITE, GO_CT)])
    assert "does not compile" in str(e.value) and "nope" in str(e.value)
    with pytest.raises(fd.GoceanStop):      # a read-only field is not assignable
        kernel_type("ro", "__device__ void ro_code(int ji, int jj, dlesm_cfield o) { o(ji, jj) = 1.0; }",
                    [field_arg(GO_READ, GO_CT)])


def test_metadata_is_validated():
    ok = "__device__ void m_code(int ji, int jj, dlesm_field o) { o(ji, jj) = 1.0; }"
    with pytest.raises(fd.GoceanStop):      # GO_SUM is only valid for globals (argument_mod.f90:44-45)
        kernel_type("m", ok, [field_arg(GO_SUM, GO_CT)])
    with pytest.raises(fd.GoceanStop):      # reductions are not supported
        kernel_type("m", ok, [field_arg(GO_WRITE, GO_CT), scalar_arg(GO_R_SCALAR, GO_SUM)])
    with pytest.raises(fd.GoceanStop):      # grid_type holds no time step
        kernel_type("m", ok, [field_arg(GO_WRITE, GO_CT), grid_arg(GO_TIME_STEP)])
    with pytest.raises(fd.GoceanStop):      # GO_EXTERNAL_PTS
        kernel_type("m", ok, [field_arg(GO_WRITE, GO_CT)], iterates_over=1)
    with pytest.raises(fd.GoceanStop):
        kernel_type("not an identifier", ok, [field_arg(GO_WRITE, GO_CT)])


# ---------------------------------------------------------------------------------------------
# GPU: the kernels against the numpy oracle
# ---------------------------------------------------------------------------------------------
def oracle_model(grid, arrays, model):
    """the oracle's Model over this grid's own metric arrays, mask and field bounds"""
    metrics = {k: grid.host_array(v) for k, v in NO.GRID_NAMES.items()}
    boxes = {}
    for pts, nm in ((fd.GO_T_POINTS, "t"), (fd.GO_U_POINTS, "u"), (fd.GO_V_POINTS, "v")):
        fb = grid.field_bounds(pts)
        boxes[nm] = NO.Box(*fb.internal.box())
        boxes[nm + "_all"] = NO.Box(*fb.whole.box())
    return NO.Model(arrays, metrics, grid.get_tmask(), boxes, model.rdt, model.cbfr, model.visc, N.OMEGA, N.D2R, N.G)


def make_model(n_x, n_y, alignment, monkeypatch, dx=1000.0, dy=1200.0):
    if alignment:
        monkeypatch.setenv("DL_ESM_ALIGNMENT", str(alignment))
    else:
        monkeypatch.delenv("DL_ESM_ALIGNMENT", raising=False)
    fd.gocean_initialise(1, 1, device=0)
    g = fd.grid_type(fd.GO_ARAKAWA_C, BCS_EXT, fd.GO_OFFSET_NE)
    g.decompose(n_x, n_y)
    tm, state = N.global_setup(n_x, n_y)
    fd.grid_init(g, dx, dy, N.local_tmask(g, tm))
    local = N.local_state(g, state)
    model = N.NemoLite2D(g)
    model.set_state(local)
    return g, model, oracle_model(g, local, model)


EXACT = ("sshn_u", "sshn_v", "hu", "hv", "ht", "ssha_u", "ssha_v")


def compare(model, om, step, exact_all=False):
    for nm in model.f:
        got, want = model.f[nm].get_data(), om.a[nm]
        if exact_all:
            assert np.array_equal(got, want), (nm, step, np.argwhere(got != want)[:3].tolist())
        else:
            # sin() in momentum_{u,v} and bc_ssh: device libm vs host libm (last-bit differences)
            err = np.abs(got - want)
            assert np.all(err <= RTOL * np.maximum(np.abs(want), 1e-3)), (nm, step, float(err.max()))


@pytest.mark.gpu
@pytest.mark.parametrize("n_x,n_y,alignment", [(20, 16, None), (67, 45, 64), (130, 77, 8), (256, 192, 64)])
def test_nemolite2d_steps_match_oracle(n_x, n_y, alignment, monkeypatch):
    g, model, om = make_model(n_x, n_y, alignment, monkeypatch)
    launches0 = fd.kernel_launch_count()
    for step in range(1, 6):
        model.step(step)
        om.step(step)
        compare(model, om, step)
    assert fd.kernel_launch_count() - launches0 >= 5 * 11      # eleven kernels per invoke
    # the fields evolved (the test is not comparing two copies of the initial state)
    assert not np.array_equal(om.a["un"], N.local_state(g, N.global_setup(n_x, n_y)[1])["un"])
    model.free()


@pytest.mark.gpu
def test_kernels_without_sin_are_bit_exact(monkeypatch):
    """continuity, bc_solid, copy (GO_ALL_PTS), next_ssh{u,v}: one kernel at a time"""
    g, model, om = make_model(67, 45, 64, monkeypatch)
    f, k, a, b, gr, tm = model.f, model.k, om.a, om.box, om.g, om.tmask
    k["continuity"]([f[n] for n in ("ssha", "sshn", "sshn_u", "sshn_v", "hu", "hv", "un", "vn")], (model.rdt,))
    NO.continuity(b["t"], a["ssha"], a["sshn"], a["sshn_u"], a["sshn_v"], a["hu"], a["hv"], a["un"], a["vn"],
                  model.rdt, gr["e12t"])
    k["bc_solid_u"]([f["ua"]])
    NO.bc_solid_u(b["u"], a["ua"], tm)
    k["bc_solid_v"]([f["va"]])
    NO.bc_solid_v(b["v"], a["va"], tm)
    k["copy"]([f["sshn"], f["ssha"]])
    NO.field_copy(b["t_all"], a["sshn"], a["ssha"])
    k["next_sshu"]([f["sshn_u"], f["sshn"]])
    NO.next_sshu(b["u"], a["sshn_u"], a["sshn"], tm, gr["e12t"], gr["e12u"])
    k["next_sshv"]([f["sshn_v"], f["sshn"]])
    NO.next_sshv(b["v"], a["sshn_v"], a["sshn"], tm, gr["e12t"], gr["e12v"])
    compare(model, om, 0, exact_all=True)
    assert (a["ua"] == 0.0).any() and not np.array_equal(a["ssha"], N.local_state(g, N.global_setup(67, 45)[1])["ssha"])
    model.free()


@pytest.mark.gpu
def test_device_layer_launch_and_argument_checks(monkeypatch):
    """dlesm_kernel_launch_dev (what a Fortran PSy layer calls with fld%device_ptr and
    grid%*_device) gives the same result as the host-mirror launch; wrong argument counts and a
    kernel written for another index_offset are refused"""
    g, model, om = make_model(40, 30, 64, monkeypatch)
    f, k = model.f, model.k
    names = ("ssha", "sshn", "sshn_u", "sshn_v", "hu", "hv", "un", "vn")
    k["continuity"]([f[n] for n in names], (model.rdt,))
    want = f["ssha"].get_data().copy()
    fd.set_field(f["ssha"], -1.0)
    i = f["ssha"].internal
    k["continuity"].launch_dev([f[n].device_ptr for n in names] + [g.device_array("area_t")], (model.rdt,), (),
                               (i.xstart, i.xstop, i.ystart, i.ystop))
    got = f["ssha"].get_data()
    sl = (slice(i.xstart - 1, i.xstop), slice(i.ystart - 1, i.ystop))
    assert np.array_equal(got[sl], want[sl])
    with pytest.raises(fd.GoceanStop):
        k["continuity"]([f[n] for n in names[:-1]], (model.rdt,))
    with pytest.raises(fd.GoceanStop):
        k["continuity"]([f[n] for n in names], ())
    sw_only = kernel_type("swk", "__device__ void swk_code(int ji, int jj, dlesm_field o) { o(ji, jj) = 1.0; }",
                          [field_arg(GO_WRITE, GO_CT)], GO_INTERNAL_PTS, fd.GO_OFFSET_SW)
    with pytest.raises(fd.GoceanStop):
        sw_only([f["ssha"]])
    model.free()

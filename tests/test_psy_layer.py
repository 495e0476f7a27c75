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

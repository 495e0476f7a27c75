"""A NEMOLite2D-style application over fd_lib fields: the second workload the
reference is written for ("e.g. NEMO code", src/grid_mod.f90:55-56) -- GO_OFFSET_NE, external
boundary conditions, a T-mask with land, curvilinear-style grid metrics (e1t = dx_t, e2u = dy_u,
e12t = area_t, gphiu, ...), which is also the only configuration the reference's distributed
memory path supports.

The kernels are NOT part of the reference; they follow the public NEMOLite2D
benchmark of PSycloneBench as this project restates it (oracle/nemolite2d_oracle.py is the
normative statement: parity unpinned).  What this file shows is the reference's *operator API*
at work on the GPU: each kernel is a per-point `code(ji, jj, ...)` plus go_arg metadata
(src/kernel_mod.f90, src/argument_mod.f90, cf. src/infrastructure_mod.f90:13-41), dispatched by
the library (dlesm_kernel_create / dlesm_kernel_launch) over device-resident r2d_fields.
"""
from __future__ import annotations

import numpy as np

import dl_esm_inf_b200 as fd
from dl_esm_inf_b200.api import (GO_CT, GO_CU, GO_CV, GO_EVERY, GO_GRID_AREA_T, GO_GRID_AREA_U, GO_GRID_AREA_V,
                                 GO_GRID_DX_T, GO_GRID_DX_U, GO_GRID_DX_V, GO_GRID_DY_T, GO_GRID_DY_U, GO_GRID_DY_V,
                                 GO_GRID_LAT_U, GO_GRID_LAT_V, GO_GRID_MASK_T, GO_INTERNAL_PTS, GO_ALL_PTS,
                                 GO_R_SCALAR, GO_READ, GO_READWRITE, GO_WRITE, field_arg, grid_arg, kernel_type,
                                 scalar_arg)

# physical parameters of the benchmark's namelist
G = 9.80665
OMEGA = 7.292116e-05
D2R = 3.14159265358979323846 / 180.0

CONTINUITY = r"""
__device__ void continuity_code(int ji, int jj, dlesm_field ssha, dlesm_cfield sshn, dlesm_cfield sshn_u,
                                dlesm_cfield sshn_v, dlesm_cfield hu, dlesm_cfield hv, dlesm_cfield un,
                                dlesm_cfield vn, double rdt, dlesm_cfield e12t)
{
    const double rtmp1 = (sshn_u(ji, jj) + hu(ji, jj)) * un(ji, jj);
    const double rtmp2 = (sshn_u(ji - 1, jj) + hu(ji - 1, jj)) * un(ji - 1, jj);
    const double rtmp3 = (sshn_v(ji, jj) + hv(ji, jj)) * vn(ji, jj);
    const double rtmp4 = (sshn_v(ji, jj - 1) + hv(ji, jj - 1)) * vn(ji, jj - 1);
    ssha(ji, jj) = sshn(ji, jj) + (rtmp2 - rtmp1 + rtmp4 - rtmp3) * rdt / e12t(ji, jj);
}
"""
CONTINUITY_META = [
    field_arg(GO_WRITE, GO_CT), field_arg(GO_READ, GO_CT), field_arg(GO_READ, GO_CU), field_arg(GO_READ, GO_CV),
    field_arg(GO_READ, GO_CU), field_arg(GO_READ, GO_CV), field_arg(GO_READ, GO_CU), field_arg(GO_READ, GO_CV),
    scalar_arg(GO_R_SCALAR), grid_arg(GO_GRID_AREA_T)]

MOMENTUM_U = r"""
__device__ void momentum_u_code(int ji, int jj, dlesm_field ua, dlesm_cfield un, dlesm_cfield vn, dlesm_cfield hu,
                                dlesm_cfield hv, dlesm_cfield ht, dlesm_cfield ssha_u, dlesm_cfield sshn,
                                dlesm_cfield sshn_u, dlesm_cfield sshn_v, double rdt, double cbfr, double visc,
                                double omega, double d2r, double g, dlesm_imask tmask, dlesm_cfield e1u,
                                dlesm_cfield e1v, dlesm_cfield e1t, dlesm_cfield e2u, dlesm_cfield e2t,
                                dlesm_cfield e12u, dlesm_cfield gphiu)
{
    if (tmask(ji, jj) + tmask(ji + 1, jj) <= 0) return;       // outside the computational domain
    if (tmask(ji, jj) <= 0 || tmask(ji + 1, jj) <= 0) return; // boundary u: left to the BC kernels
    const double u_e = 0.5 * (un(ji, jj) + un(ji + 1, jj)) * e2t(ji + 1, jj);
    const double depe = ht(ji + 1, jj) + sshn(ji + 1, jj);
    const double u_w = 0.5 * (un(ji, jj) + un(ji - 1, jj)) * e2t(ji, jj);
    const double depw = ht(ji, jj) + sshn(ji, jj);
    const double v_sc = 0.5 * (vn(ji, jj - 1) + vn(ji + 1, jj - 1));
    const double v_s = 0.5 * v_sc * (e1v(ji, jj - 1) + e1v(ji + 1, jj - 1));
    const double deps = 0.5 * (hv(ji, jj - 1) + sshn_v(ji, jj - 1) + hv(ji + 1, jj - 1) + sshn_v(ji + 1, jj - 1));
    const double v_nc = 0.5 * (vn(ji, jj) + vn(ji + 1, jj));
    const double v_n = 0.5 * v_nc * (e1v(ji, jj) + e1v(ji + 1, jj));
    const double depn = 0.5 * (hv(ji, jj) + sshn_v(ji, jj) + hv(ji + 1, jj) + sshn_v(ji + 1, jj));
    // advection (first-order upwind)
    const double uu_w = (0.5 - copysign(0.5, u_w)) * un(ji, jj) + (0.5 + copysign(0.5, u_w)) * un(ji - 1, jj);
    const double uu_e = (0.5 + copysign(0.5, u_e)) * un(ji, jj) + (0.5 - copysign(0.5, u_e)) * un(ji + 1, jj);
    double uu_s, uu_n;
    if (tmask(ji, jj - 1) <= 0 || tmask(ji + 1, jj - 1) <= 0)
        uu_s = (0.5 - copysign(0.5, v_s)) * un(ji, jj);
    else
        uu_s = (0.5 - copysign(0.5, v_s)) * un(ji, jj) + (0.5 + copysign(0.5, v_s)) * un(ji, jj - 1);
    if (tmask(ji, jj + 1) <= 0 || tmask(ji + 1, jj + 1) <= 0)
        uu_n = (0.5 + copysign(0.5, v_n)) * un(ji, jj);
    else
        uu_n = (0.5 + copysign(0.5, v_n)) * un(ji, jj) + (0.5 - copysign(0.5, v_n)) * un(ji, jj + 1);
    const double adv = uu_w * u_w * depw - uu_e * u_e * depe + uu_s * v_s * deps - uu_n * v_n * depn;
    // viscosity
    const double dudx_e = (un(ji + 1, jj) - un(ji, jj)) / e1t(ji + 1, jj) * (ht(ji + 1, jj) + sshn(ji + 1, jj));
    const double dudx_w = (un(ji, jj) - un(ji - 1, jj)) / e1t(ji, jj) * (ht(ji, jj) + sshn(ji, jj));
    double dudy_s = 0.0, dudy_n = 0.0;
    if (!(tmask(ji, jj - 1) <= 0 || tmask(ji + 1, jj - 1) <= 0))
        dudy_s = (un(ji, jj) - un(ji, jj - 1)) / (e2u(ji, jj) + e2u(ji, jj - 1)) *
                 (hu(ji, jj) + sshn_u(ji, jj) + hu(ji, jj - 1) + sshn_u(ji, jj - 1));
    if (!(tmask(ji, jj + 1) <= 0 || tmask(ji + 1, jj + 1) <= 0))
        dudy_n = (un(ji, jj + 1) - un(ji, jj)) / (e2u(ji, jj) + e2u(ji, jj + 1)) *
                 (hu(ji, jj) + sshn_u(ji, jj) + hu(ji, jj + 1) + sshn_u(ji, jj + 1));
    const double vis = ((dudx_e - dudx_w) * e2u(ji, jj) + (dudy_n - dudy_s) * e1u(ji, jj) * 0.5) * visc;
    // Coriolis, pressure gradient
    const double cor = 0.5 * (2.0 * omega * sin(gphiu(ji, jj) * d2r) * (v_sc + v_nc)) * e12u(ji, jj) *
                       (hu(ji, jj) + sshn_u(ji, jj));
    const double hpg = -g * (hu(ji, jj) + sshn_u(ji, jj)) * e2u(ji, jj) * (sshn(ji + 1, jj) - sshn(ji, jj));
    // linear bottom friction, implicit
    ua(ji, jj) = (un(ji, jj) * (hu(ji, jj) + sshn_u(ji, jj)) + rdt * (adv + vis + cor + hpg) / e12u(ji, jj)) /
                 (hu(ji, jj) + ssha_u(ji, jj)) / (1.0 + cbfr * rdt);
}
"""
MOMENTUM_U_META = (
    [field_arg(GO_READWRITE, GO_CU), field_arg(GO_READ, GO_CU), field_arg(GO_READ, GO_CV), field_arg(GO_READ, GO_CU),
     field_arg(GO_READ, GO_CV), field_arg(GO_READ, GO_CT), field_arg(GO_READ, GO_CU), field_arg(GO_READ, GO_CT),
     field_arg(GO_READ, GO_CU), field_arg(GO_READ, GO_CV)]
    + [scalar_arg(GO_R_SCALAR)] * 6
    + [grid_arg(GO_GRID_MASK_T), grid_arg(GO_GRID_DX_U), grid_arg(GO_GRID_DX_V), grid_arg(GO_GRID_DX_T),
       grid_arg(GO_GRID_DY_U), grid_arg(GO_GRID_DY_T), grid_arg(GO_GRID_AREA_U), grid_arg(GO_GRID_LAT_U)])

MOMENTUM_V = r"""
__device__ void momentum_v_code(int ji, int jj, dlesm_field va, dlesm_cfield un, dlesm_cfield vn, dlesm_cfield hu,
                                dlesm_cfield hv, dlesm_cfield ht, dlesm_cfield ssha_v, dlesm_cfield sshn,
                                dlesm_cfield sshn_u, dlesm_cfield sshn_v, double rdt, double cbfr, double visc,
                                double omega, double d2r, double g, dlesm_imask tmask, dlesm_cfield e1v,
                                dlesm_cfield e1t, dlesm_cfield e2u, dlesm_cfield e2v, dlesm_cfield e2t,
                                dlesm_cfield e12v, dlesm_cfield gphiv)
{
    if (tmask(ji, jj) + tmask(ji, jj + 1) <= 0) return;
    if (tmask(ji, jj) <= 0 || tmask(ji, jj + 1) <= 0) return;
    const double v_n = 0.5 * (vn(ji, jj) + vn(ji, jj + 1)) * e1t(ji, jj + 1);
    const double depn = ht(ji, jj + 1) + sshn(ji, jj + 1);
    const double v_s = 0.5 * (vn(ji, jj) + vn(ji, jj - 1)) * e1t(ji, jj);
    const double deps = ht(ji, jj) + sshn(ji, jj);
    const double u_wc = 0.5 * (un(ji - 1, jj) + un(ji - 1, jj + 1));
    const double u_w = 0.5 * u_wc * (e2u(ji - 1, jj) + e2u(ji - 1, jj + 1));
    const double depw = 0.5 * (hu(ji - 1, jj) + sshn_u(ji - 1, jj) + hu(ji - 1, jj + 1) + sshn_u(ji - 1, jj + 1));
    const double u_ec = 0.5 * (un(ji, jj) + un(ji, jj + 1));
    const double u_e = 0.5 * u_ec * (e2u(ji, jj) + e2u(ji, jj + 1));
    const double depe = 0.5 * (hu(ji, jj) + sshn_u(ji, jj) + hu(ji, jj + 1) + sshn_u(ji, jj + 1));
    const double vv_s = (0.5 - copysign(0.5, v_s)) * vn(ji, jj) + (0.5 + copysign(0.5, v_s)) * vn(ji, jj - 1);
    const double vv_n = (0.5 + copysign(0.5, v_n)) * vn(ji, jj) + (0.5 - copysign(0.5, v_n)) * vn(ji, jj + 1);
    double vv_w, vv_e;
    if (tmask(ji - 1, jj) <= 0 || tmask(ji - 1, jj + 1) <= 0)
        vv_w = (0.5 - copysign(0.5, u_w)) * vn(ji, jj);
    else
        vv_w = (0.5 - copysign(0.5, u_w)) * vn(ji, jj) + (0.5 + copysign(0.5, u_w)) * vn(ji - 1, jj);
    if (tmask(ji + 1, jj) <= 0 || tmask(ji + 1, jj + 1) <= 0)
        vv_e = (0.5 + copysign(0.5, u_e)) * vn(ji, jj);
    else
        vv_e = (0.5 + copysign(0.5, u_e)) * vn(ji, jj) + (0.5 - copysign(0.5, u_e)) * vn(ji + 1, jj);
    const double adv = vv_w * u_w * depw - vv_e * u_e * depe + vv_s * v_s * deps - vv_n * v_n * depn;
    const double dvdy_n = (vn(ji, jj + 1) - vn(ji, jj)) / e2t(ji, jj + 1) * (ht(ji, jj + 1) + sshn(ji, jj + 1));
    const double dvdy_s = (vn(ji, jj) - vn(ji, jj - 1)) / e2t(ji, jj) * (ht(ji, jj) + sshn(ji, jj));
    double dvdx_w = 0.0, dvdx_e = 0.0;
    if (!(tmask(ji - 1, jj) <= 0 || tmask(ji - 1, jj + 1) <= 0))
        dvdx_w = (vn(ji, jj) - vn(ji - 1, jj)) / (e1v(ji, jj) + e1v(ji - 1, jj)) *
                 (hv(ji, jj) + sshn_v(ji, jj) + hv(ji - 1, jj) + sshn_v(ji - 1, jj));
    if (!(tmask(ji + 1, jj) <= 0 || tmask(ji + 1, jj + 1) <= 0))
        dvdx_e = (vn(ji + 1, jj) - vn(ji, jj)) / (e1v(ji, jj) + e1v(ji + 1, jj)) *
                 (hv(ji, jj) + sshn_v(ji, jj) + hv(ji + 1, jj) + sshn_v(ji + 1, jj));
    const double vis = ((dvdy_n - dvdy_s) * e1v(ji, jj) + (dvdx_e - dvdx_w) * e2v(ji, jj) * 0.5) * visc;
    const double cor = -0.5 * (2.0 * omega * sin(gphiv(ji, jj) * d2r) * (u_ec + u_wc)) * e12v(ji, jj) *
                       (hv(ji, jj) + sshn_v(ji, jj));
    const double hpg = -g * (hv(ji, jj) + sshn_v(ji, jj)) * e1v(ji, jj) * (sshn(ji, jj + 1) - sshn(ji, jj));
    va(ji, jj) = (vn(ji, jj) * (hv(ji, jj) + sshn_v(ji, jj)) + rdt * (adv + vis + cor + hpg) / e12v(ji, jj)) /
                 (hv(ji, jj) + ssha_v(ji, jj)) / (1.0 + cbfr * rdt);
}
"""
MOMENTUM_V_META = (
    [field_arg(GO_READWRITE, GO_CV), field_arg(GO_READ, GO_CU), field_arg(GO_READ, GO_CV), field_arg(GO_READ, GO_CU),
     field_arg(GO_READ, GO_CV), field_arg(GO_READ, GO_CT), field_arg(GO_READ, GO_CV), field_arg(GO_READ, GO_CT),
     field_arg(GO_READ, GO_CU), field_arg(GO_READ, GO_CV)]
    + [scalar_arg(GO_R_SCALAR)] * 6
    + [grid_arg(GO_GRID_MASK_T), grid_arg(GO_GRID_DX_V), grid_arg(GO_GRID_DX_T), grid_arg(GO_GRID_DY_U),
       grid_arg(GO_GRID_DY_V), grid_arg(GO_GRID_DY_T), grid_arg(GO_GRID_AREA_V), grid_arg(GO_GRID_LAT_V)])

BC_SSH = r"""
__device__ void bc_ssh_code(int ji, int jj, int istep, dlesm_field ssha, double rdt, dlesm_imask tmask)
{
    const double amp_tide = 0.2, omega_tide = 2.0 * 3.14159265358979323846 / (12.42 * 3600.0);
    const double rtime = (double)istep * rdt;
    if (tmask(ji, jj) <= 0) return;
    if (tmask(ji, jj - 1) < 0 || tmask(ji, jj + 1) < 0 || tmask(ji + 1, jj) < 0 || tmask(ji - 1, jj) < 0)
        ssha(ji, jj) = amp_tide * sin(omega_tide * rtime);
}
"""
BC_SSH_META = [scalar_arg(fd.api.GO_I_SCALAR), field_arg(GO_READWRITE, GO_CT), scalar_arg(GO_R_SCALAR),
               grid_arg(GO_GRID_MASK_T)]

BC_SOLID_U = r"""
__device__ void bc_solid_u_code(int ji, int jj, dlesm_field ua, dlesm_imask tmask)
{
    if (tmask(ji, jj) * tmask(ji + 1, jj) == 0) ua(ji, jj) = 0.0;
}
"""
BC_SOLID_V = r"""
__device__ void bc_solid_v_code(int ji, int jj, dlesm_field va, dlesm_imask tmask)
{
    if (tmask(ji, jj) * tmask(ji, jj + 1) == 0) va(ji, jj) = 0.0;
}
"""
NEXT_SSHU = r"""
__device__ void next_sshu_code(int ji, int jj, dlesm_field sshn_u, dlesm_cfield sshn, dlesm_imask tmask,
                               dlesm_cfield e12t, dlesm_cfield e12u)
{
    if (tmask(ji, jj) + tmask(ji + 1, jj) <= 0) return;
    if (tmask(ji, jj) * tmask(ji + 1, jj) > 0) {
        const double rtmp1 = e12t(ji, jj) * sshn(ji, jj) + e12t(ji + 1, jj) * sshn(ji + 1, jj);
        sshn_u(ji, jj) = 0.5 * rtmp1 / e12u(ji, jj);
    } else if (tmask(ji, jj) <= 0) {
        sshn_u(ji, jj) = sshn(ji + 1, jj);
    } else if (tmask(ji + 1, jj) <= 0) {
        sshn_u(ji, jj) = sshn(ji, jj);
    }
}
"""
NEXT_SSHV = r"""
__device__ void next_sshv_code(int ji, int jj, dlesm_field sshn_v, dlesm_cfield sshn, dlesm_imask tmask,
                               dlesm_cfield e12t, dlesm_cfield e12v)
{
    if (tmask(ji, jj) + tmask(ji, jj + 1) <= 0) return;
    if (tmask(ji, jj) * tmask(ji, jj + 1) > 0) {
        const double rtmp1 = e12t(ji, jj) * sshn(ji, jj) + e12t(ji, jj + 1) * sshn(ji, jj + 1);
        sshn_v(ji, jj) = 0.5 * rtmp1 / e12v(ji, jj);
    } else if (tmask(ji, jj) <= 0) {
        sshn_v(ji, jj) = sshn(ji, jj + 1);
    } else if (tmask(ji, jj + 1) <= 0) {
        sshn_v(ji, jj) = sshn(ji, jj);
    }
}
"""
# the reference's one kernel (src/infrastructure_mod.f90:13-41): ITERATES_OVER = GO_ALL_PTS
FIELD_COPY = r"""
__device__ void field_copy_code(int ji, int jj, dlesm_field output, dlesm_cfield input)
{
    output(ji, jj) = input(ji, jj);
}
"""


class NemoLite2D:
    """The benchmark's time step as one `invoke` of eleven kernels over thirteen r2d_fields,
    plus the halo exchanges a distributed-memory PSy layer inserts."""

    POINTS = {"sshn": fd.GO_T_POINTS, "ssha": fd.GO_T_POINTS, "ht": fd.GO_T_POINTS,
              "sshn_u": fd.GO_U_POINTS, "ssha_u": fd.GO_U_POINTS, "hu": fd.GO_U_POINTS, "un": fd.GO_U_POINTS,
              "ua": fd.GO_U_POINTS,
              "sshn_v": fd.GO_V_POINTS, "ssha_v": fd.GO_V_POINTS, "hv": fd.GO_V_POINTS, "vn": fd.GO_V_POINTS,
              "va": fd.GO_V_POINTS}

    def __init__(self, grid, rdt=20.0, cbfr=0.00015, visc=0.1):
        self.grid, self.rdt, self.cbfr, self.visc = grid, rdt, cbfr, visc
        ne = fd.GO_OFFSET_NE
        mask = grid_arg(GO_GRID_MASK_T)
        self.k = {
            "continuity": kernel_type("continuity", CONTINUITY, CONTINUITY_META, GO_INTERNAL_PTS, ne),
            "momentum_u": kernel_type("momentum_u", MOMENTUM_U, MOMENTUM_U_META, GO_INTERNAL_PTS, ne),
            "momentum_v": kernel_type("momentum_v", MOMENTUM_V, MOMENTUM_V_META, GO_INTERNAL_PTS, ne),
            "bc_ssh": kernel_type("bc_ssh", BC_SSH, BC_SSH_META, GO_INTERNAL_PTS, ne),
            "bc_solid_u": kernel_type("bc_solid_u", BC_SOLID_U, [field_arg(GO_READWRITE, GO_CU), mask],
                                      GO_INTERNAL_PTS, ne),
            "bc_solid_v": kernel_type("bc_solid_v", BC_SOLID_V, [field_arg(GO_READWRITE, GO_CV), mask],
                                      GO_INTERNAL_PTS, ne),
            "next_sshu": kernel_type("next_sshu", NEXT_SSHU,
                                     [field_arg(GO_READWRITE, GO_CU), field_arg(GO_READ, GO_CT), mask,
                                      grid_arg(GO_GRID_AREA_T), grid_arg(GO_GRID_AREA_U)], GO_INTERNAL_PTS, ne),
            "next_sshv": kernel_type("next_sshv", NEXT_SSHV,
                                     [field_arg(GO_READWRITE, GO_CV), field_arg(GO_READ, GO_CT), mask,
                                      grid_arg(GO_GRID_AREA_T), grid_arg(GO_GRID_AREA_V)], GO_INTERNAL_PTS, ne),
            # src/infrastructure_mod.f90:13-26: GO_ALL_PTS, GO_OFFSET_ANY
            "copy": kernel_type("field_copy", FIELD_COPY,
                                [field_arg(GO_WRITE, GO_EVERY), field_arg(GO_READ, GO_EVERY)], GO_ALL_PTS,
                                fd.GO_OFFSET_ANY),
        }
        self.f = {n: fd.r2d_field(grid, p) for n, p in self.POINTS.items()}

    def set_state(self, arrays):
        for n, a in arrays.items():
            self.f[n].set_data(np.asfortranarray(a, dtype=np.float64))

    def step(self, istep):
        f, k = self.f, self.k
        phys = (self.rdt, self.cbfr, self.visc, OMEGA, D2R, G)
        k["continuity"]([f[n] for n in ("ssha", "sshn", "sshn_u", "sshn_v", "hu", "hv", "un", "vn")], (self.rdt,))
        k["momentum_u"]([f[n] for n in ("ua", "un", "vn", "hu", "hv", "ht", "ssha_u", "sshn", "sshn_u", "sshn_v")],
                        phys)
        k["momentum_v"]([f[n] for n in ("va", "un", "vn", "hu", "hv", "ht", "ssha_v", "sshn", "sshn_u", "sshn_v")],
                        phys)
        k["bc_ssh"]([f["ssha"]], (self.rdt,), (istep,))
        k["bc_solid_u"]([f["ua"]])
        k["bc_solid_v"]([f["va"]])
        k["copy"]([f["un"], f["ua"]])
        k["copy"]([f["vn"], f["va"]])
        k["copy"]([f["sshn"], f["ssha"]])
        if fd.get_num_ranks() > 1:
            fd.halo_exchange_fields([f["un"], f["vn"], f["sshn"]])
        k["next_sshu"]([f["sshn_u"], f["sshn"]])
        k["next_sshv"]([f["sshn_v"], f["sshn"]])
        if fd.get_num_ranks() > 1:
            fd.halo_exchange_fields([f["sshn_u"], f["sshn_v"]])

    def free(self):
        for fl in self.f.values():
            fl.free()
        for kk in self.k.values():
            kk.free()


def global_setup(n_x, n_y, seed=7):
    """Synthetic basin on the (n_x, n_y) global T grid, 1-based global indices as array index + 1:
    land ring (tmask 0) around the domain with an open southern boundary (tmask -1 below wet
    cells), one island; smooth depth; small random initial elevation and currents."""
    rng = np.random.default_rng(seed)
    tm = np.ones((n_x + 2, n_y + 2), dtype=np.int32, order="F")      # incl. the boundary ring
    tm[0, :] = tm[-1, :] = tm[:, -1] = 0
    tm[:, 0] = -1
    tm[0, 0] = tm[-1, 0] = 0
    tm[1, :] = tm[-2, :] = tm[:, -2] = 0                             # solid walls W, E, N
    ix, iy = n_x // 3, n_y // 2
    tm[ix:ix + max(2, n_x // 8), iy:iy + max(2, n_y // 8)] = 0       # island
    x = np.arange(n_x + 2)[:, None] / (n_x + 1.0)
    y = np.arange(n_y + 2)[None, :] / (n_y + 1.0)
    ht = np.asfortranarray(100.0 + 20.0 * np.sin(3.0 * x) * np.cos(2.0 * y))
    state = {"ht": ht, "hu": ht + 0.5, "hv": ht - 0.5}
    for nm, amp in (("sshn", 0.05), ("un", 0.01), ("vn", 0.01)):
        state[nm] = np.asfortranarray(amp * rng.uniform(-1.0, 1.0, size=(n_x + 2, n_y + 2)))
    state["sshn_u"] = np.asfortranarray(0.5 * state["sshn"])
    state["sshn_v"] = np.asfortranarray(0.25 * state["sshn"])
    state["ssha"] = state["sshn"].copy(order="F")
    state["ssha_u"] = state["sshn_u"].copy(order="F")
    state["ssha_v"] = state["sshn_v"].copy(order="F")
    state["ua"] = state["un"].copy(order="F")
    state["va"] = state["vn"].copy(order="F")
    return tm, state


def _local_index(grid):
    sub = grid.subdomain
    li = sub.internal
    ji = np.arange(li.xstart - 1, li.xstop + 2)
    jj = np.arange(li.ystart - 1, li.ystop + 2)
    return ji, jj, sub.glob.xstart + (ji - li.xstart), sub.glob.ystart + (jj - li.ystart)


def local_tmask(grid, tm):
    """this rank's part of the global mask, shaped like the tmask argument of grid_init
    (src/grid_mod.f90:330): call after grid%decompose, before grid_init.  Index (i, j) of the
    global array is global point (i, j), the boundary ring sitting at 0 and n+1."""
    sub = grid.subdomain
    ji, jj, gi, gj = _local_index(grid)
    ltm = np.zeros((sub.glob.nx, sub.glob.ny), dtype=np.int32, order="F")
    ltm[np.ix_(ji - 1, jj - 1)] = tm[np.ix_(gi, gj)]
    return ltm


def local_state(grid, state):
    """this rank's part of the global fields as (grid.nx, grid.ny) arrays: call after grid_init"""
    ji, jj, gi, gj = _local_index(grid)
    out = {}
    for nm, a in state.items():
        loc = np.zeros((grid.nx, grid.ny), order="F")
        loc[np.ix_(ji - 1, jj - 1)] = a[np.ix_(gi, gj)]
        out[nm] = loc
    return out

"""CPU oracle (TEST INFRASTRUCTURE -- never imported by the product) for the NEMOLite2D-style
kernels of examples/nemolite2d.py: the same per-point formulas as Fortran-style sweeps
`do jj = ystart, ystop; do ji = xstart, xstop` (SURVEY CS-3) over (nx, ny) column-major arrays,
written with numpy slices; every expression keeps the association of the kernel source, one
IEEE rounding per operation (numpy float64 never contracts to FMA).

Parity unpinned: the reference ships no stencil; the formulas are this project's
restatement of the public NEMOLite2D benchmark and this file is their normative statement.
sin() comes from libm here and from CUDA's math library on the device, so kernels containing it
(momentum, bc_ssh) are compared within 1e-12 relative, everything else bit for bit.
"""
from __future__ import annotations

import numpy as np


class Box:
    """1-based inclusive sweep bounds (a field's `internal` or `whole` region)"""

    def __init__(self, xstart, xstop, ystart, ystop):
        self.xs, self.xe, self.ys, self.ye = xstart, xstop, ystart, ystop

    def s(self, a, di=0, dj=0):
        """a(ji+di, jj+dj) for every (ji, jj) of the sweep"""
        return a[self.xs - 1 + di:self.xe + di, self.ys - 1 + dj:self.ye + dj]


def continuity(b, ssha, sshn, sshn_u, sshn_v, hu, hv, un, vn, rdt, e12t):
    s = b.s
    rtmp1 = (s(sshn_u) + s(hu)) * s(un)
    rtmp2 = (s(sshn_u, -1, 0) + s(hu, -1, 0)) * s(un, -1, 0)
    rtmp3 = (s(sshn_v) + s(hv)) * s(vn)
    rtmp4 = (s(sshn_v, 0, -1) + s(hv, 0, -1)) * s(vn, 0, -1)
    s(ssha)[...] = s(sshn) + (rtmp2 - rtmp1 + rtmp4 - rtmp3) * rdt / s(e12t)


def momentum_u(b, ua, un, vn, hu, hv, ht, ssha_u, sshn, sshn_u, sshn_v, rdt, cbfr, visc, omega, d2r, g, tmask,
               e1u, e1v, e1t, e2u, e2t, e12u, gphiu):
    s = b.s
    cs = np.copysign
    with np.errstate(all="ignore"):
        active = ~((s(tmask) + s(tmask, 1, 0) <= 0) | (s(tmask) <= 0) | (s(tmask, 1, 0) <= 0))
        u_e = 0.5 * (s(un) + s(un, 1, 0)) * s(e2t, 1, 0)
        depe = s(ht, 1, 0) + s(sshn, 1, 0)
        u_w = 0.5 * (s(un) + s(un, -1, 0)) * s(e2t)
        depw = s(ht) + s(sshn)
        v_sc = 0.5 * (s(vn, 0, -1) + s(vn, 1, -1))
        v_s = 0.5 * v_sc * (s(e1v, 0, -1) + s(e1v, 1, -1))
        deps = 0.5 * (s(hv, 0, -1) + s(sshn_v, 0, -1) + s(hv, 1, -1) + s(sshn_v, 1, -1))
        v_nc = 0.5 * (s(vn) + s(vn, 1, 0))
        v_n = 0.5 * v_nc * (s(e1v) + s(e1v, 1, 0))
        depn = 0.5 * (s(hv) + s(sshn_v) + s(hv, 1, 0) + s(sshn_v, 1, 0))
        uu_w = (0.5 - cs(0.5, u_w)) * s(un) + (0.5 + cs(0.5, u_w)) * s(un, -1, 0)
        uu_e = (0.5 + cs(0.5, u_e)) * s(un) + (0.5 - cs(0.5, u_e)) * s(un, 1, 0)
        south_land = (s(tmask, 0, -1) <= 0) | (s(tmask, 1, -1) <= 0)
        north_land = (s(tmask, 0, 1) <= 0) | (s(tmask, 1, 1) <= 0)
        uu_s = np.where(south_land, (0.5 - cs(0.5, v_s)) * s(un),
                        (0.5 - cs(0.5, v_s)) * s(un) + (0.5 + cs(0.5, v_s)) * s(un, 0, -1))
        uu_n = np.where(north_land, (0.5 + cs(0.5, v_n)) * s(un),
                        (0.5 + cs(0.5, v_n)) * s(un) + (0.5 - cs(0.5, v_n)) * s(un, 0, 1))
        adv = uu_w * u_w * depw - uu_e * u_e * depe + uu_s * v_s * deps - uu_n * v_n * depn
        dudx_e = (s(un, 1, 0) - s(un)) / s(e1t, 1, 0) * (s(ht, 1, 0) + s(sshn, 1, 0))
        dudx_w = (s(un) - s(un, -1, 0)) / s(e1t) * (s(ht) + s(sshn))
        dudy_s = np.where(south_land, 0.0, (s(un) - s(un, 0, -1)) / (s(e2u) + s(e2u, 0, -1)) *
                          (s(hu) + s(sshn_u) + s(hu, 0, -1) + s(sshn_u, 0, -1)))
        dudy_n = np.where(north_land, 0.0, (s(un, 0, 1) - s(un)) / (s(e2u) + s(e2u, 0, 1)) *
                          (s(hu) + s(sshn_u) + s(hu, 0, 1) + s(sshn_u, 0, 1)))
        vis = ((dudx_e - dudx_w) * s(e2u) + (dudy_n - dudy_s) * s(e1u) * 0.5) * visc
        cor = 0.5 * (2.0 * omega * np.sin(s(gphiu) * d2r) * (v_sc + v_nc)) * s(e12u) * (s(hu) + s(sshn_u))
        hpg = -g * (s(hu) + s(sshn_u)) * s(e2u) * (s(sshn, 1, 0) - s(sshn))
        new = (s(un) * (s(hu) + s(sshn_u)) + rdt * (adv + vis + cor + hpg) / s(e12u)) / (s(hu) + s(ssha_u)) / \
              (1.0 + cbfr * rdt)
        s(ua)[...] = np.where(active, new, s(ua))


def momentum_v(b, va, un, vn, hu, hv, ht, ssha_v, sshn, sshn_u, sshn_v, rdt, cbfr, visc, omega, d2r, g, tmask,
               e1v, e1t, e2u, e2v, e2t, e12v, gphiv):
    s = b.s
    cs = np.copysign
    with np.errstate(all="ignore"):
        active = ~((s(tmask) + s(tmask, 0, 1) <= 0) | (s(tmask) <= 0) | (s(tmask, 0, 1) <= 0))
        v_n = 0.5 * (s(vn) + s(vn, 0, 1)) * s(e1t, 0, 1)
        depn = s(ht, 0, 1) + s(sshn, 0, 1)
        v_s = 0.5 * (s(vn) + s(vn, 0, -1)) * s(e1t)
        deps = s(ht) + s(sshn)
        u_wc = 0.5 * (s(un, -1, 0) + s(un, -1, 1))
        u_w = 0.5 * u_wc * (s(e2u, -1, 0) + s(e2u, -1, 1))
        depw = 0.5 * (s(hu, -1, 0) + s(sshn_u, -1, 0) + s(hu, -1, 1) + s(sshn_u, -1, 1))
        u_ec = 0.5 * (s(un) + s(un, 0, 1))
        u_e = 0.5 * u_ec * (s(e2u) + s(e2u, 0, 1))
        depe = 0.5 * (s(hu) + s(sshn_u) + s(hu, 0, 1) + s(sshn_u, 0, 1))
        vv_s = (0.5 - cs(0.5, v_s)) * s(vn) + (0.5 + cs(0.5, v_s)) * s(vn, 0, -1)
        vv_n = (0.5 + cs(0.5, v_n)) * s(vn) + (0.5 - cs(0.5, v_n)) * s(vn, 0, 1)
        west_land = (s(tmask, -1, 0) <= 0) | (s(tmask, -1, 1) <= 0)
        east_land = (s(tmask, 1, 0) <= 0) | (s(tmask, 1, 1) <= 0)
        vv_w = np.where(west_land, (0.5 - cs(0.5, u_w)) * s(vn),
                        (0.5 - cs(0.5, u_w)) * s(vn) + (0.5 + cs(0.5, u_w)) * s(vn, -1, 0))
        vv_e = np.where(east_land, (0.5 + cs(0.5, u_e)) * s(vn),
                        (0.5 + cs(0.5, u_e)) * s(vn) + (0.5 - cs(0.5, u_e)) * s(vn, 1, 0))
        adv = vv_w * u_w * depw - vv_e * u_e * depe + vv_s * v_s * deps - vv_n * v_n * depn
        dvdy_n = (s(vn, 0, 1) - s(vn)) / s(e2t, 0, 1) * (s(ht, 0, 1) + s(sshn, 0, 1))
        dvdy_s = (s(vn) - s(vn, 0, -1)) / s(e2t) * (s(ht) + s(sshn))
        dvdx_w = np.where(west_land, 0.0, (s(vn) - s(vn, -1, 0)) / (s(e1v) + s(e1v, -1, 0)) *
                          (s(hv) + s(sshn_v) + s(hv, -1, 0) + s(sshn_v, -1, 0)))
        dvdx_e = np.where(east_land, 0.0, (s(vn, 1, 0) - s(vn)) / (s(e1v) + s(e1v, 1, 0)) *
                          (s(hv) + s(sshn_v) + s(hv, 1, 0) + s(sshn_v, 1, 0)))
        vis = ((dvdy_n - dvdy_s) * s(e1v) + (dvdx_e - dvdx_w) * s(e2v) * 0.5) * visc
        cor = -0.5 * (2.0 * omega * np.sin(s(gphiv) * d2r) * (u_ec + u_wc)) * s(e12v) * (s(hv) + s(sshn_v))
        hpg = -g * (s(hv) + s(sshn_v)) * s(e1v) * (s(sshn, 0, 1) - s(sshn))
        new = (s(vn) * (s(hv) + s(sshn_v)) + rdt * (adv + vis + cor + hpg) / s(e12v)) / (s(hv) + s(ssha_v)) / \
              (1.0 + cbfr * rdt)
        s(va)[...] = np.where(active, new, s(va))


def bc_ssh(b, istep, ssha, rdt, tmask):
    s = b.s
    amp_tide, omega_tide = 0.2, 2.0 * 3.14159265358979323846 / (12.42 * 3600.0)
    rtime = float(istep) * rdt
    hit = (s(tmask) > 0) & ((s(tmask, 0, -1) < 0) | (s(tmask, 0, 1) < 0) | (s(tmask, 1, 0) < 0) |
                            (s(tmask, -1, 0) < 0))
    s(ssha)[...] = np.where(hit, amp_tide * np.sin(omega_tide * rtime), s(ssha))


def bc_solid_u(b, ua, tmask):
    s = b.s
    s(ua)[...] = np.where(s(tmask) * s(tmask, 1, 0) == 0, 0.0, s(ua))


def bc_solid_v(b, va, tmask):
    s = b.s
    s(va)[...] = np.where(s(tmask) * s(tmask, 0, 1) == 0, 0.0, s(va))


def _next_ssh(b, out, sshn, tmask, e12t, e12x, di, dj):
    s = b.s
    with np.errstate(all="ignore"):
        t0, t1 = s(tmask), s(tmask, di, dj)
        avg = 0.5 * (s(e12t) * s(sshn) + s(e12t, di, dj) * s(sshn, di, dj)) / s(e12x)
        new = np.where(t0 * t1 > 0, avg, np.where(t0 <= 0, s(sshn, di, dj), np.where(t1 <= 0, s(sshn), s(out))))
        s(out)[...] = np.where(t0 + t1 <= 0, s(out), new)


def next_sshu(b, sshn_u, sshn, tmask, e12t, e12u):
    _next_ssh(b, sshn_u, sshn, tmask, e12t, e12u, 1, 0)


def next_sshv(b, sshn_v, sshn, tmask, e12t, e12v):
    _next_ssh(b, sshn_v, sshn, tmask, e12t, e12v, 0, 1)


def field_copy(b, output, inp):
    """src/infrastructure_mod.f90:32-41 over GO_ALL_PTS"""
    b.s(output)[...] = b.s(inp)


GRID_NAMES = {"e1t": "dx_t", "e1u": "dx_u", "e1v": "dx_v", "e2t": "dy_t", "e2u": "dy_u", "e2v": "dy_v",
              "e12t": "area_t", "e12u": "area_u", "e12v": "area_v", "gphiu": "gphiu", "gphiv": "gphiv"}


class Model:
    """The benchmark's invoke (same kernel order as examples.nemolite2d.NemoLite2D.step, without
    the halo exchanges: one serial domain) on host arrays.
    arrays: name -> (nx, ny) array; metrics: e1t.. -> (nx, ny) array; boxes: 't','u','v' internal
    regions and 't_all','u_all','v_all' whole regions as Box."""

    def __init__(self, arrays, metrics, tmask, boxes, rdt, cbfr, visc, omega, d2r, g):
        self.a = {k: np.array(v, order="F", copy=True) for k, v in arrays.items()}
        self.g, self.tmask, self.box = metrics, tmask, boxes
        self.rdt, self.phys = rdt, (rdt, cbfr, visc, omega, d2r, g)

    @classmethod
    def serial_global(cls, n_x, n_y, tm, state, dx, dy, rdt, cbfr, visc, omega, d2r, g):
        """one serial NE / external-BC domain of n_x x n_y points (src/parallel_mod.f90:272-290 with
        one tile, halo width 1; src/grid_mod.f90:368-385 with DL_ESM_ALIGNMENT=1): arrays are
        (n_x + 3, n_y + 3), internal region 2:n+1, the caller's ring-included global arrays sit
        at 1:n+2."""
        nx, ny = n_x + 3, n_y + 3

        def pad(a, dtype):
            out = np.zeros((nx, ny), dtype=dtype, order="F")
            out[:n_x + 2, :n_y + 2] = a
            return out
        arrays = {k: pad(v, np.float64) for k, v in state.items()}
        tmask = pad(tm, np.int32)
        tmask[n_x + 2, :] = tmask[n_x + 1, :]       # high-x / high-y fill of grid_init (:416-431)
        tmask[:, n_y + 2] = tmask[:, n_y + 1]
        const = {"e1t": dx, "e1u": dx, "e1v": dx, "e2t": dy, "e2u": dy, "e2v": dy, "e12t": dx * dy, "e12u": dx * dy,
                 "e12v": dx * dy, "gphiu": 50.0, "gphiv": 50.0}      # src/grid_mod.f90:461-503
        metrics = {k: np.full((nx, ny), v, order="F") for k, v in const.items()}
        boxes = {}
        for nm in ("t", "u", "v"):                  # NE offset, external BCs: src/field_mod.f90:755-1000
            boxes[nm] = Box(2, n_x + 1, 2, n_y + 1)
            boxes[nm + "_all"] = Box(1, n_x + 2, 1, n_y + 2)
        return cls(arrays, metrics, tmask, boxes, rdt, cbfr, visc, omega, d2r, g)

    def step(self, istep):
        a, g, b, tm = self.a, self.g, self.box, self.tmask
        continuity(b["t"], a["ssha"], a["sshn"], a["sshn_u"], a["sshn_v"], a["hu"], a["hv"], a["un"], a["vn"],
                   self.rdt, g["e12t"])
        momentum_u(b["u"], a["ua"], a["un"], a["vn"], a["hu"], a["hv"], a["ht"], a["ssha_u"], a["sshn"],
                   a["sshn_u"], a["sshn_v"], *self.phys, tm, g["e1u"], g["e1v"], g["e1t"], g["e2u"], g["e2t"],
                   g["e12u"], g["gphiu"])
        momentum_v(b["v"], a["va"], a["un"], a["vn"], a["hu"], a["hv"], a["ht"], a["ssha_v"], a["sshn"],
                   a["sshn_u"], a["sshn_v"], *self.phys, tm, g["e1v"], g["e1t"], g["e2u"], g["e2v"], g["e2t"],
                   g["e12v"], g["gphiv"])
        bc_ssh(b["t"], istep, a["ssha"], self.rdt, tm)
        bc_solid_u(b["u"], a["ua"], tm)
        bc_solid_v(b["v"], a["va"], tm)
        field_copy(b["u_all"], a["un"], a["ua"])
        field_copy(b["v_all"], a["vn"], a["va"])
        field_copy(b["t_all"], a["sshn"], a["ssha"])
        next_sshu(b["u"], a["sshn_u"], a["sshn"], tm, g["e12t"], g["e12u"])
        next_sshv(b["v"], a["sshn_v"], a["sshn"], tm, g["e12t"], g["e12v"])

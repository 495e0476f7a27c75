"""TEST INFRASTRUCTURE (only tests/ may import this; never the product).

An INDEPENDENT restatement of the published shallow-water benchmark the GOcean "shallow"
application implements -- P. N. Swarztrauber's NCAR `shallow.f` (Sadourny 1975's potential-
enstrophy conserving C-grid scheme, leapfrog in time with a Robert-Asselin filter) -- written
from the published loop nests, on the bare periodic n_x x n_y core with numpy index rolls: no
halo cells, no regions, no fd_lib containers, no SURVEY a-S table.  The reference repository
holds no stencil code, so oracle/dlesm_oracle_sw.c is "parity unpinned" by it;
this file is the second opinion that pins the oracle's stencil arithmetic to the published
algorithm: tests/test_oracle_sw.py demands oracle == this, and the GPU tests demand CUDA ==
oracle bit for bit.

shallow.f, loop 100 / 200 / 300 (arrays there are (M+1, N+1) with the periodic copies made by
separate loops; here the wrap is the roll):

    CU(I+1,J)   = .5*(P(I+1,J)+P(I,J))*U(I+1,J)
    CV(I,J+1)   = .5*(P(I,J+1)+P(I,J))*V(I,J+1)
    Z(I+1,J+1)  = (FSDX*(V(I+1,J+1)-V(I,J+1)) - FSDY*(U(I+1,J+1)-U(I+1,J)))
                  / (P(I,J)+P(I+1,J)+P(I+1,J+1)+P(I,J+1))
    H(I,J)      = P(I,J)+.25*(U(I+1,J)*U(I+1,J)+U(I,J)*U(I,J)+V(I,J+1)*V(I,J+1)+V(I,J)*V(I,J))

    UNEW(I+1,J) = UOLD(I+1,J) + TDTS8*(Z(I+1,J+1)+Z(I+1,J))*(CV(I+1,J+1)+CV(I,J+1)+CV(I,J)+CV(I+1,J))
                  - TDTSDX*(H(I+1,J)-H(I,J))
    VNEW(I,J+1) = VOLD(I,J+1) - TDTS8*(Z(I+1,J+1)+Z(I,J+1))*(CU(I+1,J+1)+CU(I,J+1)+CU(I,J)+CU(I+1,J))
                  - TDTSDY*(H(I,J+1)-H(I,J))
    PNEW(I,J)   = POLD(I,J) - TDTSDX*(CU(I+1,J)-CU(I,J)) - TDTSDY*(CV(I,J+1)-CV(I,J))

    first cycle:  TDT = TDT+TDT ; UOLD = U ; U = UNEW (same for V, P)
    later cycles: UOLD = U+ALPHA*(UNEW-2.*U+UOLD) ; U = UNEW (same for V, P)

with FSDX = 4/DX, FSDY = 4/DY, TDTS8 = TDT/8, TDTSDX = TDT/DX, TDTSDY = TDT/DY.

Index convention.  In shallow.f, U(I+1,J) sits on the face between P(I,J) and P(I+1,J), V(I,J+1)
between P(I,J) and P(I,J+1), Z(I+1,J+1) on the corner shared by P(I,J) .. P(I+1,J+1).  Renaming
I+1 -> i for U, CU, Z and J+1 -> j for V, CV, Z puts every quantity at "its own" (i, j) with the
T point (i, j) to the north-east of the F point (i, j): GO_OFFSET_SW, the stagger the product's
shallow-water layer requires (src/grid_mod.f90:52-60).  `W(a)` below is a(i-1, j), `S(a)` is
a(i, j-1), `E`, `N` the opposite neighbours.
"""
from __future__ import annotations

import numpy as np


def W(a):
    return np.roll(a, 1, axis=0)


def E(a):
    return np.roll(a, -1, axis=0)


def S(a):
    return np.roll(a, 1, axis=1)


def N(a):
    return np.roll(a, -1, axis=1)


class PublishedShallow:
    """state on the periodic core: u, v, p and uold, vold, pold, each (n_x, n_y)"""

    def __init__(self, u, v, p, dx=1.0e5, dy=1.0e5, dt=90.0, alpha=0.001):
        self.u, self.v, self.p = (np.array(a, dtype=np.float64, order="F") for a in (u, v, p))
        self.uold, self.vold, self.pold = self.u.copy(), self.v.copy(), self.p.copy()
        self.dx, self.dy, self.dt, self.tdt, self.alpha = dx, dy, dt, dt, alpha
        self.ncycle = 0

    def diagnostics(self):
        """cu, cv at their U / V points, z at F points, h at T points (renamed indices, see above)"""
        u, v, p = self.u, self.v, self.p
        fsdx, fsdy = 4.0 / self.dx, 4.0 / self.dy
        cu = 0.5 * (p + W(p)) * u
        cv = 0.5 * (p + S(p)) * v
        z = (fsdx * (v - W(v)) - fsdy * (u - S(u))) / (S(W(p)) + S(p) + p + W(p))
        h = p + 0.25 * (E(u) * E(u) + u * u + N(v) * N(v) + v * v)
        return cu, cv, z, h

    def step(self):
        cu, cv, z, h = self.diagnostics()
        tdts8, tdtsdx, tdtsdy = self.tdt / 8.0, self.tdt / self.dx, self.tdt / self.dy
        unew = self.uold + tdts8 * (N(z) + z) * (N(cv) + N(W(cv)) + W(cv) + cv) - tdtsdx * (h - W(h))
        vnew = self.vold - tdts8 * (E(z) + z) * (E(cu) + cu + S(cu) + S(E(cu))) - tdtsdy * (h - S(h))
        pnew = self.pold - tdtsdx * (E(cu) - cu) - tdtsdy * (N(cv) - cv)
        if self.ncycle == 0:
            self.tdt = self.tdt + self.tdt
            self.uold, self.vold, self.pold = self.u, self.v, self.p
        else:
            a = self.alpha
            self.uold = self.u + a * (unew - 2.0 * self.u + self.uold)
            self.vold = self.v + a * (vnew - 2.0 * self.v + self.vold)
            self.pold = self.p + a * (pnew - 2.0 * self.p + self.pold)
        self.u, self.v, self.p = unew, vnew, pnew
        self.ncycle += 1


# ---------------------------------------------------------------------------------------
# Invariants of the Sadourny (1975) potential-enstrophy conserving scheme on the periodic core.
# The spatial discretisation conserves mass and potential enstrophy exactly (energy to second
# order); leapfrog + Robert-Asselin adds a time-truncation drift, so tests bound the drift.
# ---------------------------------------------------------------------------------------
def invariants(u, v, p, dx=1.0e5, dy=1.0e5):
    """(mass, total energy, potential enstrophy) of a state on the periodic core (n_x, n_y)"""
    u, v, p = (np.asarray(a, dtype=np.float64) for a in (u, v, p))
    mass = float(np.sum(p, dtype=np.longdouble))
    # kinetic energy at T points from the four surrounding faces (the Bernoulli function's average),
    # potential energy 0.5 p^2 (p is g*h: the benchmark's "pressure")
    ke = 0.25 * (E(u) * E(u) + u * u + N(v) * N(v) + v * v)
    energy = float(np.sum(p * ke + 0.5 * p * p, dtype=np.longdouble))
    # potential vorticity q = zeta / pbar at F points; potential enstrophy 0.5 * q^2 * pbar
    pbar = 0.25 * (S(W(p)) + S(p) + p + W(p))
    zeta = (v - W(v)) / dx - (u - S(u)) / dy
    q = zeta / pbar
    enstrophy = float(np.sum(0.5 * q * q * pbar, dtype=np.longdouble))
    return mass, energy, enstrophy

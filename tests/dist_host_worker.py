"""CPU multi-process check of the host-side distributed logic (gloo, world_size >= 2).

Every rank builds its own decomposition, grid, field bounds and map_comms tables through the
product's C ABI (host-only entry points: no CUDA is touched), derives the per-peer message
plan the device exchange uses (dlesm_comm_plan_build), and then really exchanges halos with
its peers over gloo -- numpy pack in plan order, one buffer per peer, numpy unpack.  The
result must equal the oracle's exchange_generic run over all ranks (bit for bit, edges and
corners).  Also checks that the tables of different ranks pair up (every send has the matching
receive on the peer) and that the NCCL-id bootstrap plumbing (128 bytes from rank 1 to all)
works over the process group.

This file is test infrastructure: the numpy pack/unpack here stands in for the CUDA pack/unpack
kernels and gloo for NCCL, so that the plan logic is exercised by real processes on a CPU box.
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import dl_esm_inf_b200 as fd  # noqa: E402
from dl_esm_inf_b200.api import comm_plan  # noqa: E402
from oracle import oracle as O  # noqa: E402

RANK = int(os.environ.get("RANK", 0))
WORLD = int(os.environ.get("WORLD_SIZE", 1))
BCS_EXT = (fd.GO_BC_EXTERNAL, fd.GO_BC_EXTERNAL, fd.GO_BC_NONE)
BCS_PER = (fd.GO_BC_PERIODIC, fd.GO_BC_PERIODIC, fd.GO_BC_NONE)
NCHECK = 0


def ok(cond, msg):
    global NCHECK
    if not cond:
        raise AssertionError("rank %d: %s" % (RANK + 1, msg))
    NCHECK += 1


def pack(a, c, slots, send=True):
    out = []
    for s in slots:
        if send:
            i0, j0, nx, ny = c.isrcsend[s], c.jsrcsend[s], c.nxsend[s], c.nysend[s]
        else:
            i0, j0, nx, ny = c.idesrecv[s], c.jdesrecv[s], c.nxrecv[s], c.nyrecv[s]
        out.append(a[i0 - 1:i0 - 1 + nx, j0 - 1:j0 - 1 + ny].flatten(order="F"))   # i fastest
    return np.concatenate(out) if out else np.zeros(0)


def unpack(a, c, slots, buf):
    q = 0
    for s in slots:
        i0, j0, nx, ny = c.idesrecv[s], c.jdesrecv[s], c.nxrecv[s], c.nyrecv[s]
        a[i0 - 1:i0 - 1 + nx, j0 - 1:j0 - 1 + ny] = buf[q:q + nx * ny].reshape((nx, ny), order="F")
        q += nx * ny


def gloo_exchange(a, c, plan):
    """what dlesm_dev_halo_exchange does, with gloo for NCCL and numpy for the kernels"""
    ops, recvs, selfbuf = [], [], None
    for g in range(plan.npeers):
        peer = plan.peer[g]
        sslots = [plan.send_slot[g][q] for q in range(plan.nsend[g])]
        rslots = [plan.recv_slot[g][q] for q in range(plan.nrecv[g])]
        sbuf = torch.from_numpy(pack(a, c, sslots))
        ok(sbuf.numel() == plan.send_n[g], "send_n of the plan")
        if peer == RANK:
            selfbuf = (rslots, sbuf.numpy().copy())
            continue
        rbuf = torch.zeros(int(plan.recv_n[g]), dtype=torch.float64)
        if sbuf.numel():
            ops.append(dist.P2POp(dist.isend, sbuf, peer))
        if rbuf.numel():
            ops.append(dist.P2POp(dist.irecv, rbuf, peer))
        recvs.append((rslots, rbuf))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    for rslots, rbuf in recvs:
        unpack(a, c, rslots, rbuf.numpy())
    if selfbuf:
        unpack(a, c, *selfbuf)


def tables_pair_up(c):
    """every send of mine has the matching receive on the peer: same size, same direction"""
    mine = [(c.destination[s], c.dirsend[s], c.nxsend[s], c.nysend[s]) for s in range(c.nsend)]
    theirs = [(c.source[r], c.dirrecv[r], c.nxrecv[r], c.nyrecv[r]) for r in range(c.nrecv)]
    gathered_s, gathered_r = [None] * WORLD, [None] * WORLD
    dist.all_gather_object(gathered_s, mine)
    dist.all_gather_object(gathered_r, theirs)
    for dst, d, nx, ny in mine:
        ok((RANK, d, nx, ny) in gathered_r[dst], "send to rank %d dir %d has no matching receive" % (dst + 1, d))
    for src, d, nx, ny in theirs:
        ok((RANK, d, nx, ny) in gathered_s[src], "receive from rank %d dir %d has no matching send" % (src + 1, d))


def case(nx, ny, offset, bcs, periodic, alignment):
    if alignment:
        os.environ["DL_ESM_ALIGNMENT"] = str(alignment)
    else:
        os.environ.pop("DL_ESM_ALIGNMENT", None)
    fd.set_extensions(fd.EXT_PERIODIC_DISTMEM if periodic else 0)
    g = fd.grid_type(fd.GO_ARAKAWA_C, bcs, offset)
    g.decompose(nx, ny)
    try:
        # grid_init installs the tables into the device context (host memory only)
        fd.grid_init(g, 1.0, 1.0)
    except fd.GoceanStop as e:  # pragma: no cover
        raise AssertionError("grid_init failed: %s" % e)
    c = g.comms
    plan = comm_plan(c)
    tables_pair_up(c)
    rng = np.random.default_rng(1000 + nx * 7 + ny)
    if periodic:
        # The reference (and therefore the oracle's map_comms) refuses periodic BCs with more
        # than one rank; the wrap-around tables are this project's extension and
        # are defined by the serial semantics: after the exchange every halo cell, corners
        # included, holds the value of the periodic global field at the wrapped position.
        G = np.asfortranarray(rng.uniform(-1, 1, size=(nx, ny)))
        sub = g.subdomain
        li = sub.internal
        a = np.full((g.nx, g.ny), -7.0, order="F")
        gi = sub.glob.xstart + np.arange(li.nx) - 1
        gj = sub.glob.ystart + np.arange(li.ny) - 1
        a[li.xstart - 1:li.xstop, li.ystart - 1:li.ystop] = G[np.ix_(gi, gj)]
        gloo_exchange(a, c, plan)
        ji = np.arange(li.xstart - 1, li.xstop + 2)
        jj = np.arange(li.ystart - 1, li.ystop + 2)
        wi = (sub.glob.xstart + (ji - li.xstart) - 1) % nx
        wj = (sub.glob.ystart + (jj - li.ystart) - 1) % ny
        ok(np.array_equal(a[np.ix_(ji - 1, jj - 1)], G[np.ix_(wi, wj)]),
           "periodic wrap-around exchange (%dx%d) does not reproduce the serial periodic halo" % (nx, ny))
    else:
        # the oracle's view of every rank
        ogs, arrs = [], []
        for r in range(1, WORLD + 1):
            O.reset_saved_state()
            og = O.Grid(O.GO_ARAKAWA_C, bcs, offset)
            og.decompose(nx, ny, rank=r, nranks=WORLD)
            og.init(1.0, 1.0, alignment=alignment or 1, dist_mem_enabled=True, metrics=False)
            ogs.append(og)
        ok((g.nx, g.ny) == (ogs[RANK].nx, ogs[RANK].ny), "allocation shape")
        oc = ogs[RANK].comms
        for name in ("nsend", "nrecv"):
            ok(getattr(c, name) == getattr(oc, name), "%s differs from the oracle" % name)
        for og in ogs:
            arrs.append(np.asfortranarray(rng.uniform(-1, 1, size=(og.nx, og.ny))))
        mine = arrs[RANK].copy(order="F")
        O.exchange_all(arrs, [og.comms for og in ogs])
        gloo_exchange(mine, c, plan)
        ok(np.array_equal(mine, arrs[RANK]),
           "gloo exchange by the product's plan differs from the oracle (%dx%d)" % (nx, ny))
    os.environ.pop("DL_ESM_ALIGNMENT", None)
    fd.set_extensions(0)


def case_depth(nx, ny, depth):
    """extension: tables for halo_exchange(depth) on a grid decomposed with halo_width = depth.
    Known answer: every halo cell within `depth` that lies inside the global domain ends up
    holding the global field's value at its position (edges and corners)."""
    os.environ.pop("DL_ESM_ALIGNMENT", None)
    g = fd.grid_type(fd.GO_ARAKAWA_C, BCS_EXT, fd.GO_OFFSET_NE)
    g.decompose(nx, ny, halo_width=depth)
    fd.grid_init(g, 1.0, 1.0)
    c = fd.map_comms(g.decomp, halo_depths=(depth, depth))
    plan = comm_plan(c)
    tables_pair_up(c)
    G = np.asfortranarray(np.random.default_rng(77 + nx + ny).uniform(-1, 1, size=(nx, ny)))
    sub = g.subdomain
    li = sub.internal
    ok(li.xstart == depth + 1, "halo_width shifts the internal region")
    a = np.full((g.nx, g.ny), -7.0, order="F")
    gi = sub.glob.xstart + np.arange(li.nx) - 1
    gj = sub.glob.ystart + np.arange(li.ny) - 1
    a[li.xstart - 1:li.xstop, li.ystart - 1:li.ystop] = G[np.ix_(gi, gj)]
    want = a.copy(order="F")
    gloo_exchange(a, c, plan)
    for ji in range(li.xstart - depth, li.xstop + depth + 1):
        for jj in range(li.ystart - depth, li.ystop + depth + 1):
            gx, gy = sub.glob.xstart + ji - li.xstart, sub.glob.ystart + jj - li.ystart
            if 1 <= gx <= nx and 1 <= gy <= ny:
                want[ji - 1, jj - 1] = G[gx - 1, gy - 1]
    ok(np.array_equal(a, want), "depth-%d exchange (%dx%d) does not reproduce the global field" % (depth, nx, ny))
    # deeper than the decomposition's halo width is refused (reference: ierr from map_comms)
    try:
        fd.map_comms(g.decomp, halo_depths=(depth + 1, depth + 1))
        ok(False, "depth beyond the halo width must be refused")
    except fd.GoceanStop:
        ok(True, "")
    fd.gocean_initialise(RANK + 1, WORLD)            # go_decompose's SAVEd halo width back to 1


def case_depth_periodic(nx, ny, depth=2):
    """the tables the two-step sweep runs on (both extensions): periodic wrap AND depth 2 on a tile
    with halo_width 2 -- used by its halo_exchange(depth 2) and, entry for entry, as its push
    boxes.  Known answer: every cell within `depth` of the internal region, corners included,
    holds the periodic global field at the wrapped position; source boxes lie inside the internal
    region (so pushing them from the sweep equals exchanging them afterwards)."""
    os.environ.pop("DL_ESM_ALIGNMENT", None)
    fd.set_extensions(fd.EXT_PERIODIC_DISTMEM | fd.EXT_HALO_DEPTH)
    g = fd.grid_type(fd.GO_ARAKAWA_C, (fd.GO_BC_PERIODIC, fd.GO_BC_PERIODIC, fd.GO_BC_NONE), fd.GO_OFFSET_SW)
    g.decompose(nx, ny, halo_width=depth)
    fd.grid_init(g, 1.0, 1.0)
    c = fd.map_comms(g.decomp, pbc=(True, True), halo_depths=(depth, depth))
    plan = comm_plan(c)
    tables_pair_up(c)
    sub = g.subdomain
    li = sub.internal
    for q in range(c.nsend):
        ok(li.xstart <= c.isrcsend[q] and c.isrcsend[q] + c.nxsend[q] - 1 <= li.xstop and
           li.ystart <= c.jsrcsend[q] and c.jsrcsend[q] + c.nysend[q] - 1 <= li.ystop,
           "send box %d leaves the internal region" % q)
    G = np.asfortranarray(np.random.default_rng(99 + nx + ny).uniform(-1, 1, size=(nx, ny)))
    a = np.full((g.nx, g.ny), -7.0, order="F")
    gi = sub.glob.xstart + np.arange(li.nx) - 1
    gj = sub.glob.ystart + np.arange(li.ny) - 1
    a[li.xstart - 1:li.xstop, li.ystart - 1:li.ystop] = G[np.ix_(gi, gj)]
    gloo_exchange(a, c, plan)
    ji = np.arange(li.xstart - depth, li.xstop + depth + 1)
    jj = np.arange(li.ystart - depth, li.ystop + depth + 1)
    wi = (sub.glob.xstart + (ji - li.xstart) - 1) % nx
    wj = (sub.glob.ystart + (jj - li.ystart) - 1) % ny
    ok(np.array_equal(a[np.ix_(ji - 1, jj - 1)], G[np.ix_(wi, wj)]),
       "periodic depth-%d exchange (%dx%d) does not reproduce the serial periodic halo" % (depth, nx, ny))
    fd.set_extensions(0)
    fd.gocean_initialise(RANK + 1, WORLD)            # go_decompose's SAVEd halo width back to 1


def case_tileset(nx, ny, tiles_per_rank, periodic, depth=1):
    """extension F3: several sub-domains per rank (src/decomposition_mod.f90:64-67; the \\TODO of
    src/parallel_comms_mod.f90:295).  Each rank owns proc_subdomains(:, rank); its exchange plan
    (dlesm_tileset_plan) turns same-rank neighbours into copies and groups the rest per peer RANK.
    Driven over gloo with numpy for the copy kernel.  Known answer: every halo cell that lies inside
    the (periodic: wrapped) global domain holds the global field's value at its position."""
    from dl_esm_inf_b200.api import tileset_plan
    os.environ.pop("DL_ESM_ALIGNMENT", None)
    ndom = tiles_per_rank * WORLD
    d = fd.go_decompose(nx, ny, ndomains=ndom, halo_width=depth)
    ok(d.domperrank == tiles_per_rank, "domperrank")
    owned = [d.proc_subdomains(s, RANK + 1) for s in range(1, tiles_per_rank + 1)]
    ok(owned == list(range(RANK * tiles_per_rank + 1, (RANK + 1) * tiles_per_rank + 1)),
       "proc_subdomains(:, rank) as go_decompose fills it (src/parallel_mod.f90:141-152)")
    rows, (send_total, recv_total) = tileset_plan(d, RANK + 1, WORLD, pbc=(periodic, periodic), depth=depth)
    subs = d.subdomains
    G = np.asfortranarray(np.random.default_rng(5 + nx * 3 + ny).uniform(-1, 1, size=(nx, ny)))
    arrs = []
    for tid in owned:
        sb = subs[tid - 1]
        li = sb.internal
        a = np.full((sb.glob.nx, sb.glob.ny), -7.0, order="F")     # the tile's full extent incl. halos
        gi = sb.glob.xstart + np.arange(li.nx) - 1
        gj = sb.glob.ystart + np.arange(li.ny) - 1
        a[li.xstart - 1:li.xstop, li.ystart - 1:li.ystop] = G[np.ix_(gi, gj)]
        arrs.append(a)
    before = [a.copy() for a in arrs]
    sendbuf, recvbuf = np.zeros(send_total), np.zeros(recv_total)
    peers_s, peers_r = {}, {}
    for kind, peer, st, dt_, si0, sj0, di0, dj0, w, h, off, dst_id, src_id, dirn in rows.tolist():
        if kind == 0:       # same-rank neighbour: a copy that never leaves the device
            ok(peer == RANK, "copy segments stay on this rank")
            arrs[dt_][di0 - 1:di0 - 1 + w, dj0 - 1:dj0 - 1 + h] = before[st][si0 - 1:si0 - 1 + w, sj0 - 1:sj0 - 1 + h]
        elif kind == 1:
            sendbuf[off:off + w * h] = before[st][si0 - 1:si0 - 1 + w, sj0 - 1:sj0 - 1 + h].flatten(order="F")
            lo, n = peers_s.get(peer, (off, 0))
            peers_s[peer] = (min(lo, off), n + w * h)
        else:
            lo, n = peers_r.get(peer, (off, 0))
            peers_r[peer] = (min(lo, off), n + w * h)
    ok(set(peers_s) == set(peers_r), "a rank sends to exactly the ranks it receives from")
    ops, views = [], []
    for peer, (lo, n) in sorted(peers_s.items()):
        ops.append(dist.P2POp(dist.isend, torch.from_numpy(sendbuf[lo:lo + n].copy()), peer))
    for peer, (lo, n) in sorted(peers_r.items()):
        t = torch.zeros(n, dtype=torch.float64)
        views.append((lo, n, t))
        ops.append(dist.P2POp(dist.irecv, t, peer))
    if ops:
        for w_ in dist.batch_isend_irecv(ops):
            w_.wait()
    for lo, n, t in views:
        recvbuf[lo:lo + n] = t.numpy()
    for kind, peer, st, dt_, si0, sj0, di0, dj0, w, h, off, dst_id, src_id, dirn in rows.tolist():
        if kind == 2:
            arrs[dt_][di0 - 1:di0 - 1 + w, dj0 - 1:dj0 - 1 + h] = recvbuf[off:off + w * h].reshape((w, h), order="F")
    for tid, a in zip(owned, arrs):
        sb = subs[tid - 1]
        li = sb.internal
        for ji in range(li.xstart - depth, li.xstop + depth + 1):
            for jj in range(li.ystart - depth, li.ystop + depth + 1):
                gx, gy = sb.glob.xstart + ji - li.xstart, sb.glob.ystart + jj - li.ystart
                if periodic:
                    want = G[(gx - 1) % nx, (gy - 1) % ny]
                elif 1 <= gx <= nx and 1 <= gy <= ny:
                    want = G[gx - 1, gy - 1]
                else:
                    want = -7.0
                ok(a[ji - 1, jj - 1] == want, "tile %d cell (%d,%d) after the tile-set exchange (%dx%d, %d tiles "
                   "per rank, periodic=%s, depth %d)" % (tid, ji, jj, nx, ny, tiles_per_rank, periodic, depth))
    fd.gocean_initialise(RANK + 1, WORLD)            # go_decompose's SAVEd halo width back to 1


def main():
    dist.init_process_group("gloo")
    fd.gocean_initialise(RANK + 1, WORLD)            # host only: no device, no NCCL
    ok(fd.get_rank() == RANK + 1 and fd.get_num_ranks() == WORLD, "rank bookkeeping")
    ok(fd.on_master() == (RANK == 0), "on_master")
    # NCCL-id plumbing: 128 bytes from rank 1 to everyone through the process group
    blob = torch.zeros(128, dtype=torch.uint8)
    if RANK == 0:
        blob = torch.arange(128, dtype=torch.uint8)
    dist.broadcast(blob, 0)
    ok(bytes(blob.numpy().tobytes()) == bytes(range(128)), "id broadcast")
    for shape in ((10, 4), (4, 10), (10, 10), (37, 23), (64, 48)):
        case(*shape, fd.GO_OFFSET_NE, BCS_EXT, False, None)
    case(10, 10, fd.GO_OFFSET_NE, BCS_EXT, False, 8)
    for shape in ((16, 16), (33, 47), (64, 96)):
        case(*shape, fd.GO_OFFSET_SW, BCS_PER, True, 2)
    for shape in ((10, 10), (37, 23), (64, 48)):
        case_depth(*shape, 2)
    case_depth(40, 36, 3)
    for shape in ((16, 24), (37, 23), (64, 48)):
        case_depth_periodic(*shape)
    for tpr in (2, 3):
        case_tileset(48, 36, tpr, False)
        case_tileset(48, 36, tpr, True)
    case_tileset(37, 61, 2, True)
    case_tileset(64, 48, 2, True, depth=2)
    case_tileset(40, 40, 4, False, depth=2)
    dist.barrier()
    if RANK == 0:
        print("HOST DIST OK %d checks on %d ranks" % (NCHECK, WORLD), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

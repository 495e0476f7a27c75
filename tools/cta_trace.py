"""Per-CTA timeline of ONE launch of a fused sweep (diagnostic; dlesm_debug_cta_trace_begin/end).

    python tools/cta_trace.py --size 8192 --sweep two --halo push
    python -m torch.distributed.run --nproc-per-node 4 ... tools/cta_trace.py --size 16384 --scaling strong

Which CTAs make the tail of a short sweep?  Classes: interior, x-edge (first / last strip: under halo push
these lanes store the neighbours' halo columns on every row), y-edge (first / last segment), corner.  Prints
per rank: launch span, per-class CTA count and duration percentiles, the wait for the neighbours, and what
the last CTAs to finish were.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=8192)
    ap.add_argument("--scaling", choices=["weak", "strong"], default="weak")
    ap.add_argument("--sweep", choices=["one", "two"], default="two")
    ap.add_argument("--halo", choices=["push", "exchange"], default="push")
    ap.add_argument("--warm", type=int, default=9)
    ap.add_argument("--launches", type=int, default=1,
                    help="sweeps enqueued back to back while tracing; every launch overwrites the table, so the LAST "
                         "one is what is reported (the host is ahead of the device by then: steady state)")
    ap.add_argument("--dump", default="")
    ap.add_argument("--tmask", action="store_true", help="BASELINE config 3's land islands (bench.config3_mask_for)")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    os.environ["DL_ESM_ALIGNMENT"] = "64"

    import dl_esm_inf_b200 as fd
    from dl_esm_inf_b200 import _lib
    L = _lib.lib()
    dist = torch = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    fd.gocean_initialise(rank + 1, world, device=local)
    if world > 1:
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt = torch.frombuffer(bytearray(fd.comm_unique_id()), dtype=torch.uint8).cuda()
        dist.broadcast(idt, 0)
        fd.gocean_initialise(rank + 1, world, device=local, comm_id=bytes(idt.cpu().numpy().tobytes()))

    two = args.sweep == "two"
    n = args.size
    probe = fd.go_decompose(n, n)
    gx, gy = (n, n) if args.scaling == "strong" else (n * probe.nx, n * probe.ny)
    fd.gocean_initialise(rank + 1, world)
    ext = fd.EXT_PERIODIC_DISTMEM if world > 1 else 0
    bcs = (fd.GO_BC_PERIODIC, fd.GO_BC_PERIODIC, fd.GO_BC_NONE)
    grid = fd.grid_type(fd.GO_ARAKAWA_C, bcs, fd.GO_OFFSET_SW)
    if two:
        fd.set_extensions(ext | (fd.EXT_HALO_DEPTH if world > 1 else 0))
        grid.decompose(gx, gy, halo_width=2)
    else:
        fd.set_extensions(ext)
        grid.decompose(gx, gy)
    tm = None
    if args.tmask:
        import bench
        tm = bench.config3_mask_for(grid, gx, gy, np)
    fd.grid_init(grid, 1.0e5, 1.0e5, tm)
    sw = fd.shallow_water(grid, dt=90.0, alpha=0.001, mode=fd.SW_FUSED, use_tmask=bool(args.tmask))
    if two:
        sw.set_two_step(True, 0)
    sw.init_device()
    if args.halo == "push":
        sw.set_halo_mode(fd.HALO_PUSH)
    sw.step(args.warm)          # odd: the first (unfiltered) step + pairs
    fd.dev_sync()
    g = sw.launch_grid(two_step=two)
    ncta = int(g[0]) * int(g[1])
    if dist is not None:
        dist.barrier()
    assert L.dlesm_debug_cta_trace_begin(ncta) == 0
    sw.step((2 if two else 1) * args.launches)
    buf = (C.c_longlong * (4 * ncta))()
    assert L.dlesm_debug_cta_trace_end(buf, ncta) == 0
    t = np.frombuffer(buf, dtype=np.int64).reshape(ncta, 4).copy()
    sub = grid.subdomain
    gxb, gyb = int(g[0]), int(g[1])
    bx = np.arange(ncta) % gxb
    by = np.arange(ncta) // gxb
    start, end, sm, waited = t[:, 0], t[:, 1], t[:, 2], t[:, 3]
    t0 = start.min()
    dur = (end - start) * 1e-3
    xe = (bx == 0) | (bx == gxb - 1)
    ye = (by == 0) | (by == gyb - 1)
    cls = np.where(xe & ye, 3, np.where(xe, 1, np.where(ye, 2, 0)))
    names = ["interior", "x-edge", "y-edge", "corner"]
    rec = {"rank": rank + 1, "tile": "%dx%d" % (sub.internal.nx, sub.internal.ny), "grid": [gxb, gyb], "ctas": ncta,
           "sweep": args.sweep, "halo": args.halo, "span_us": float((end.max() - t0) * 1e-3),
           "sms_used": int(len(np.unique(sm)))}
    for c, name in enumerate(names):
        m = cls == c
        if m.any():
            d = dur[m]
            rec[name] = {"n": int(m.sum()), "dur_us_p50": float(np.median(d)), "dur_us_p95": float(np.percentile(d, 95)),
                         "dur_us_max": float(d.max()), "wait_us_max": float(((waited - start)[m]).max() * 1e-3),
                         "last_end_us": float((end[m].max() - t0) * 1e-3)}
    # when does the bulk finish, and what runs in the tail?
    e_us = (end - t0) * 1e-3
    s_us = (start - t0) * 1e-3
    order = np.argsort(e_us)
    rec["end_us_p50_p90_p99_max"] = [float(np.percentile(e_us, q)) for q in (50, 90, 99, 100)]
    rec["last_start_us"] = float(s_us.max())
    rec["tail_ctas"] = [{"bx": int(bx[q]), "by": int(by[q]), "class": names[cls[q]], "start_us": float(s_us[q]),
                         "dur_us": float(dur[q])} for q in order[-6:]]
    # slow CTAs (the land tiles of a split masked sweep): how many, how long, when they start and end
    med = float(np.median(dur))
    slow = dur > 1.3 * med
    rec["slow_ctas"] = {"n": int(slow.sum()), "dur_us_p50": float(np.median(dur[slow])) if slow.any() else None,
                        "start_us_p10_p50_p90": [float(np.percentile(s_us[slow], q)) for q in (10, 50, 90)] if slow.any() else None,
                        "end_us_p50_max": [float(np.percentile(e_us[slow], 50)), float(e_us[slow].max())] if slow.any() else None}
    rec["fast_ctas"] = {"n": int((~slow).sum()), "dur_us_p50": float(np.median(dur[~slow])),
                        "end_us_p50_max": [float(np.percentile(e_us[~slow], 50)), float(e_us[~slow].max())]}
    # occupancy profile: CTAs running at 10 points of the span
    span = rec["span_us"]
    rec["running_at_tenths"] = [int(((s_us <= f * span) & (e_us > f * span)).sum()) for f in np.linspace(0.05, 0.95, 10)]
    out = [rec]
    if dist is not None:
        out = [None] * world
        dist.all_gather_object(out, rec)
    if rank == 0:
        for r in out:
            print(json.dumps(r))
    if args.dump and rank == 0:
        np.save(args.dump, t)
    sw.free()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

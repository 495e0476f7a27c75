"""BASELINE config 5: halo_exchange + field_checksum microbench (width-1/2 halos, local tiles
2048^2 .. 32768^2), plus the per-kernel figures of SURVEY 8(d): copy_field 16 B/pt, set_field 8,
field_checksum 8, the un-fused shallow-water sweeps 24..40, the periodic halo fill.

    python tools/microbench.py [--sizes 2048,4096,...] [--reps 20]
    python -m torch.distributed.run --nproc-per-node N ... tools/microbench.py     (N > 1: NCCL exchange)

One JSON line per measurement on rank 0.  Timing: CUDA events on the library's stream
(dlesm_event_record), `reps` back-to-back calls after 3 warm-up calls, max over ranks.  Inputs are
>= 8 x the 126 MB L2 except at 2048^2 (34 MB/field: L2-resident, flagged "l2_resident").
HBM fraction is against MEASURED_PEAKS.json's copy bandwidth; halo traffic is reported against
900 GB/s per direction of NVLink 5.
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
    ap.add_argument("--sizes", default="2048,4096,8192,16384,32768")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--alignment", type=int, default=64)
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    os.environ["DL_ESM_ALIGNMENT"] = str(args.alignment)

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
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            peak = float(json.load(fh)["hbm_gbs"])
    except Exception:
        peak = 6550.0

    def timed(fn, reps):
        for _ in range(3):
            fn()
        fd.dev_sync()
        if dist is not None:
            dist.barrier()
        L.dlesm_event_record(6)
        for _ in range(reps):
            fn()
        L.dlesm_event_record(7)
        ms = C.c_double(0.0)
        assert L.dlesm_event_elapsed_ms(6, 7, C.byref(ms)) == 0
        el = ms.value
        if dist is not None:
            t = torch.tensor([el], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            el = float(t.item())
        return el / reps * 1e3          # microseconds per call

    def emit(**kw):
        if rank == 0:
            print(json.dumps(kw), flush=True)

    def hbm(op, n, bytes_per_pt, us, points, **extra):
        gbs = bytes_per_pt * points / (us * 1e-6) / 1e9
        emit(op=op, local_tile="%dx%d" % (n, n), n_gpus=world, us_per_call=us, bytes_per_point=bytes_per_pt,
             achieved_gbs=gbs, hbm_frac=gbs / peak, peak_gbs=peak, **extra)

    for n in [int(x) for x in args.sizes.split(",")]:
        # ---------- field ops + un-fused sweeps on one periodic SW tile per GPU ----------
        fd.set_extensions(fd.EXT_PERIODIC_DISTMEM)
        probe = fd.go_decompose(n, n)
        g = fd.grid_type(fd.GO_ARAKAWA_C, (fd.GO_BC_PERIODIC, fd.GO_BC_PERIODIC, fd.GO_BC_NONE), fd.GO_OFFSET_SW)
        g.decompose(n * probe.nx, n * probe.ny)
        fd.grid_init(g, 1.0e5, 1.0e5)
        alloc_pts = g.nx * g.ny
        l2 = {"l2_resident": alloc_pts * 8 < 126e6}
        reps = args.reps if n <= 16384 else max(5, args.reps // 2)
        a, b = fd.r2d_field(g, fd.GO_T_POINTS), fd.r2d_field(g, fd.GO_T_POINTS)
        fd.set_field(a, 1.0)
        i = a.internal
        hbm("set_field", n, 8, timed(lambda: fd.set_field(b, 2.5), reps), alloc_pts, **l2)
        hbm("copy_field", n, 16, timed(lambda: fd.copy_field(a, b), reps), alloc_pts, **l2)
        val = [0.0]

        def cks():
            val[0] = fd.field_checksum(a)
        us = timed(cks, reps)
        want = float(n) * n * world
        assert abs(val[0] - want) <= 1e-12 * want, (val[0], want)      # test_gsum's known answer
        hbm("field_checksum", n, 8, us, i.nx * i.ny, includes="device reduction + %s + D2H of the result"
            % ("ncclAllReduce" if world > 1 else "no allreduce"), known_answer_ok=True, **l2)
        if world == 1:
            hbm("periodic_halo_fill", n, 16, timed(a.apply_periodic_halos, reps), 2 * i.nx + 2 * (i.ny + 2),
                note="halo cells only: latency-bound", **l2)
        # PSy layer: generated sweeps (NVRTC) -- the reference's copy kernel over GO_ALL_PTS and the
        # NEMOLite2D continuity kernel (9 arrays read, 1 written: 80 B/pt) on this grid
        from dl_esm_inf_b200.api import (GO_ALL_PTS, GO_CT, GO_CU, GO_CV, GO_EVERY, GO_GRID_AREA_T, GO_INTERNAL_PTS,
                                         GO_R_SCALAR, GO_READ, GO_WRITE, field_arg, grid_arg, kernel_type, scalar_arg)
        kc = kernel_type("field_copy", "__device__ void field_copy_code(int ji, int jj, dlesm_field o, dlesm_cfield i)"
                         " { o(ji, jj) = i(ji, jj); }", [field_arg(GO_WRITE, GO_EVERY), field_arg(GO_READ, GO_EVERY)],
                         GO_ALL_PTS)
        w = a.whole
        hbm("psy_field_copy", n, 16, timed(lambda: kc([b, a]), reps), w.nx * w.ny, how="generated sweep (NVRTC)", **l2)
        kc.free()
        if n <= 16384:
            cont = """__device__ void cont_code(int ji, int jj, dlesm_field ssha, dlesm_cfield sshn, dlesm_cfield sshn_u,
                dlesm_cfield sshn_v, dlesm_cfield hu, dlesm_cfield hv, dlesm_cfield un, dlesm_cfield vn, double rdt,
                dlesm_cfield e12t) {
                const double r1 = (sshn_u(ji, jj) + hu(ji, jj)) * un(ji, jj);
                const double r2 = (sshn_u(ji - 1, jj) + hu(ji - 1, jj)) * un(ji - 1, jj);
                const double r3 = (sshn_v(ji, jj) + hv(ji, jj)) * vn(ji, jj);
                const double r4 = (sshn_v(ji, jj - 1) + hv(ji, jj - 1)) * vn(ji, jj - 1);
                ssha(ji, jj) = sshn(ji, jj) + (r2 - r1 + r4 - r3) * rdt / e12t(ji, jj); }"""
            kk = kernel_type("cont", cont, [field_arg(GO_WRITE, GO_CT), field_arg(GO_READ, GO_CT),
                                            field_arg(GO_READ, GO_CU), field_arg(GO_READ, GO_CV),
                                            field_arg(GO_READ, GO_CU), field_arg(GO_READ, GO_CV),
                                            field_arg(GO_READ, GO_CU), field_arg(GO_READ, GO_CV),
                                            scalar_arg(GO_R_SCALAR), grid_arg(GO_GRID_AREA_T)], GO_INTERNAL_PTS)
            extra = [fd.r2d_field(g, fd.GO_T_POINTS) for _ in range(6)]
            for q, f in enumerate(extra):
                fd.set_field(f, 1.0 + 0.1 * q)
            flds = [b, a] + extra
            hbm("psy_continuity", n, 80, timed(lambda: kk(flds, (20.0,)), reps), i.nx * i.ny,
                how="generated sweep (NVRTC): 9 arrays read (incl. area_t), 1 written", **l2)
            kk.free()
            for f in extra:
                f.free()
        a.free()
        b.free()
        if n <= 16384:
            sw = fd.shallow_water(g, mode=fd.SW_UNFUSED)
            sw.init_device()
            pts = i.nx * i.ny
            hbm("sw_step_unfused", n, 368, timed(lambda: sw.step(1), max(3, reps // 4)), pts,
                note="one launch per reference-style kernel + 3 copy_field (368 B/pt model)", **l2)
            sw.free()
            sw = fd.shallow_water(g, mode=fd.SW_FUSED)
            sw.init_device()
            if world > 1:
                sw.set_halo_mode(fd.HALO_PUSH)
            hbm("sw_step_fused", n, 96, timed(lambda: sw.step(1), reps), pts, **l2)
            sw.free()
        fd.set_extensions(0)

        # ---------- halo_exchange, NE offset / external BCs (the reference's dist-mem setup) ----------
        if world > 1:
            for depth in (1, 2):
                fd.gocean_initialise(rank + 1, world)          # SAVEd halo width back to 1
                fd.set_extensions(fd.EXT_HALO_DEPTH)
                g2 = fd.grid_type(fd.GO_ARAKAWA_C, (fd.GO_BC_EXTERNAL, fd.GO_BC_EXTERNAL, fd.GO_BC_NONE),
                                  fd.GO_OFFSET_NE)
                g2.decompose(n * probe.nx, n * probe.ny, halo_width=depth)
                fd.grid_init(g2, 1.0, 1.0)
                c = fd.map_comms(g2.decomp, halo_depths=(depth, depth))
                sent = sum(c.nxsend[s] * c.nysend[s] for s in range(c.nsend)) * 8
                recv = sum(c.nxrecv[s] * c.nyrecv[s] for s in range(c.nrecv)) * 8
                for nf in (1, 3):
                    fl = [fd.r2d_field(g2, fd.GO_T_POINTS) for _ in range(nf)]
                    for f in fl:
                        fd.set_field(f, float(rank + 1))
                    for mode, how in ((fd.EXCHANGE_DIRECT, "direct: send kernel stores the patches into the neighbours' "
                                       "receive slots over NVLink + flag, receive kernel waits and unpacks"),
                                      (fd.EXCHANGE_NCCL, "pack kernel -> grouped ncclSend/ncclRecv -> unpack kernel")):
                        fd.set_exchange_mode(mode)
                        us = timed(lambda: fd.halo_exchange_fields(fl, depth), reps)
                        emit(op="halo_exchange", local_tile="%dx%d" % (n, n), n_gpus=world, depth=depth, fields=nf,
                             tiles="%dx%d" % (probe.nx, probe.ny), us_per_call=us, rank1_bytes_sent=sent * nf,
                             rank1_bytes_received=recv * nf, messages=c.nsend,
                             nvlink_frac_of_900GBs_per_dir=max(sent, recv) * nf / (us * 1e-6) / 900e9,
                             mode="direct" if mode == fd.EXCHANGE_DIRECT else "nccl", how=how)
                    fd.set_exchange_mode(fd.EXCHANGE_DIRECT)
                    for f in fl:
                        f.free()
                fd.set_extensions(0)
            fd.gocean_initialise(rank + 1, world)
    if dist is not None:
        dist.barrier()
        fd.gocean_finalise()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

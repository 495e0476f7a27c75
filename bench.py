#!/usr/bin/env python
"""bench.py -- grid-point updates/s of the shallow-water step over fd_lib fields on B200.

    python bench.py --gpus N --steps K --warmup W              (our arm)
    python bench.py --impl reference --gpus N --steps K --warmup W   (reference arm: CPU)

Workload (BASELINE.json metric/config): FP64 shallow-water C-grid step, double periodic,
GO_OFFSET_SW, 16384^2 points PER GPU (weak scaling: the global domain grows with the
go_decompose tile grid 1x1, 1x2, 2x2, 2x4), DL_ESM_ALIGNMENT=64, synthetic classic "shallow"
initial state.  One bench step = one model time step of the fused sweep + halo fill /
exchange.  One JSON line on stdout (rank 0).

Besides the timed K steps the line carries, all outside the timed region:
  parity     a pre-flight parity leg on the SAME ranks and the same halo path as the timed leg: the
             reference's distributed known answers (example/model.f90, tests/dist_mem/test_halos.f90,
             test_gsum.f90, test_reduction.f90) and a 1024x1536 shallow-water domain (with and without
             a tmask, one-step and two-step sweep), gathered to rank 1 and compared bit for bit with the
             serial oracle; any mismatch ends the run with exit code 1
  sustained  >= --sustain-s seconds of back-to-back steps with their own clocks record (the driver's
             --steps 20 is a 40 ms burst, shorter than the board's power-management time constant)
  config3    BASELINE configs[2]: 4096^2 with tmask land islands, value + roofline + an oracle
             comparison of every field at full size (N=1)
  cpu_baseline  and the assertion that the GPU checksums after the same number of steps agree with
             the CPU arm's within 1e-12 at full size (N=1)
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "grid-point updates/s per shallow-water step"
UNIT = "points/s"
BYTES_PER_POINT = 96.0          # 6 FP64 reads + 6 FP64 writes per internal point
FALLBACK_HBM_GBS = 6650.0       # B200_PROFILING.md fallback when MEASURED_PEAKS.json is absent


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


# ---------------------------------------------------------------------------------------------
# clocks sampler (NVML) during the timed region
# ---------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    def __init__(self, device_index, period=0.02):
        super().__init__(daemon=True)
        self.samples, self.period, self.stop_flag = [], period, False
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.max_mhz = None

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self.stop_flag:
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                try:
                    reasons = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    reasons = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append((time.time(), mhz, reasons))
            except Exception:
                pass
            time.sleep(self.period)

    def nvlink_kib(self):
        """cumulative NVLink payload / raw traffic of this GPU, all links: (data_tx, data_rx, raw_tx, raw_rx)
        in KiB (NVML_FI_DEV_NVLINK_THROUGHPUT_*), or None where the driver does not report them"""
        if not self.ok:
            return None
        try:
            nv = self.nv
            ids = [nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX, nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX,
                   nv.NVML_FI_DEV_NVLINK_THROUGHPUT_RAW_TX, nv.NVML_FI_DEV_NVLINK_THROUGHPUT_RAW_RX]
            vals = nv.nvmlDeviceGetFieldValues(self.h, ids)
            out = []
            for v in vals:
                if v.nvmlReturn != 0:
                    return None
                out.append(int(v.value.ullVal))
            return out
        except Exception:
            return None

    def summary(self, t0, t1):
        if not self.ok:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable"]}
        inside = [s for s in self.samples if t0 <= s[0] <= t1] or self.samples[-3:]
        mhz = sorted(s[1] for s in inside)
        bits = 0
        for s in inside:
            bits |= s[2]
        names = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
                 0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
                 0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}
        reasons = [n for b, n in names.items() if bits & b and n != "gpu_idle"]
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None, "sm_max_mhz": self.max_mhz,
                "reasons": reasons, "samples": len(inside)}


def measured_hbm_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the oracle port of the reference's CPU formulation
# ---------------------------------------------------------------------------------------------
WORKLOAD = ("shallow-water C-grid step, %dx%d points per GPU (weak scaling over go_decompose tiles), double periodic, "
            "GO_OFFSET_SW, FP64, DL_ESM_ALIGNMENT=%d, fused mode")


def cpu_run(global_n, nthreads, steps, warmup):
    """returns points/s, seconds, the reference-style checksums (double-precision sum(abs()) per
    sub-domain + global_sum) and the same sums accumulated in long double"""
    from oracle import oracle as O
    secs, cks = O.sw_run_threads(global_n, global_n, nthreads, steps, warmup)
    return global_n * global_n * steps / secs, secs, cks, O.sw_last_checksums_ld()


def pick_cpu_sample(nthreads, steps, warmup, budget_s):
    """largest square sample whose (steps+warmup) steps fit the time budget"""
    rate = cpu_run(1024, nthreads, 3, 1)[0]              # calibration: points/s
    best = 1024
    for n in (2048, 4096, 8192, 16384):
        est = n * n * (steps + warmup) / rate * 1.3 + 13 * n * n * 8 / 4e9   # + first-touch init
        try:
            avail = os.sysconf("SC_AVPHYS_PAGES") * os.sysconf("SC_PAGE_SIZE")
        except (ValueError, OSError):
            avail = 1 << 62
        if est <= budget_s and 13 * n * n * 8 * 1.2 < avail:
            best = n
    return best


def arm_config(args, workload=None):
    """the `config` object of BOTH arms, key for key and value for value (the driver compares them): what is
    measured and how L2 is kept out of it.  Everything that belongs to one arm only goes to `config_notes`."""
    return {"workload": workload or WORKLOAD % (args.size, args.size, args.alignment),
            "l2": "GPU arm: inputs larger than L2, never flushed (each of the 9-12 resident fields is %.2f GB per GPU); "
                  "CPU arm: a bounded sample of the same workload, see cpu_baseline.sample"
                  % (args.size * args.size * 8 / 1e9)}


def run_reference_arm(args):
    rank = env_int("RANK", 0)
    if rank != 0:
        return 0        # rank 0 alone runs the CPU arm
    ncores = os.cpu_count() or 1
    n = pick_cpu_sample(ncores, args.steps, args.warmup, budget_s=90.0)
    value, secs, cks, cks_ld = cpu_run(n, ncores, args.steps, args.warmup)
    sample = ("%dx%d double-periodic sample of the %d^2-per-GPU workload, %d timed steps, %d threads "
              "(one go_decompose sub-domain each, exchange_generic-style pack/copy/unpack)%s"
              % (n, n, args.size, args.steps, ncores,
                 "; at --gpus %d the GPU arm's domain is %d such tiles: the CPU arm times ONE tile's worth of "
                 "points on this host (throughput metric, same points/s unit)" % (args.gpus, args.gpus)
                 if args.gpus > 1 else ""))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * secs / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": arm_config(args),
        "config_notes": {"reference_sample": sample,
                         "reference_build": "oracle port of the reference's Fortran loops (no Fortran/MPI toolchain "
                                            "in this image: oracle/_ref cannot be built)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": ncores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "checksums_p_u_v": list(cks), "checksums_p_u_v_long_double": list(cks_ld),
    }
    print(json.dumps(line), flush=True)
    return 0


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------
def config3_mask_for(grid, gx, gy, np):
    """BASELINE config 3: land on 64x64 islands where mod(i,512) in [200,263] and mod(j,512) in
    [300,363] (~1.6 % land), periodic; this tile's part over its whole extent incl. halos"""
    sb = grid.subdomain
    hw = sb.internal.xstart - 1
    gi = (sb.glob.xstart - 1 - hw + np.arange(sb.glob.nx)) % gx + 1
    gj = (sb.glob.ystart - 1 - hw + np.arange(sb.glob.ny)) % gy + 1
    li = (gi % 512 >= 200) & (gi % 512 <= 263)
    lj = (gj % 512 >= 300) & (gj % 512 <= 363)
    return np.asfortranarray(np.where(li[:, None] & lj[None, :], 0, 1).astype(np.int32))


def run_parity(args, fd, dist, torch, world, rank):
    """Pre-flight parity leg, outside every timed region, on the ranks and the halo path of the timed
    leg.  Cases and checks are those of tests/dist_worker.py (the reference's distributed tests +
    the periodic shallow-water model against the serial oracle, fields gathered with
    gather_inner_data and compared bit for bit)."""
    tdir = os.path.join(ROOT, "tests")
    if tdir not in sys.path:
        sys.path.insert(0, tdir)
    import dist_worker as DW
    DW.COLLECT = True
    t0 = time.time()
    cases = []

    def run(name, fn, *a, **k):
        fn(*a, **k)
        cases.append(name)

    # the reference's own known answers on `world` ranks
    run("example/model.f90:54-109", DW.case_example_model)
    run("tests/dist_mem/test_gsum.f90:107-132", DW.case_gsum)
    for shape in ((10, 4), (4, 10), (10, 10), (37, 23)):
        run("tests/dist_mem/test_halos.f90:153-265 %dx%d" % shape, DW.case_halos, *shape)
    run("test_halos.f90 64x48 DL_ESM_ALIGNMENT=64", DW.case_halos, 64, 48, alignment=64)
    run("tests/dist_mem/test_reduction.f90:128-192", DW.case_reduction, 10)
    run("parallel_utils_mod gather", DW.case_gather)
    # the timed path against the serial oracle: 1024x1536, DL_ESM_ALIGNMENT=64
    push = fd.HALO_PUSH if (world > 1 or args.halo == "push") and args.halo != "exchange" else None
    hname = "halo push" if push is not None else ("periodic halo kernel" if world == 1 else "NCCL exchange")
    nx, ny, A = 1024, 1536, args.alignment
    for masked in (False, True):
        tag = " + tmask" if masked else ""
        run("shallow water %dx%d one-step sweep, %s%s" % (nx, ny, hname, tag), DW.case_shallow_water,
            nx, ny, 7, A, fd.SW_FUSED, halo_mode=push, masked=masked)
        run("shallow water %dx%d two-step sweep, %s%s" % (nx, ny, hname, tag), DW.case_shallow_water,
            nx, ny, 8, A, fd.SW_FUSED, halo_mode=push, two_step=True, masked=masked)
    if push is not None:
        run("shallow water %dx%d one-step sweep, exchange path" % (nx, ny), DW.case_shallow_water,
            nx, ny, 5, A, fd.SW_FUSED)
    run("shallow water %dx%d two-step sweep, %s, contracted arithmetic (1e-12)" % (nx, ny, hname),
        DW.case_shallow_water, nx, ny, 8, A, fd.SW_FUSED, halo_mode=push, two_step=True, arith=fd.ARITH_CONTRACTED)
    run("shallow water %dx%d one-step sweep, %s, contracted arithmetic (1e-12)" % (nx, ny, hname),
        DW.case_shallow_water, nx, ny, 7, A, fd.SW_FUSED, halo_mode=push, arith=fd.ARITH_CONTRACTED)
    fd.dev_sync()
    nfail, ncheck = len(DW.FAILURES), DW.NCHECK
    if dist is not None:
        t = torch.tensor([nfail, ncheck], dtype=torch.int64, device="cuda")
        dist.all_reduce(t)
        nfail, ncheck = int(t[0].item()), int(t[1].item())
    for msg in DW.FAILURES[:10]:
        print("PARITY MISMATCH " + msg, file=sys.stderr, flush=True)
    # leave the library as the timed legs expect it
    fd.gocean_initialise(rank + 1, world)
    fd.set_arith(fd.ARITH_EXACT)
    os.environ["DL_ESM_ALIGNMENT"] = str(args.alignment)
    return {"checked": True, "ranks": world, "cases": len(cases), "checks": ncheck, "mismatches": nfail,
            "seconds": time.time() - t0, "halo_path": hname,
            "oracle": "serial CPU oracle (oracle/), itself pinned on the reference's known answers and on the "
                      "published shallow.f scheme (tests/test_oracle_golden.py, tests/test_oracle_sw.py)",
            "case_list": cases}


def run_ours(args):
    os.environ["DL_ESM_ALIGNMENT"] = str(args.alignment)
    rank, world, local_rank = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if world != args.gpus and world > 1:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, world))
    import numpy as np
    import torch

    import dl_esm_inf_b200 as fd
    from dl_esm_inf_b200 import _lib

    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    L = _lib.lib()
    fd.gocean_initialise(rank + 1, world, device=local_rank)
    if world > 1:
        # ship rank 1's NCCL id to everyone through torch.distributed (plumbing only)
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt = torch.frombuffer(bytearray(fd.comm_unique_id()), dtype=torch.uint8).cuda()
        dist.broadcast(idt, 0)
        fd.gocean_initialise(rank + 1, world, device=local_rank, comm_id=bytes(idt.cpu().numpy().tobytes()))

    def barrier():
        fd.dev_sync()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    def gather_ranks(obj):
        """every rank's value on every rank (small Python objects; outside the timed regions)"""
        if dist is None:
            return [obj]
        out = [None] * world
        dist.all_gather_object(out, obj)
        return out

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- parity first: a bench line of a wrong kernel is worth nothing ----
    parity = {"checked": False, "mismatches": None, "why": "--no-parity"}
    if not args.no_parity:
        parity = run_parity(args, fd, dist, torch, world, rank)
        if parity["mismatches"]:
            if rank == 0:
                print(json.dumps({"metric": METRIC, "value": None, "parity": parity,
                                  "error": "parity leg found mismatches: no number is reported"}), flush=True)
            return 1

    sampler = ClockSampler(local_rank)
    sampler.start()
    peak, peak_src = measured_hbm_peak()
    ext_base = fd.EXT_PERIODIC_DISTMEM if world > 1 else 0   # periodic wrap across ranks (SURVEY F4 extension)

    def timed(two_step, run=True, size=None, tmask=None, scaling=None, steps=None, sustain_s=None, arith=None,
              halo=None):
        """one timed run of K steps: the one-step sweep (96 B/point per step) or the two-step sweep
        (temporal blocking: 96 B/point per TWO steps, tiles with halo width 2)"""
        # weak scaling: size^2 points per GPU; strong: size^2 points in total (configs[3] names both).
        # The tile grid is go_decompose's own choice either way.
        n = size or args.size
        tmask = args.tmask if tmask is None else tmask
        scaling = scaling or args.scaling
        steps = steps or args.steps
        sustain_s = args.sustain_s if sustain_s is None else sustain_s
        fd.gocean_initialise(rank + 1, world)               # go_decompose's SAVEd halo width back to 1
        fd.set_arith(fd.ARITH_CONTRACTED if (arith or args.arith) == "fma" else fd.ARITH_EXACT)
        probe = fd.go_decompose(n, n)                       # tile grid for `world` ranks on a square domain
        if scaling == "strong":
            gx, gy = n, n
        else:
            gx, gy = n * probe.nx, n * probe.ny
        bcs = (fd.GO_BC_PERIODIC, fd.GO_BC_PERIODIC, fd.GO_BC_NONE)
        grid = fd.grid_type(fd.GO_ARAKAWA_C, bcs, fd.GO_OFFSET_SW)
        if two_step:
            fd.set_extensions(ext_base | (fd.EXT_HALO_DEPTH if world > 1 else 0))
            grid.decompose(gx, gy, halo_width=2)      # the two-step sweep needs two halo cells
        else:
            fd.set_extensions(ext_base)
            grid.decompose(gx, gy)
        tm = config3_mask_for(grid, gx, gy, np) if tmask else None
        fd.grid_init(grid, 1.0e5, 1.0e5, tm)
        sub = grid.subdomain
        local_points = sub.internal.nx * sub.internal.ny
        total_points = gx * gy

        sw = fd.shallow_water(grid, dt=90.0, alpha=0.001, mode=fd.SW_FUSED if args.mode == "fused" else fd.SW_UNFUSED,
                              use_tmask=bool(tmask))
        sw.set_variant(args.variant)
        if two_step:
            sw.set_two_step(True, args.variant2)
        sw.init_device()
        halo = halo or args.halo
        if halo == "auto":
            halo = "push" if (world > 1 and args.mode == "fused") else "exchange"
        if halo == "push":
            sw.set_halo_mode(fd.HALO_PUSH)              # collective: maps the neighbours' arrays (NVLink peer memory)
        if not run:                                      # set-up only (the e2e leg's one-step tile)
            return dict(sw=sw, grid=grid, total_points=total_points)
        mass0 = fd.field_checksum(sw.field("p"))

        def region(nsteps):
            """nsteps back-to-back model steps bracketed by barrier + sync; device time (CUDA events on the
            library's stream), max over ranks; the dominant kernel timed by its own event pairs"""
            L.dlesm_profile_enable(1)
            L.dlesm_profile_read(None, None, 1)
            # (the NVML counter reads take milliseconds and serialise across the ranks' processes: they stay
            # OUTSIDE the barriers -- taken between the barrier and the first event they started the ranks up
            # to 34 ms apart at N=4, which the max over ranks then charged to the steps)
            nvl0 = sampler.nvlink_kib() if world > 1 else None
            barrier()
            launches0 = fd.kernel_launch_count()
            w0 = time.time()
            L.dlesm_event_record(0)
            sw.step(nsteps)
            L.dlesm_event_record(1)
            ms = C.c_double(0.0)
            assert L.dlesm_event_elapsed_ms(0, 1, C.byref(ms)) == 0
            barrier()
            w1 = time.time()
            launches = fd.kernel_launch_count() - launches0
            kms, kn = C.c_double(0.0), C.c_longlong(0)
            L.dlesm_profile_read(C.byref(kms), C.byref(kn), 1)
            L.dlesm_profile_enable(0)
            el = max_over_ranks(ms.value)
            nvl1 = sampler.nvlink_kib() if world > 1 else None
            nvl = None
            if nvl0 and nvl1:
                # the counters tick at their own pace: this is a lower bound on the region's traffic unless
                # the region is long (the sustained leg); reported per step beside the algorithmic figure
                nvl = [(b - a) * 1024.0 / nsteps for a, b in zip(nvl0, nvl1)]
            clocks = sampler.summary(w0, w1)
            # what each rank saw: its own sweep time (with the in-sweep hand-shake a rank that waits for a
            # slower neighbour shows the wait here) and its own GPU's clocks -- a slow GPU explains itself
            per_rank = None
            if world > 1:
                per_rank = gather_ranks({"rank": rank + 1, "kernel_ms": kms.value / max(1, kn.value),
                                         "elapsed_ms_per_step": ms.value / nsteps, "sm_mhz": clocks.get("sm_mhz"),
                                         "reasons": clocks.get("reasons")})
            return dict(nvlink_bytes_per_step=nvl, elapsed_ms=el, launches=int(launches), kernel_ms_total=kms.value, kernel_launches=kn.value,
                        clocks=clocks, value=total_points * nsteps / (el * 1e-3), steps=nsteps, per_rank=per_rank)

        def roofline_of(r):
            """the dominant kernel (fused sweep) against the measured HBM copy bandwidth.  Bytes per launch =
            the DRAM minimum of one launch: 96 B x the points it sweeps (6 reads + 6 writes of FP64,
            SURVEY 8(d)) -- ONE pass over the state whether the launch advances one model step or two."""
            if r["kernel_launches"] <= 0:
                return None
            k_ms = r["kernel_ms_total"] / r["kernel_launches"]
            spl = r["steps"] / r["kernel_launches"]           # model steps per launch: 2 with temporal blocking
            bytes_launch = BYTES_PER_POINT * local_points + (local_points if tmask else 0)   # + 1 B/pt byte mask
            achieved = bytes_launch / (k_ms * 1e-3) / 1e9
            rf = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                  "traffic": None, "peak_source": peak_src,
                  "kernel": ("two-step fused shallow-water sweep fused2s_tma_kernel (variant2 %d)" % args.variant2)
                  if two_step else ("fused shallow-water sweep fused_tma_kernel (variant %d)" % args.variant),
                  "kernel_ms": k_ms, "kernel_share_of_step": r["kernel_ms_total"] / r["elapsed_ms"],
                  "algorithmic_bytes_per_launch": bytes_launch, "steps_per_launch": spl,
                  "frac_of_nominal_8TBs": achieved / 8000.0,
                  "effective_bytes_speedup": spl,
                  "effective_GBs_on_the_per_step_byte_model": achieved * spl,
                  "note": ("temporal blocking: one pass over the state (96 B/point) advances TWO model steps; "
                           "`frac` is that pass against the measured HBM copy bandwidth, "
                           "`effective_bytes_speedup` the factor the per-step byte model is beaten by")
                  if spl > 1.5 else "one pass over the state (96 B/point) per model step"}
            try:
                with open(os.path.join(ROOT, "profiles", "fused_traffic.json")) as fh:
                    tj = json.load(fh)
                key = ("dram_bytes_per_launch_two_step" if two_step else "dram_bytes_per_launch") + ("_masked" if tmask else "")
                if n == 16384 and world == 1:
                    rf["traffic"] = tj.get(key)
            except Exception:
                pass
            return rf

        sw.step(max(args.warmup, 3))
        R = region(steps)
        R["roofline"] = roofline_of(R)
        # parity guards on the timed state: mass conservation and finite checksums
        cks = sw.checksums()
        assert all(np.isfinite(c) for c in cks), cks
        assert abs(cks[0] - mass0) <= 1e-11 * abs(mass0), (mass0, cks[0])
        sustained = None
        if sustain_s > 0:
            ms_step = R["elapsed_ms"] / steps
            ns = max(2, int(sustain_s * 1e3 / ms_step))
            ns += ns & 1
            S = region(ns)
            rf = roofline_of(S)
            sustained = {"value": S["value"], "ms_per_step": S["elapsed_ms"] / ns, "steps": ns,
                         "seconds": S["elapsed_ms"] * 1e-3, "clocks": S["clocks"],
                         "roofline_frac": rf["frac"] if rf else None, "kernel_ms": rf["kernel_ms"] if rf else None}

        # ---- halo traffic of one step against NVLink (north_star: "halo traffic also as a fraction of
        # NVLink bandwidth"): the map_comms send/recv lists of this rank x the three new fields ----
        halo_traffic = None
        if world > 1:
            # one-step: depth-1 lists x 3 new fields every step; two-step: depth-2 lists x 6 fields
            # (u,v,p and the filtered old fields) every second step
            cm = fd.map_comms(grid.decomp, pbc=(True, True), halo_depths=(2, 2)) if two_step else grid.comms
            per_step = 8 * 6 / 2.0 if two_step else 8 * 3
            sent = sum(cm.nxsend[q] * cm.nysend[q] for q in range(cm.nsend)
                       if cm.destination[q] != rank) * per_step
            recv = sum(cm.nxrecv[q] * cm.nyrecv[q] for q in range(cm.nrecv)
                       if cm.source[q] != rank) * per_step
            step_s = R["elapsed_ms"] * 1e-3 / steps
            halo_traffic = {"sent_bytes_per_rank_per_step": sent, "recv_bytes_per_rank_per_step": recv,
                            "messages_per_rank_per_step": int(cm.nsend),
                            "nvlink_peak_GBs_both_directions": 1800.0,
                            "nvlink_frac_of_step": (sent + recv) / step_s / 1.8e12,
                            "nvml_nvlink_bytes_per_step_data_tx_rx_raw_tx_rx": (S if sustain_s > 0 else R).get(
                                "nvlink_bytes_per_step"),
                            "nvml_note": "NVML NVLINK_THROUGHPUT counters of this rank's GPU (all links) over the "
                                         "%s leg, per model step: data = payload, raw = with protocol overhead"
                                         % ("sustained" if sustain_s > 0 else "timed"),
                            "step_minus_sweep_us": (R["elapsed_ms"] - R["kernel_ms_total"]) / steps * 1e3,
                            "per_rank": R.get("per_rank")}

        R.update(cks=cks, halo_traffic=halo_traffic, halo=halo, grid=grid, sw=sw, probe=probe, sub=sub, gx=gx,
                 gy=gy, local_points=local_points, total_points=total_points, two_step=two_step,
                 sustained=sustained, n=n, tmask=bool(tmask), arith=(arith or args.arith))
        return R

    def halo_text(r):
        if r["two_step"] and r["halo"] == "push":
            return ("push: the two-step sweep stores the neighbours' halo cells of its six outputs, two cells "
                    "deep, over NVLink peer memory; flag hand-shake per sweep")
        if r["two_step"]:
            return ("periodic halo kernel, two cells deep, every second step" if world == 1 else
                    "halo_exchange(depth 2) of six fields every second step: pack -> NCCL send/recv -> unpack")
        if r["halo"] == "push":
            return ("push: the fused sweep stores the neighbours' halo cells over NVLink peer memory, "
                    "flag hand-shake per step")
        return "periodic halo kernel" if world == 1 else "pack -> NCCL send/recv -> unpack"

    def sub_record(r):
        return {"value": r["value"], "ms_per_step": r["elapsed_ms"] / r["steps"], "steps": r["steps"],
                "roofline": r["roofline"], "clocks": r["clocks"], "gpu_launches": r["launches"],
                "halo": halo_text(r), "halo_traffic": r["halo_traffic"], "checksums_p_u_v": list(r["cks"]),
                "sustained": r["sustained"], "arithmetic": r["arith"]}

    sweep = "two" if args.two_step else args.sweep
    if args.mode != "fused":
        sweep = "one"                                # the two-step sweep is a fused-mode path
    one = two = None
    if sweep in ("one", "both"):
        one = timed(False)
    main = one
    # ---- e2e: the same step driven with HOST buffers through the field hooks (one-step tile) ----
    e2e = None
    if args.e2e_steps > 0:
        src = one if one is not None else timed(False, run=False)
        e2e = run_e2e(args, fd, L, src["sw"], src["grid"], src["total_points"], dist, torch)
        if one is None:
            src["sw"].free()
        if sweep in ("two", "both"):
            # the headline e2e: TWO model steps per PCIe round trip (dlesm_sw_step_host_n, the bands go
            # through the two-step sweep); the one-step-per-trip figure stays beside it
            src2 = timed(True, run=False, halo="exchange")
            e2 = run_e2e(args, fd, L, src2["sw"], src2["grid"], src2["total_points"], dist, torch, per_trip=2)
            if world == 1 and not args.no_deep_trips:
                # for the record, not the headline: more model steps per round trip (skewed passes of the
                # two-step sweep over the row bands); the bus is crossed once per direction per CALL
                e2["more_steps_per_round_trip"] = [
                    run_e2e(args, fd, L, src2["sw"], src2["grid"], src2["total_points"], dist, torch, per_trip=k)
                    for k in (4, 8)]
            src2["sw"].free()
            for k in ("pcie_GBs_per_direction_both_busy_measured",):
                if k in e2e:
                    e2[k] = e2e[k]
            if e2.get("pcie_GBs_per_direction_both_busy_measured"):
                e2["frac_of_pcie"] = e2["pcie_GBs_per_direction_achieved"] / e2["pcie_GBs_per_direction_both_busy_measured"]
            e2["one_step_per_round_trip"] = e2e
            e2e = e2
    if one is not None:
        one["sw"].free()
        one["sw"] = None
    if sweep in ("two", "both"):
        two = timed(True)
        two["sw"].free()
        two["sw"] = None
        main = two
    # the other arithmetic of the headline sweep, for the record (same process, same box)
    other = None
    if args.both_arith and main is not None and args.mode == "fused":
        other = timed(main["two_step"], arith="exact" if args.arith == "fma" else "fma")
        other["sw"].free()
        other["sw"] = None

    # ---- BASELINE configs[2] (N=1): 4096^2, tmask land islands, DL_ESM_ALIGNMENT=64; full-size oracle check ----
    config3 = None
    if world == 1 and not args.no_config3 and args.mode == "fused":
        config3 = run_config3(args, fd, timed, sub_record, np)

    # ---- BASELINE configs[3] "(strong/weak)": the same --size^2 domain split over the N GPUs ----
    strong = None
    if world > 1 and args.scaling == "weak" and not args.no_strong and args.mode == "fused":
        strong = {}
        for label, two_step in (("two_step", True), ("one_step", False)):
            r = timed(two_step, scaling="strong", steps=max(args.steps, 40), sustain_s=0.0)
            r["sw"].free()
            r["sw"] = None
            strong[label] = sub_record(r)
            strong[label]["tile"] = "%dx%d per GPU, global %dx%d" % (r["sub"].internal.nx, r["sub"].internal.ny,
                                                                   r["gx"], r["gy"])
    # ---- BASELINE configs[4]: halo_exchange + field_checksum microbench on small and large local tiles ----
    config5 = None
    if not args.no_config5:
        config5 = run_config5(args, fd, L, C, world, rank, max_over_ranks, barrier)

    sampler.stop_flag = True

    # ---- cpu baseline (rank 0, N=1 only): bounded sample on the box's host cores; its checksums are the
    # full-size parity check of the GPU path ----
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        ncores = os.cpu_count() or 1
        ncpu = pick_cpu_sample(ncores, 6, 1, budget_s=20.0)
        cval, csecs, ccks, ccks_ld = cpu_run(ncpu, ncores, 6, 1)
        cpu = {"value": cval, "unit": UNIT, "cores": ncores, "kind": "port",
               "sample": "%dx%d double-periodic sample, 6 timed steps, %d threads (oracle port of the "
                         "reference's un-fused loops + exchange_generic-style halo exchange)" % (ncpu, ncpu, ncores),
               "checksums_p_u_v_after_7_steps": list(ccks), "checksums_p_u_v_long_double": list(ccks_ld)}
        cpu["serial_fd_lib_1core_1024sq"] = cpu_run(1024, 1, 6, 1)[0]
        # the GPU path from the same initial state, same number of steps (1 warm-up + 6 timed), same size
        full = {}
        for label, two_step in (("one_step_sweep", False), ("two_step_sweep", True)):
            r = timed(two_step, run=False, size=ncpu, tmask=False, scaling="weak", arith="exact")
            r["sw"].step(7)
            got = r["sw"].checksums()
            r["sw"].free()
            rel = [abs(a - b) / abs(b) for a, b in zip(got, ccks_ld)]
            full[label] = {"checksums_p_u_v": list(got), "max_rel_diff_vs_cpu_long_double": max(rel)}
            if max(rel) > 1e-12:
                parity["mismatches"] = (parity.get("mismatches") or 0) + 1
        parity["full_size"] = {"size": ncpu, "steps": 7, "tolerance": 1e-12,
                               "what": "GPU checksums (p, u, v) after 7 steps from the classic initial state vs the "
                                       "CPU arm's (long-double accumulation of the same sums) at %dx%d" % (ncpu, ncpu),
                               **full}

    rc = 0
    if rank == 0:
        R = main
        sub, probe, grid = R["sub"], R["probe"], R["grid"]
        line = {
            "metric": METRIC, "value": R["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": R["elapsed_ms"] / args.steps, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": arm_config(args, None if args.scaling == "weak" and args.mode == "fused" and not args.tmask else
                                 "shallow-water C-grid step %dx%d per GPU (%s scaling), double periodic, GO_OFFSET_SW, "
                                 "FP64, DL_ESM_ALIGNMENT=%d, mode=%s%s"
                                 % (sub.internal.nx, sub.internal.ny, args.scaling, args.alignment, args.mode,
                                    ", tmask with 64x64 land islands (1.6 % land)" if args.tmask else "")),
            "config_notes": {
                       "tile": "%dx%d per GPU, global %dx%d, tiles %dx%d" % (sub.internal.nx, sub.internal.ny, R["gx"],
                                                                          R["gy"], probe.nx, probe.ny),
                       "l2": "inputs larger than L2 (%d fields x %.2f GB per GPU)"
                             % (12 if R["two_step"] else 9, grid.nx * grid.ny * 8 / 1e9),
                       "fused_variant": args.variant,
                       "arithmetic": "contracted (FMA; within 1e-12 of the oracle)" if R["arith"] == "fma" else
                                     "exact (one rounding per operation; bit-identical to the oracle)",
                       "sweep": ("two model steps per sweep (temporal blocking, halo width 2, variant2 %d): the "
                                 "intermediate state never reaches HBM" % args.variant2)
                                if R["two_step"] else "one model step per sweep",
                       "parallelism": "tile-per-GPU %dx%d" % (probe.nx, probe.ny),
                       "halo": halo_text(R)},
            "clocks": R["clocks"], "gpu_launches": R["launches"], "roofline": R["roofline"], "cpu_baseline": cpu,
            "e2e": e2e, "halo_traffic": R["halo_traffic"], "checksums_p_u_v": list(R["cks"]),
            "sustained": R["sustained"], "parity": parity,
        }

        def pick(d, *path):
            for k in path:
                d = d.get(k) if isinstance(d, dict) else None
            return d
        # the figures a reviewer looks for first, at the top level at every N (the sub-records below carry the detail)
        line["summary"] = {
            "parity": {"ranks": parity.get("ranks"), "cases": parity.get("cases"), "checks": parity.get("checks"),
                       "mismatches": parity.get("mismatches"), "full_size_checked": "full_size" in parity},
            "sustained_value": pick(R, "sustained", "value"), "sustained_frac": pick(R, "sustained", "roofline_frac"),
            "burst_frac": pick(R, "roofline", "frac"),
            "one_step_value": one["value"] if (R["two_step"] and one is not None) else None,
            "one_step_frac": pick(one, "roofline", "frac") if (R["two_step"] and one is not None) else None,
            "strong_two_step": pick(strong, "two_step", "value"), "strong_one_step": pick(strong, "one_step", "value"),
            "strong_two_step_frac": pick(strong, "two_step", "roofline", "frac"),
            "config3_two_step_frac": pick(config3, "two_step", "roofline", "frac"),
            "config3_one_step_frac": pick(config3, "one_step", "roofline", "frac"),
            "config3_mismatches": pick(config3, "parity", "mismatches"),
            "e2e_value": pick(e2e, "value"), "e2e_frac_of_pcie": pick(e2e, "frac_of_pcie"),
            "stencil_parity_note": "stencils are compared with the project's oracle (pinned on the published shallow.f; "
                                   "the reference ships no stencil): vs oracle, unpinned by the reference",
        }
        if R["two_step"] and one is not None:
            # the one-step formulation (the reference's: every step reads and writes the state once),
            # measured in the same process just before
            line["one_step"] = sub_record(one)
        if other is not None:
            line["other_arithmetic"] = sub_record(other)
        if config3 is not None:
            line["config3"] = config3
        if strong is not None:
            line["strong"] = strong
        if config5 is not None:
            line["config5"] = config5
        print(json.dumps(line), flush=True)
        if parity.get("mismatches"):
            print("PARITY MISMATCH at full size: see parity.full_size / config3.parity", file=sys.stderr, flush=True)
            rc = 1
    if dist is not None:
        dist.barrier()
        fd.gocean_finalise()
        dist.destroy_process_group()
    return rc


def run_config5(args, fd, L, C, world, rank, max_over_ranks, barrier):
    """BASELINE configs[4]: "Halo-exchange + field_checksum microbench (width-1/2 halos, 2048^2-32768^2 local
    tiles, NE offset, external BCs -- the reference's dist-mem set-up)".  CUDA events on the library's stream,
    20 calls after 3 warm-up calls, max over ranks.  halo_exchange (N > 1) in both modes: direct (patches
    stored into the neighbours' receive slots over NVLink + flag hand-shake) and NCCL."""
    out = []
    reps = 20
    host_issue_us = [0.0]

    def timed_us(fn):
        for _ in range(3):
            fn()
        fd.dev_sync()
        barrier()
        L.dlesm_event_record(6)
        h0 = time.perf_counter()
        for _ in range(reps):
            fn()
        host_issue_us[0] = (time.perf_counter() - h0) / reps * 1e6   # host time to ISSUE a call (no sync)
        L.dlesm_event_record(7)
        ms = C.c_double(0.0)
        assert L.dlesm_event_elapsed_ms(6, 7, C.byref(ms)) == 0
        fd.dev_sync()
        return max_over_ranks(ms.value) / reps * 1e3

    bcs = (fd.GO_BC_EXTERNAL, fd.GO_BC_EXTERNAL, fd.GO_BC_NONE)
    for n in ((2048, 4096, 8192, 16384, 32768) if args.config5_full else (2048, 4096, 8192, 16384)):
        fd.gocean_initialise(rank + 1, world)
        probe = fd.go_decompose(n, n)
        for depth in ((1, 2) if world > 1 else (1,)):
            fd.gocean_initialise(rank + 1, world)              # go_decompose's SAVEd halo width back to 1
            fd.set_extensions(fd.EXT_HALO_DEPTH)
            g = fd.grid_type(fd.GO_ARAKAWA_C, bcs, fd.GO_OFFSET_NE)
            g.decompose(n * probe.nx, n * probe.ny, halo_width=depth)
            fd.grid_init(g, 1.0, 1.0)
            f = fd.r2d_field(g, fd.GO_T_POINTS)
            fd.set_field(f, 1.0)
            rec = {"local_tile": "%dx%d" % (n, n), "halo_width": depth}
            if depth == 1:
                i = f.internal
                us = timed_us(lambda: fd.field_checksum(f))
                rec["field_checksum_us"] = us
                rec["field_checksum_GBs"] = 8.0 * i.nx * i.ny / (us * 1e-6) / 1e9
                want = float(i.nx * i.ny) * world
                got = fd.field_checksum(f)
                assert abs(got - want) <= 1e-12 * want, (got, want)      # known answer: N ranks x points x 1.0
            if world > 1:
                c = g.comms if depth == 1 else fd.map_comms(g.decomp, halo_depths=(depth, depth))
                rec["halo_bytes_sent_rank1"] = sum(c.nxsend[q] * c.nysend[q] for q in range(c.nsend)) * 8
                for mode, name in ((fd.EXCHANGE_DIRECT, "direct"), (fd.EXCHANGE_NCCL, "nccl")):
                    fd.set_exchange_mode(mode)
                    rec["halo_exchange_%s_us" % name] = timed_us(lambda: f.halo_exchange(depth))
                    rec["halo_exchange_%s_host_issue_us" % name] = host_issue_us[0]
                fd.set_exchange_mode(fd.EXCHANGE_DIRECT)
            out.append(rec)
            f.free()
            fd.set_extensions(0)
    fd.gocean_initialise(rank + 1, world)
    return out


def run_config3(args, fd, timed, sub_record, np):
    """BASELINE configs[2]: "shallow-water step 4096^2 with tmask land points and DL_ESM_ALIGNMENT=64
    padding" -- both sweeps timed like the headline, and every prognostic field after 5 steps compared
    bit for bit with the serial oracle at full size."""
    from oracle import oracle as O
    n, out = 4096, {}
    for label, two_step in (("one_step", False), ("two_step", True)):
        r = timed(two_step, size=n, tmask=True, scaling="weak", steps=max(args.steps, 20),
                  sustain_s=min(args.sustain_s, 1.0))
        r["sw"].free()
        r["sw"] = None
        out[label] = sub_record(r)
    # oracle comparison at full size (outside any timed region)
    t0 = time.time()
    i = np.arange(1, n + 1)[:, None]
    j = np.arange(1, n + 1)[None, :]
    land = ((i % 512 >= 200) & (i % 512 <= 263)) & ((j % 512 >= 300) & (j % 512 <= 363))
    gmask = np.where(land, 0, 1).astype(np.int32)
    osw = O.ShallowWater(n, n, alignment=args.alignment, tmask=np.asfortranarray(np.pad(gmask, 1, mode="wrap")))
    oi = osw.fb["t"].internal
    # the oracle's own initial state (host libm sin/cos), so that the comparison is bit for bit
    core0 = {k: osw.f[k][oi.xstart - 1:oi.xstop, oi.ystart - 1:oi.ystop].copy() for k in ("u", "v", "p")}
    for _ in range(5):
        osw.step()
    oi = osw.fb["t"].internal
    osl = (slice(oi.xstart - 1, oi.xstop), slice(oi.ystart - 1, oi.ystop))
    par = {"size": n, "steps": 5, "fields": 6, "mismatches": 0, "what": "u,v,p,uold,vold,pold after 5 steps vs the "
           "serial oracle, bit for bit, 4096x4096 with the config-3 tmask"}
    for label, two_step in (("one_step", False), ("two_step", True)):
        r = timed(two_step, run=False, size=n, tmask=True, scaling="weak", arith="exact")
        sw = r["sw"]
        ii = sw.field("p").internal
        hw = ii.xstart - 1
        init = []
        for k in ("u", "v", "p"):
            a = np.zeros((r["grid"].nx, r["grid"].ny), order="F")
            a[0:n + 2 * hw, 0:n + 2 * hw] = np.pad(core0[k], hw, mode="wrap")
            init.append(a)
        sw.set_state(*init)
        sw.step(5)
        gsl = (slice(ii.xstart - 1, ii.xstop), slice(ii.ystart - 1, ii.ystop))
        bad = 0
        for name in ("u", "v", "p", "uold", "vold", "pold"):
            if not np.array_equal(sw.field(name).get_data()[gsl], osw.f[name][osl]):
                bad += 1
        par[label + "_fields_differing"] = bad
        par["mismatches"] += bad
        sw.free()
    par["seconds"] = time.time() - t0
    out["parity"] = par
    out["workload"] = "shallow-water C-grid step 4096x4096, tmask with 64x64 land islands (1.6 % land), DL_ESM_ALIGNMENT=64"
    return out


def run_e2e(args, fd, L, sw, grid, total_points, dist, torch, per_trip=1):
    """The step through the host-facing API with HOST buffers.  Every step the six prognostic
    fields (u,v,p,uold,vold,pold) go host -> device from pinned memory and the six results come
    back, all inside the timed region.  Two forms are timed:
      * pipelined (the headline `value`): dlesm_sw_step_host streams the tile through the device
        in row bands, upload / sweep / download overlapped on three streams;
      * hooks_serial: the reference's own calling pattern -- write_to_device_c x6, step,
        read_from_device_c x6 back to back on one stream (what fld%write_to_device /
        fld%read_from_device do, src/field_mod.f90:407-525)."""
    import numpy as np
    names = ("u", "v", "p", "uold", "vold", "pold")
    nx, ny = grid.nx, grid.ny
    nbytes = nx * ny * 8
    if per_trip >= 2 and sw.ncycle_hint == 0:
        sw.ncycle_hint = 3
        sw.step(3)                # the forward start-up step, then one pair: two-cell halos of all six fields valid
    flds = [sw.field(nm) for nm in names]
    host = []
    for f in flds:
        p = C.c_void_p()
        assert L.dlesm_host_alloc(nbytes, C.byref(p)) == 0, L.dlesm_last_error()
        a = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_double)), shape=(nx * ny,)).reshape((nx, ny), order="F")
        host.append((p, a))
        assert L.dlesm_field_read_from_device(f._h, a.ctypes.data_as(C.POINTER(C.c_double)), 1, 1, nx, ny, True) == 0
    ptrs = [p for p, _ in host]

    def step_hooks():
        fl = [sw.field(nm) for nm in names]          # handles follow the buffer rotation
        for f, (p, a) in zip(fl, host):
            L.dlesm_write_to_device_c(p, f.device_ptr, 1, 1, nx, ny, False)
        sw.step(1)
        fl = [sw.field(nm) for nm in names]
        for k, (f, (p, a)) in enumerate(zip(fl, host)):
            L.dlesm_read_from_device_c(f.device_ptr, p, 1, 1, nx, ny, k == len(fl) - 1)

    def step_pipelined():
        sw.step_host(ptrs, ptrs, band_rows=args.band_rows, nsteps=per_trip)

    def timed(fn, nsteps, slot):
        fn()                                          # warm-up
        fd.dev_sync()
        if dist is not None:
            dist.barrier()
        t0 = time.perf_counter()
        L.dlesm_event_record(slot)
        for _ in range(nsteps):
            fn()
        L.dlesm_event_record(slot + 1)
        ms = C.c_double(0.0)
        assert L.dlesm_event_elapsed_ms(slot, slot + 1, C.byref(ms)) == 0
        fd.dev_sync()
        # step_host returns only when the results are in host memory: the host clock and the
        # device events bracket the same work; report the larger
        el = max(ms.value, (time.perf_counter() - t0) * 1e3)
        if dist is not None:
            t = torch.tensor([el], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            el = float(t.item())
        return el

    trips = args.e2e_steps
    nmodel = trips * per_trip
    el = timed(step_pipelined, trips, 2)
    cks = [float(np.abs(a[1:-1, 1:-1]).sum()) for _, a in host[:3]] if nx * ny <= (1 << 22) else None
    out = {"value": total_points * nmodel / (el * 1e-3), "unit": UNIT,
           "h2d_bytes_per_step": 6 * nbytes // per_trip, "d2h_bytes_per_step": 6 * nbytes // per_trip,
           "steps": nmodel, "round_trips": trips, "model_steps_per_round_trip": per_trip,
           "ms_per_step": el / nmodel,
           "what": ("per round trip: dlesm_sw_step_host_n with HOST (pinned) buffers: u,v,p,uold,vold,pold "
                    "host->device, %s, the six results device->host, pipelined in row bands (upload / sweep / "
                    "download overlapped, both PCIe directions busy)"
                    % ("one model step" if per_trip == 1 else "TWO model steps in one pass of the two-step sweep"
                       if per_trip == 2 else "%d model steps in %d skewed passes of the two-step sweep"
                       % (per_trip, per_trip // 2)))}
    if per_trip == 1:
        el_hooks = timed(step_hooks, 1, 4)
        out["hooks_serial"] = {"value": total_points / (el_hooks * 1e-3), "ms_per_step": el_hooks,
                               "what": "write_to_device_c x6, step, read_from_device_c x6 back to back (the "
                                       "reference's hook calling pattern)"}
    if cks is not None:
        out["host_checksums_u_v_p"] = cks
    for p, _ in host:
        L.dlesm_host_free(p)
    # the bound of this path is the PCIe link, both directions at once: measure it the plain way
    # (two 1 GiB pinned buffers, one copy per direction on its own stream, CUDA events) and state
    # the step's traffic against it
    out["pcie_GBs_per_direction_achieved"] = 6 * nbytes / (el / trips * 1e-3) / 1e9
    if per_trip != 1:
        return out
    try:
        nb = 1 << 30
        h_up = torch.empty(nb, dtype=torch.uint8).pin_memory()
        h_dn = torch.empty(nb, dtype=torch.uint8).pin_memory()
        d_up = torch.empty(nb, dtype=torch.uint8, device="cuda")
        d_dn = torch.empty(nb, dtype=torch.uint8, device="cuda")
        s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()
        best = 0.0
        for it in range(4):
            torch.cuda.synchronize()
            e0, e1, e2, e3 = (torch.cuda.Event(enable_timing=True) for _ in range(4))
            with torch.cuda.stream(s_up):
                e0.record()
                d_up.copy_(h_up, non_blocking=True)
                e1.record()
            with torch.cuda.stream(s_dn):
                e2.record()
                h_dn.copy_(d_dn, non_blocking=True)
                e3.record()
            torch.cuda.synchronize()
            ms_both = max(e0.elapsed_time(e1), e2.elapsed_time(e3))
            if it > 0:
                best = max(best, nb / (ms_both * 1e-3) / 1e9)
        out["pcie_GBs_per_direction_both_busy_measured"] = best
        out["frac_of_pcie"] = out["pcie_GBs_per_direction_achieved"] / best if best > 0 else None
        del h_up, h_dn, d_up, d_dn
    except Exception as exc:          # the measurement is informational
        out["pcie_note"] = "bidirectional copy probe failed: %s" % exc
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--size", type=int, default=16384, help="points per GPU in each dimension")
    ap.add_argument("--alignment", type=int, default=64)
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--mode", default="fused", choices=["fused", "unfused"])
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--halo", default="auto", choices=["auto", "push", "exchange"],
                    help="halo fill of the fused step: push (in-kernel peer stores) or exchange (NCCL); auto = push when N>1")
    ap.add_argument("--band-rows", type=int, default=0, help="row-band height of the pipelined host step (0 = default)")
    ap.add_argument("--scaling", choices=["weak", "strong"], default="weak",
                    help="weak: --size^2 points per GPU (default, the metric's configuration); "
                         "strong: --size^2 points in total, split by go_decompose")
    ap.add_argument("--sweep", choices=["one", "two", "both"], default="both",
                    help="one: one model step per sweep (96 B/point/step); two: two steps per sweep (temporal "
                         "blocking); both (default): measure one, then two, report two as `value` and one "
                         "under `one_step`")
    ap.add_argument("--tmask", action="store_true",
                    help="config 3: pass a tmask with 64x64 land islands (~1.6 %% land) to grid_init")
    ap.add_argument("--two-step", action="store_true",
                    help="two model steps per sweep (temporal blocking; tile decomposed with halo_width=2)")
    ap.add_argument("--variant2", type=int, default=0, help="kernel variant of the two-step sweep")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the pre-flight parity leg (profiling runs)")
    ap.add_argument("--no-config3", action="store_true", help="skip the config-3 (4096^2 + tmask) sub-record")
    ap.add_argument("--no-deep-trips", action="store_true", help="skip the 4 / 8 steps-per-round-trip e2e sub-records")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling sub-record (N > 1)")
    ap.add_argument("--no-config5", action="store_true", help="skip the halo_exchange / field_checksum microbench")
    ap.add_argument("--config5-full", action="store_true", help="config 5 up to 32768^2 local tiles (8.6 GB per field)")
    ap.add_argument("--sustain-s", type=float, default=3.0,
                    help="seconds of back-to-back steps of the sustained leg (0 = none)")
    ap.add_argument("--arith", choices=["exact", "fma"], default="exact",
                    help="arithmetic of the fused sweeps: exact = one rounding per operation (bit-identical to the "
                         "oracle), fma = contracted (within 1e-12)")
    ap.add_argument("--both-arith", action="store_true", help="also time the headline sweep with the other arithmetic")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())

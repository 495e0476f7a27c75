"""Builds libdlesm_b200.so (hand-written sm_100a kernels + C++ host + C ABI) in-tree.

    python -m dl_esm_inf_b200.build [--force] [-v]

nvcc cross-compiles for sm_100a without a GPU.  Every translation unit is compiled to its own
object (in parallel; objects are cached under build/obj keyed by a hash of the source, the headers
and the flags) and the objects are linked into the shared library.

Arithmetic: the library is built with -fmad=false -- every FP64 operation rounds once, like the
reference's Fortran built without FMA and like the CPU oracle, so stencil results are
bit-identical.  dlesm_fused.cu is compiled a SECOND time with -fmad=true
(-DDLESM_CONTRACTED): the contracted-arithmetic twin of the fused sweeps, selected at run time
(dlesm_set_arith / DLESM_ARITH=fma) and held to north_star's 1e-12 relative tolerance instead of
bit equality.

The library carries a hash of everything it was built from ("DLESM_SRC_HASH=<hex>", also returned
by dlesm_source_hash()); build() compares it with the sources on disk and rebuilds on mismatch,
so a stale or foreign .so is never taken for the current code.
"""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libdlesm_b200.so")
OBJ_DIR = os.path.join(ROOT, "build", "obj")
HEADERS = ["dlesm_host.hpp", "dlesm_kernels.cuh", "dlesm_internal.hpp",
           os.path.join("..", "..", "include", "dlesm_b200.h")]

COMMON = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
          "-Xcompiler", "-fPIC,-Wall,-Wextra,-fvisibility=default"]
# (object stem, source, extra flags)
UNITS = [
    ("dlesm_host", "dlesm_host.cpp", ["-fmad=false"]),
    ("dlesm_kernels", "dlesm_kernels.cu", ["-fmad=false"]),
    ("dlesm_fused", "dlesm_fused.cu", ["-fmad=false"]),
    ("dlesm_fused_fma", "dlesm_fused.cu", ["-fmad=true", "-DDLESM_CONTRACTED=1"]),
    ("dlesm_runtime", "dlesm_runtime.cu", ["-fmad=false"]),
    ("dlesm_psy", "dlesm_psy.cu", ["-fmad=false"]),
    ("dlesm_static_scratch", "dlesm_static_scratch.cu", ["-fmad=false"]),
]
MARKER = b"DLESM_SRC_HASH="


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.sep not in cand or os.path.exists(cand)):
            return cand
    return "nvcc"


def _read(path: str) -> bytes:
    with open(path, "rb") as fh:
        return fh.read()


def _headers_blob() -> bytes:
    return b"".join(_read(os.path.join(CSRC, h)) for h in HEADERS)


def source_hash() -> str:
    """hash of every source, header and flag the library is built from"""
    h = hashlib.sha256()
    h.update(_headers_blob())
    for stem, src, extra in UNITS:
        h.update(stem.encode() + b"\0" + " ".join(COMMON + extra).encode() + b"\0")
        h.update(_read(os.path.join(CSRC, src)))
    h.update(_read(os.path.join(CSRC, "dlesm_version.cpp")))
    return h.hexdigest()[:32]


def embedded_hash(lib_path: str = LIB):
    """the DLESM_SRC_HASH the .so on disk was built with (None: no library / no marker)"""
    try:
        blob = _read(lib_path)
    except OSError:
        return None
    k = blob.find(MARKER)
    if k < 0:
        return None
    return blob[k + len(MARKER):k + len(MARKER) + 32].decode("ascii", "replace")


def needs_build() -> bool:
    return embedded_hash() != source_hash()


def _compile(stem, src, extra, hdr_blob, verbose):
    path = os.path.join(CSRC, src)
    flags = COMMON + extra
    key = hashlib.sha256(hdr_blob + _read(path) + " ".join(flags).encode()).hexdigest()[:16]
    obj = os.path.join(OBJ_DIR, "%s-%s.o" % (stem, key))
    if os.path.exists(obj):
        return obj, ""
    cmd = [_nvcc()] + flags + (["-Xptxas", "-v"] if verbose else []) + ["-c", path, "-o", obj + ".tmp"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("nvcc failed compiling %s (%s)" % (src, stem))
    os.replace(obj + ".tmp", obj)
    # drop older objects of the same unit
    for f in os.listdir(OBJ_DIR):
        if f.startswith(stem + "-") and f.endswith(".o") and os.path.join(OBJ_DIR, f) != obj:
            os.remove(os.path.join(OBJ_DIR, f))
    return obj, res.stdout + res.stderr


def build(force: bool = False, verbose: bool = False) -> str:
    want = source_hash()
    if not force and embedded_hash() == want:
        return LIB
    os.makedirs(OBJ_DIR, exist_ok=True)
    if force:
        for f in os.listdir(OBJ_DIR):
            if f.endswith(".o"):
                os.remove(os.path.join(OBJ_DIR, f))
    hdr = _headers_blob()
    objs, logs = [], []
    with cf.ThreadPoolExecutor(max_workers=min(len(UNITS), os.cpu_count() or 2)) as pool:
        futs = [pool.submit(_compile, stem, src, extra, hdr, verbose) for stem, src, extra in UNITS]
        for f in futs:
            obj, log = f.result()
            objs.append(obj)
            logs.append(log)
    # the version unit carries the hash: always recompiled (a second of g++)
    vobj = os.path.join(OBJ_DIR, "dlesm_version.o")
    cmd = [_nvcc()] + COMMON + ["-DDLESM_SOURCE_HASH=\"%s\"" % want, "-c",
                                os.path.join(CSRC, "dlesm_version.cpp"), "-o", vobj]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("nvcc failed compiling dlesm_version.cpp")
    cmd = [_nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", LIB + ".tmp"] + objs + \
          [vobj, "-lnccl", "-ldl"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("nvcc failed linking libdlesm_b200.so")
    os.replace(LIB + ".tmp", LIB)
    if verbose:
        print("\n".join(logs))
    got = embedded_hash()
    if got != want:
        raise RuntimeError("libdlesm_b200.so does not carry the expected source hash (%s vs %s)" % (got, want))
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))

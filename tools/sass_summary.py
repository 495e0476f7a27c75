#!/usr/bin/env python
"""Per-kernel SASS instruction histogram of libdlesm_b200.so (cuobjdump -sass), keyed by the library's
source hash: the committed proof of what the sm_100a build contains (TMA loads UTMALDG, mbarrier SYNCS,
FP64 pipe DADD/DMUL/DFMA, stores STG, shared loads LDS ...).

    python tools/sass_summary.py            # writes profiles/sass_summary_<hash>.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dl_esm_inf_b200 import build as B  # noqa: E402

WATCH = ["UTMALDG", "UTMASTG", "UTMAPF", "SYNCS", "DADD", "DMUL", "DFMA", "MUFU", "LDG", "STG", "LDS", "STS", "SHFL",
         "BAR", "ATOM*", "RED*", "UTC*", "HMMA", "DMMA", "LDL", "STL"]


def count(cnt, w):
    """occurrences of mnemonic w; a trailing * matches every mnemonic with that prefix (ATOMG, ATOMS ...)"""
    if w.endswith("*"):
        return sum(v for k, v in cnt.items() if k.startswith(w[:-1]))
    return cnt[w]


def main():
    lib = B.build()
    h = B.embedded_hash(lib)
    out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    kernels, cur, arch = collections.OrderedDict(), None, set()
    for ln in out.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            cur = kernels.setdefault(m.group(1), collections.Counter())
            continue
        m = re.search(r"arch = (sm_\w+)", ln)
        if m:
            arch.add(m.group(1))
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", ln)
        if m and cur is not None:
            cur[m.group(1)] += 1
    demangle = subprocess.run(["c++filt"], input="\n".join(kernels), capture_output=True, text=True).stdout.splitlines()
    path = os.path.join(ROOT, "profiles", "sass_summary_%s.txt" % h)
    tot = collections.Counter()
    with open(path, "w") as fh:
        fh.write("# cuobjdump -sass dl_esm_inf_b200/libdlesm_b200.so   source hash %s   arch %s   %d kernels\n" %
                 (h, ",".join(sorted(arch)), len(kernels)))
        fh.write("# columns: total instructions, then the watched mnemonics that occur (opcode root before the first '.')\n")
        for (name, cnt), pretty in zip(kernels.items(), demangle):
            short = re.sub(r"\(anonymous namespace\)::", "", pretty)
            short = re.sub(r"\(.*", "", short)[:110]
            n = sum(cnt.values())
            cells = " ".join("%s=%d" % (w, count(cnt, w)) for w in WATCH if count(cnt, w))
            fh.write("%-112s n=%-6d %s\n" % (short, n, cells))
            tot.update(cnt)
        fh.write("# whole library: n=%d %s\n" % (sum(tot.values()), " ".join("%s=%d" % (w, count(tot, w)) for w in WATCH)))
    print(path)


if __name__ == "__main__":
    main()

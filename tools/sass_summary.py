#!/usr/bin/env python
"""Per-kernel SASS mnemonic counts of the built libgrb200.so -> profiles/rN_sass_summary.txt.

    python tools/sass_summary.py [gnuradio_b200/libgrb200.so] > profiles/r2_sass_summary.txt

The .so is git-ignored (history stays source-only), so this listing is the tracked evidence that the kernels are
Blackwell-native: UTCHMMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st (tensor memory), UBLKCP = cp.async.bulk (1-D TMA),
UTMALDG/UTMASTG = tensor-map TMA, SYNCS = mbarrier, UTCBAR = tcgen05.commit, FFMA2/FADD2/FMUL2 = packed fp32.
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WANT = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTCBAR", "UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "FFMA2", "FADD2", "FMUL2",
        "FFMA", "LDS", "STS", "LDG", "STG", "SHFL", "BAR", "MUFU"]


def main():
    lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gnuradio_b200", "libgrb200.so")
    txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    dem = {}
    fn, counts, order = None, {}, []
    for line in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            fn = m.group(1)
            counts[fn] = collections.Counter()
            order.append(fn)
            continue
        if fn is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
        if m:
            op = m.group(1)
            counts[fn][op] += 1
            counts[fn]["_total"] += 1
    names = subprocess.run(["c++filt"], input="\n".join(order), capture_output=True, text=True).stdout.splitlines()
    for a, b in zip(order, names):
        dem[a] = re.sub(r"\(.*$", "", b)
    arch = re.findall(r"arch = (sm_\w+)", txt)
    print("# %s: %d kernels, arch %s" % (os.path.relpath(lib, ROOT), len(order), sorted(set(arch))))
    print("# %-64s %7s  %s" % ("kernel", "instrs", "  ".join(WANT)))
    for fn in sorted(order, key=lambda f: dem[f]):
        c = counts[fn]
        cols = "  ".join("%*d" % (len(w), c[w]) for w in WANT)
        print("%-66s %7d  %s" % (dem[fn][:66], c["_total"], cols))
    tc = [dem[f] for f in order if counts[f]["UTCHMMA"] or counts[f]["LDTM"] or counts[f]["STTM"]]
    tma = [dem[f] for f in order if counts[f]["UBLKCP"] or counts[f]["UTMALDG"]]
    print("\n# kernels using tensor memory / tcgen05 (%d): %s" % (len(tc), ", ".join(sorted(set(tc)))))
    print("# kernels using TMA bulk copies (%d): %s" % (len(tma), ", ".join(sorted(set(tma)))))


if __name__ == "__main__":
    main()

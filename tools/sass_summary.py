#!/usr/bin/env python
"""Per-kernel SASS instruction mix of libm3r_b200.so (cuobjdump -sass), written as markdown.
Usage: python tools/sass_summary.py > profiles/<round>_sass_summary.md"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "lightweight-mast3r-slam_b200", "libm3r_b200.so")
KEY = ["UBLKCP", "SYNCS", "LDGSTS", "LDG", "LDS", "STS", "STG", "LD", "ST", "ATOM", "ATOMG", "ATOMS", "RED", "SHFL", "BAR",
       "HMUL2", "HADD2", "HFMA2", "PRMT", "FFMA", "FMUL", "FADD", "MUFU", "DFMA", "DADD", "DMUL", "IMAD", "LOP3",
       "CCTL", "ERRBAR", "MEMBAR", "VOTE", "REDUX", "UTMALDG", "UTCMMA"]


def demangle(name):
    try:
        return subprocess.check_output(["c++filt", name], text=True).strip()
    except Exception:
        return name


def main():
    out = subprocess.check_output(["cuobjdump", "-sass", SO], text=True, errors="replace")
    kernels, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = demangle(m.group(1))
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_]+)((?:\.[A-Z0-9_]+)*)", line)
        if m and cur:
            op, mods = m.group(2), m.group(3)
            kernels[cur][op] += 1
            kernels[cur]["_total"] += 1
            if op in ("LDS", "LDG", "STG", "STS", "LD", "ST") and ".128" in mods:
                kernels[cur][op + ".128"] += 1
            if op == "MUFU":
                kernels[cur]["MUFU" + mods] += 1
    print(f"# SASS instruction mix per kernel (`cuobjdump -sass {os.path.relpath(SO, ROOT)}`, sm_100a)\n")
    print("Static counts (instructions in the binary, not executed counts).  What to look for: `UBLKCP` = TMA bulk copy "
          "(cp.async.bulk), `SYNCS` = mbarrier, `LDGSTS` = cp.async, `HMUL2`/`HADD2` without `HFMA2` = per-op rounded "
          "half arithmetic (bit-exactness contract of refine_matches), `DFMA` = fp64 solve, `RED`/`ATOM` = atomics.\n")
    for name, c in kernels.items():
        short = re.sub(r"\(.*", "", name)
        tot = c["_total"]
        parts = [f"{k} {c[k]}" for k in KEY if c[k]]
        extra = [f"{k} {v}" for k, v in sorted(c.items()) if (k.endswith(".128") or k.startswith("MUFU.")) and v]
        print(f"* `{short}` — {tot} instructions: " + ", ".join(parts + extra))
    return 0


if __name__ == "__main__":
    sys.exit(main())

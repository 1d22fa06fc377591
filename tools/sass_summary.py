#!/usr/bin/env python
"""SASS evidence for libnconv_b200.so, small enough to commit: per kernel the register / shared-memory use
(cuobjdump -res-usage) and a histogram of the instruction mnemonics that matter for the design claims —
packed fp32 FMA (FFMA2), 128-bit global accesses (LDG.E.128 / STG.E.128), bulk / tensor-map copies (UBLKCP,
UTMALDG), tcgen05 (UTCMMA / UTCHMMA, LDTM, UTCBAR), MUFU.RCP epilogues, shuffles, local-memory spills.

  python tools/sass_summary.py [--so path] [--match regex] > profiles/sass_summary_rNN.txt
"""
from __future__ import annotations

import argparse
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "realtime-depth-estimation-nconv_b200", "csrc", "libnconv_b200.so")

KEYS = [("FFMA2", r"^FFMA2"), ("FFMA", r"^FFMA(\.|$)"), ("FMUL2/FADD2", r"^(FMUL2|FADD2)"), ("MUFU.RCP", r"^MUFU\.RCP"),
        ("LDG.128", r"^LDG\.E(\.\w+)*\.128"), ("LDG", r"^LDG"), ("STG.128", r"^STG\.E(\.\w+)*\.128"), ("STG", r"^STG"),
        ("LDS", r"^LDS"), ("STS", r"^STS"), ("SHFL", r"^SHFL"), ("UBLKCP", r"^UBLKCP"), ("UTMALDG", r"^UTMALDG"),
        ("UTCMMA", r"^UTC[A-Z]*MMA"), ("LDTM", r"^LDTM"), ("UTCBAR", r"^UTCBAR"), ("SYNCS", r"^SYNCS"),
        ("BAR", r"^BAR"), ("LDL/STL", r"^(LDL|STL)"), ("ATOM/RED", r"^(ATOM|RED|ATOMS|ATOMG)"), ("DADD/DFMA", r"^(DADD|DFMA|DMUL)")]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def short(name: str) -> str:
    name = re.sub(r"\(.*$", "", name)
    name = name.replace("nconv::", "").replace("void ", "")
    return name if len(name) <= 110 else name[:107] + "..."


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--so", default=SO)
    ap.add_argument("--match", default=".")
    args = ap.parse_args()
    res = subprocess.run(["cuobjdump", "-res-usage", args.so], capture_output=True, text=True).stdout
    usage = {}
    cur = None
    for line in res.splitlines():
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            cur = m.group(1)
            continue
        if cur and "REG:" in line:
            g = dict(re.findall(r"(\w+):(\d+)", line))
            usage[cur] = (int(g.get("REG", 0)), int(g.get("SHARED", 0)), int(g.get("LOCAL", 0)), int(g.get("STACK", 0)))
            cur = None
    sass = subprocess.run(["cuobjdump", "-sass", args.so], capture_output=True, text=True).stdout
    hist = collections.OrderedDict()
    cur = None
    pats = [(k, re.compile(p)) for k, p in KEYS]
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            hist[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        hist[cur]["total"] += 1
        for k, p in pats:
            if p.match(op):
                hist[cur][k] += 1
    names = demangle(list(hist))
    rx = re.compile(args.match)
    print("# %s  (cuobjdump -res-usage / -sass, sm_100a)" % os.path.relpath(args.so, ROOT))
    print("# columns: instructions | REG SHARED(static) LOCAL STACK | non-zero mnemonic counts")
    for fn in sorted(hist, key=lambda f: names[f]):
        nm = short(names[fn])
        if not rx.search(nm):
            continue
        r = usage.get(fn, (0, 0, 0, 0))
        h = hist[fn]
        cols = " ".join("%s=%d" % (k, h[k]) for k, _ in KEYS if h[k])
        print("%s\n    %d instr | REG %d SHARED %d LOCAL %d STACK %d | %s" % (nm, h["total"], r[0], r[1], r[2], r[3], cols))


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""SASS listing excerpt of a CUDA-core kernel's inner FMA loop: the 120-instruction window with the most FFMA2 / FFMA,
encodings stripped.   python tools/sass_ffma_loop.py <mangled-name-substring> > profiles/sass_<kernel>_fma_loop_rNN.txt"""
import os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "realtime-depth-estimation-nconv_b200", "csrc", "libnconv_b200.so")
key = sys.argv[1]
txt = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout.split("\n")
out, on, name = [], False, None
for ln in txt:
    m = re.search(r"Function : (\S+)", ln)
    if m:
        on = key in m.group(1) and (name is None or m.group(1) == name)
        if on:
            name = m.group(1)
        continue
    if on and re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
        out.append(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", ln).rstrip())
is_fma = [1 if re.search(r"\bFFMA2?\b", l) else 0 for l in out]
W = 120
best, lo = -1, 0
run = sum(is_fma[:W])
for i in range(0, max(1, len(out) - W)):
    if run > best:
        best, lo = run, i
    run += (is_fma[i + W] if i + W < len(out) else 0) - is_fma[i]
dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
n2 = sum(1 for l in out if re.search(r"\bFFMA2\b", l)); n1 = sum(is_fma) - n2
print("# %s\n# %s\n# instructions %d..%d of %d (densest FMA window: %d of %d are FFMA2/FFMA); whole kernel: %d FFMA2, %d FFMA"
      % (dem, name, lo, lo + W, len(out), best, W, n2, n1))
print("\n".join(out[lo:lo + W]))

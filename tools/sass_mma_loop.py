#!/usr/bin/env python
"""SASS listing excerpt of a tcgen05 kernel's MMA issue loop (first to last UTCHMMA, with context), encodings stripped.
  python tools/sass_mma_loop.py <mangled-name-substring> > profiles/sass_<kernel>_mma_loop_rNN.txt"""
import os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "realtime-depth-estimation-nconv_b200", "csrc", "libnconv_b200.so")
key = sys.argv[1]
txt = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout.split("\n")
out, on, name = [], False, None
for ln in txt:
    m = re.search(r"Function : (\S+)", ln)
    if m:
        on = key in m.group(1)
        if on and name is None:
            name = m.group(1)
        elif on and m.group(1) != name:
            on = False
        continue
    if on and re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
        out.append(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", ln).rstrip())
idx = [i for i, l in enumerate(out) if "UTCHMMA" in l]
lo, hi = max(0, idx[0] - 30), min(len(out), idx[-1] + 15)
dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
print("# %s\n# %s\n# instructions %d..%d of %d (first to last UTCHMMA with context); %d UTCHMMA in the excerpt" % (dem, name, lo, hi, len(out), len(idx)))
print("\n".join(out[lo:hi]))

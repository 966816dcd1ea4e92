#!/usr/bin/env python
"""SASS instruction histogram per kernel of the built library (code-size evidence).

    python tools/sass_hist.py [maniac_mc_b200/libmaniacgpu.so] [regex of kernel names] > profiles/<tag>_sass_hist.txt
"""
import collections
import re
import subprocess
import sys

so = sys.argv[1] if len(sys.argv) > 1 else "maniac_mc_b200/libmaniacgpu.so"
pat = re.compile(sys.argv[2] if len(sys.argv) > 2 else "k_chain_run|k_sweep|k_strict")
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
cur, hist = None, {}
for ln in txt.splitlines():
    m = re.match(r"\s*Function : (\S+)", ln)
    if m:
        cur = m.group(1); hist[cur] = collections.Counter(); continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", ln)
    if m and cur:
        hist[cur][m.group(1).split(".")[0] + ("." + m.group(1).split(".")[1] if m.group(1).startswith(("LDS", "STS", "LDG", "STG", "LDL", "STL")) and "." in m.group(1) else "")] += 1
names = subprocess.run(["cu++filt"] + list(hist), capture_output=True, text=True).stdout.splitlines()
print(f"# {so}: SASS instruction counts per kernel (cuobjdump -sass); bytes = 16 x instructions")
for mangled, name in zip(hist, names):
    if not pat.search(name):
        continue
    h = hist[mangled]; n = sum(h.values())
    cls = lambda *p: sum(v for k, v in h.items() if k.startswith(p))
    print(f"\n== {name[:110]}\n   {n} instructions ({16 * n / 1024:.0f} kB): DFMA {cls('DFMA')} DMUL {cls('DMUL')} DADD {cls('DADD')} DSETP {cls('DSETP')} "
          f"MUFU {cls('MUFU')} LDS {cls('LDS')} STS {cls('STS')} LDG {cls('LDG')} STG {cls('STG')} LDL {cls('LDL')} STL {cls('STL')} "
          f"BAR {cls('BAR')} BRA {cls('BRA')} CALL {cls('CALL')} SHFL {cls('SHFL')} LDC {cls('LDC', 'ULDC')}")
    print("   top: " + ", ".join(f"{k} {v}" for k, v in h.most_common(14)))

#!/usr/bin/env python
"""SASS instruction count per source line / region of one kernel (static code size, not a profile).
Usage: sass_lines.py lib.so kernel-substring [file:lo-hi:name,...]"""
import collections, os, re, subprocess, sys, tempfile
so, pat = sys.argv[1:3]
td = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=td, check=True, stdout=subprocess.DEVNULL)
cnt = collections.Counter()
for f in os.listdir(td):
    dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(td, f)], stdout=subprocess.PIPE, text=True).stdout
    on, cur = False, None
    for ln in dis.splitlines():
        if ln.startswith("//-----") and ".text." in ln:
            on = pat in ln
        if not on:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
        if re.match(r"\s+/\*[0-9a-f]+\*/\s+\S", ln):
            cnt[cur] += 1
tot = sum(cnt.values())
print("total", tot)
if len(sys.argv) > 3:
    for spec in sys.argv[3].split(","):
        f, rng, name = spec.split(":"); lo, hi = map(int, rng.split("-"))
        print(f"{name:20s} {sum(v for k, v in cnt.items() if k and k[0] == f and lo <= k[1] <= hi):6d}")
else:
    for k, v in cnt.most_common(40): print(k, v)

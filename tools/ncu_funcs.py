#!/usr/bin/env python
"""Executed warp-instructions per device FUNCTION of one kernel (ncu --page source joined with nvdisasm line
info and the function definitions found in the source headers).
Usage: ncu_funcs.py report.ncu-rep lib.so kernel-substring game_steps_per_launch"""
import collections, csv, os, re, subprocess, sys, tempfile
rep, so, pat = sys.argv[1:4]
ws = float(sys.argv[4]) / 32 if len(sys.argv) > 4 else 1.0
td = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=td, check=True, stdout=subprocess.DEVNULL)
line_of = {}
for f in os.listdir(td):
    dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(td, f)], stdout=subprocess.PIPE, text=True).stdout
    if pat not in dis:
        continue
    on, cur = False, None
    for ln in dis.splitlines():
        if ln.startswith("//-----") and ".text." in ln:
            on = pat in ln
        if not on:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (m.group(1), int(m.group(2))); continue
        m = re.match(r"\s+/\*([0-9a-f]+)\*/\s+(\S.*);", ln)
        if m:
            line_of[int(m.group(1), 16)] = cur
marks = {}
def fn_of(k):
    if k is None:
        return "?"
    path, l = k
    base = os.path.basename(path)
    if base not in marks:
        ms = []
        try:
            for n, ln in enumerate(open(path), 1):
                m = re.match(r"\s*(static |inline |__host__ __device__ inline )*(RGS_DEV[_A-Z]*|__global__)[\w<>:, \*&\(\)]*?\s\**&?(\w+)\(", ln)
                if m:
                    ms.append((n, m.group(3)))
        except OSError:
            pass
        marks[base] = ms
    name = base
    for n, nm in marks[base]:
        if n <= l:
            name = nm
        else:
            break
    return name
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h = rows[1]
ia, ii, it, isamp = h.index("Address"), h.index("Instructions Executed"), h.index("Thread Instructions Executed"), h.index("# Samples")
per = collections.defaultdict(lambda: [0, 0, 0, 0])
base = None
for r in rows[2:]:
    if len(r) <= it or not r[ia].startswith("0x"):
        continue
    a = int(r[ia], 16)
    if base is None:
        base = a
    k = fn_of(line_of.get(a - base))
    per[k][0] += int(r[ii]); per[k][1] += int(r[it]); per[k][2] += 1; per[k][3] += int(r[isamp])
tot = sum(v[0] for v in per.values()); tots = sum(v[3] for v in per.values())
print(f"total executed warp-instructions {tot:.4g} = {tot / ws:.1f} per warp-step; avg lanes {sum(v[1] for v in per.values()) / tot:.2f}")
print(f"{'function':20s} {'inst/warp-step':>14s} {'inst%':>6s} {'samp%':>6s} {'lanes':>6s} {'sass':>5s}")
for k, v in sorted(per.items(), key=lambda kv: -kv[1][0]):
    if v[0]:
        print(f"{k:20s} {v[0] / ws:14.1f} {100 * v[0] / tot:6.1f} {100 * v[3] / max(1, tots):6.1f} {v[1] / v[0]:6.1f} {v[2]:5d}")

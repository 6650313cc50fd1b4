#!/usr/bin/env python
"""Per-source-line instruction / stall-sample profile of one kernel.

Joins `ncu --page source --csv` (per-SASS-address counters) with `nvdisasm -g -c` line info of the
cubin extracted from the shipped .so (same build), because ncu's CUDA-source view does not import
header files.  Usage: ncu_lines.py <report.ncu-rep> <lib.so> <kernel-mangled-substring> [top]
"""
import csv, re, subprocess, sys, tempfile, os, collections

rep, so, pat = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
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
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]+)\*/\s+(\S.*);", ln)
        if m:
            line_of[int(m.group(1), 16)] = cur
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
ia, ii, isamp, inoinst = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("stall_no_inst")
base = None
inst = collections.Counter(); samp = collections.Counter(); noinst = collections.Counter(); nsass = collections.Counter()
for r in rows[2:]:
    if len(r) <= ii or not r[ia].startswith("0x"):
        continue
    a = int(r[ia], 16)
    if base is None:
        base = a
    key = line_of.get(a - base)
    inst[key] += int(r[ii]); samp[key] += int(r[isamp]); noinst[key] += int(r[inoinst]); nsass[key] += 1
ti, ts = sum(inst.values()), sum(samp.values())
print(f"total inst {ti:.4g} samples {ts} sass {sum(nsass.values())}")
print(f"{'file:line':28s} {'inst%':>7s} {'samp%':>7s} {'noinst%':>8s} {'sass':>5s}")
for k, v in sorted(inst.items(), key=lambda kv: -samp[kv[0]])[:top]:
    print(f"{str(k[0])+':'+str(k[1]) if k else '?':28s} {100*v/ti:7.2f} {100*samp[k]/ts:7.2f} {100*noinst[k]/max(1,samp[k]):8.1f} {nsass[k]:5d}")
if os.environ.get("RANGES"):
    # RANGES="file:lo-hi:name,..." aggregates inst% per region
    print("---- regions")
    for spec in os.environ["RANGES"].split(","):
        f, rng, name = spec.split(":")
        lo, hi = map(int, rng.split("-"))
        ks = [k for k in inst if k and k[0] == f and lo <= k[1] <= hi]
        print(f"{name:24s} inst% {100*sum(inst[k] for k in ks)/ti:6.2f}  samp% {100*sum(samp[k] for k in ks)/ts:6.2f}  sass {sum(nsass[k] for k in ks)}")

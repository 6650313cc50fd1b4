#!/usr/bin/env python
"""tools/sass_funcs.py obj.o kernel-substring : static SASS instruction count per source FUNCTION of one kernel
(nvdisasm -g line info joined with the function definitions found in the source headers)."""
import collections, os, re, subprocess, sys
obj, pat = sys.argv[1:3]
dis = subprocess.run(["nvdisasm", "-g", "-c", obj], stdout=subprocess.PIPE, text=True).stdout
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
                m = re.match(r"\s*(static |inline |template <[^>]*>\s*|__host__ __device__ inline )*(RGS_DEV[_A-Z]*|__global__)[\w<>:, \*&\(\)]*?\s\**&?(\w+)\(", ln)
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
on, cur = False, None
cnt = collections.Counter()
for ln in dis.splitlines():
    if ln.startswith("//-----") and ".text." in ln:
        on = pat in ln
    if not on:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1), int(m.group(2))); continue
    if re.match(r"\s+/\*[0-9a-f]+\*/\s+\S.*;", ln):
        cnt[fn_of(cur)] += 1
tot = sum(cnt.values())
print("total", tot)
for k, v in cnt.most_common(45):
    print(f"{k:24s} {v:5d}")

#!/usr/bin/env python
"""Aggregate an `ncu --page source --csv` SASS dump by CUDA source line, using `nvdisasm -g` line
info of the same cubin (the CLI's source page has no per-line metrics).

usage: ncu_by_line.py <source.csv> <nvdisasm -gi -c output> <mangled kernel name substring> [top N]
Also prints a per-function table: every instruction is attributed to the innermost function of OUR
sources on its inline chain (function extents are parsed from csrc/*.cuh).
"""
import glob
import os
import csv
import re
import sys
from collections import defaultdict

src_csv, sass, kernel = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40

# --- nvdisasm: address -> (file, line) for the kernel's section
lines = open(sass).read().split("\n")
start = next(i for i, l in enumerate(lines) if l.startswith(".text.") and kernel in l)
addr_line = {}
chain, fresh = [], True
addr_chain = {}
for l in lines[start + 1:]:
    if l.startswith(".text.") or l.startswith("//--------------------- "):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        if fresh:
            chain, fresh = [], False
        chain.append((m.group(1).split("/")[-1], int(m.group(2))))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        fresh = True
        a = int(m.group(1), 16)
        addr_line[a] = (chain[0] if chain else ("?", 0), m.group(2).strip())
        addr_chain[a] = list(chain)

# function extents of our sources
funcs = {}
here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "ray-tracer_b200", "csrc")
for f in glob.glob(os.path.join(here, "*.cuh")) + glob.glob(os.path.join(here, "*.cu")):
    starts = []
    for i, l in enumerate(open(f), 1):
        m = re.match(r"^(?:template <[^>]*>\s*)?(?:static\s+)?(?:__device__|__global__)[^(]*?\b([A-Za-z_0-9]+)\s*\(", l)
        if m and not l.startswith(" "):
            starts.append((i, m.group(1)))
        elif l.startswith("__global__") or ("__global__ void" in l and not l.startswith(" ")):
            m2 = re.search(r"\b([A-Za-z_0-9]+)\s*\(", l.split("__global__")[-1].replace("__launch_bounds__(RTB_BLOCK_THREADS, RTB_MIN_BLOCKS)", ""))
            if m2:
                starts.append((i, m2.group(1)))
    funcs[os.path.basename(f)] = sorted(set(starts))


def func_of(file, line):
    best = None
    for st, name in funcs.get(file, []):
        if st <= line:
            best = name
    return best

rows = list(csv.reader(open(src_csv)))
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hdr_i]
col = {n: i for i, n in enumerate(hdr)}
body = []
for r in rows[hdr_i + 1:]:  # a filtered import prints the listing twice: stop at the second header
    if not r or not r[0].startswith("0x"):
        break
    body.append(r)
base = int(body[0][0], 16)
by_line = defaultdict(lambda: [0, 0, 0])
by_func = defaultdict(lambda: [0, 0, 0])
tot = [0, 0, 0]
missing = 0
for r in body:
    a = int(r[0], 16) - base
    inst, thr, samp = int(r[col["Instructions Executed"]]), int(r[col["Thread Instructions Executed"]]), int(r[col["# Samples"]])
    key = addr_line.get(a, (("?", 0), ""))[0]
    if a not in addr_line:
        missing += 1
    e = by_line[key]
    e[0] += inst; e[1] += thr; e[2] += samp
    names = [func_of(f, ln) for f, ln in addr_chain.get(a, [])]
    names = [n for n in names if n]
    small = {"v3", "dot", "cross", "normalised", "operator", "rng_next", "rng_to_float", "rng_normal", "consider", "code_ids", "change_direction"}
    fn = next((n for n in names if n not in small), names[0] if names else "?")
    e = by_func[fn]
    e[0] += inst; e[1] += thr; e[2] += samp
    tot[0] += inst; tot[1] += thr; tot[2] += samp
print(f"total warp instr {tot[0]:.4g}, thread instr {tot[1]:.4g}, avg threads {tot[1]/max(1,tot[0]):.2f}, samples {tot[2]}, unmapped rows {missing}")
print(f"{'file:line':40s} {'warp-inst%':>10s} {'thr-inst%':>10s} {'samples%':>9s} {'avg thr':>8s}")
for key, e in sorted(by_line.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{key[0] + ':' + str(key[1]):40s} {100*e[0]/tot[0]:10.2f} {100*e[1]/tot[1]:10.2f} {100*e[2]/max(1,tot[2]):9.2f} {e[1]/max(1,e[0]):8.1f}")
print()
print(f"{'function':40s} {'warp-inst%':>10s} {'thr-inst%':>10s} {'samples%':>9s} {'avg thr':>8s}")
for key, e in sorted(by_func.items(), key=lambda kv: -kv[1][0]):
    print(f"{key:40s} {100*e[0]/tot[0]:10.2f} {100*e[1]/tot[1]:10.2f} {100*e[2]/max(1,tot[2]):9.2f} {e[1]/max(1,e[0]):8.1f}")

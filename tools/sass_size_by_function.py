#!/usr/bin/env python
"""Static code size of one kernel by source function (innermost function of our sources on the inline chain, or the
CUDA header function when none): where the instruction-cache footprint goes.
usage: sass_size_by_function.py <nvdisasm -gi -c output> <mangled kernel substring> [top N]"""
import re
import sys
from collections import Counter

sass, kernel = sys.argv[1:3]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
lines = open(sass).read().split("\n")
start = next(i for i, l in enumerate(lines) if l.startswith(".text.") and kernel in l)
by_file_line, by_inline = Counter(), Counter()
chain, fresh, total = [], True, 0
for l in lines[start + 1:]:
    if l.startswith(".text.") or l.startswith("//--------------------- "):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', l)
    if m:
        if fresh:
            chain, fresh = [], False
        chain.append((m.group(1).split("/")[-1], int(m.group(2))))
        continue
    if re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l):
        fresh = True
        total += 1
        if chain:
            by_file_line[chain[0]] += 1
            # outermost frame inside our kernel source = the call site in the kernel
            by_inline[" <- ".join(f"{f}:{n}" for f, n in chain[:3])] += 1
print(f"{total} instructions = {total * 16} bytes")
print("-- by innermost source line")
for (f, n), c in by_file_line.most_common(top):
    print(f"{c:6d}  {f}:{n}")
print("-- by inline chain (innermost first, 3 levels)")
for k, c in by_inline.most_common(top):
    print(f"{c:6d}  {k}")

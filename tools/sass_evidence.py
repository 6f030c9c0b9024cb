#!/usr/bin/env python
"""Static evidence of the shipped kernels for profiles/: `ptxas -v` lines (registers, spills, stack) and SASS
excerpts (instruction counts, the node / triangle / wave-state load instructions) of libpt_b200's kernels.cu.

  python tools/sass_evidence.py profiles/r07       (after `make -C path-tracer_b200/csrc`)
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
B = os.path.join(ROOT, "path-tracer_b200", "csrc", "_build")
out_dir = os.path.join(ROOT, sys.argv[1] if len(sys.argv) > 1 else "profiles/r07")
os.makedirs(out_dir, exist_ok=True)

log = open(os.path.join(B, "kernels.ptxas.log")).read().splitlines()
with open(os.path.join(out_dir, "ptxas_v.txt"), "w") as f:
    f.write("# nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 --fmad=false -Xptxas -v  kernels.cu\n")
    for i, l in enumerate(log):
        if "Compiling entry function" in l:
            name = re.search(r"_ZN3ptb\d+(k_[a-z_0-9]+?)(ILi(\d+)E)?E", l)
            label = (name.group(1) + ("<%s>" % name.group(3) if name.group(3) else "")) if name else l
            f.write("%s\n  %s\n  %s\n" % (label, log[i + 2].strip(), log[i + 3].strip()))

sass = subprocess.run(["cuobjdump", "-sass", os.path.join(B, "kernels.o")], capture_output=True, text=True).stdout
kern, rows = None, collections.OrderedDict()
for l in sass.splitlines():
    m = re.search(r"Function : _ZN3ptb\d+(k_[a-z_0-9]+?)(ILi(\d+)E)?E", l)
    if m:
        kern = m.group(1) + ("<%s>" % m.group(3) if m.group(3) else "")
        rows[kern] = []
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", l)
    if m and kern:
        rows[kern].append((m.group(1), m.group(2).strip()))
with open(os.path.join(out_dir, "sass_excerpt.txt"), "w") as f:
    f.write("# cuobjdump -sass kernels.o: per kernel, instruction count, opcode histogram of the memory instructions,\n"
            "# and the wide loads (LDG.E.256 = one 32-byte chunk of a BVH node per instruction; LDG.E.128 = one float4 of a\n"
            "# triangle / wave-state record; LDL/STL.64 = one packed stack word)\n")
    for k, r in rows.items():
        if k.split("<")[0] not in ("k_extend", "k_shadow", "k_shade", "k_hit_list", "k_raygen", "k_resolve", "k_tonemap"):
            continue
        ops = collections.Counter(re.sub(r"^@!?U?P\d\s+", "", x[1]).split()[0] for x in r)
        mem = {o: c for o, c in ops.items() if re.match(r"(LDG|STG|LDL|STL|ATOMG|REDG|RED|ATOM|LDS|STS|LDC|MUFU|CALL)", o)}
        f.write("\n== %s: %d instructions (%d KB)\n   %s\n" % (k, len(r), len(r) * 16 // 1024,
                                                              ", ".join("%s x%d" % kv for kv in sorted(mem.items()))))
        shown = 0
        for a, ins in r:
            if re.search(r"LDG\.E\.(ENL2\.)?256|LDL\.64|STL\.64", ins) and shown < 10:
                f.write("   /*%s*/ %s\n" % (a, ins))
                shown += 1
print(open(os.path.join(out_dir, "ptxas_v.txt")).read())

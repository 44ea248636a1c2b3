"""Compact SASS listing of the iteration kernels for profiles/: per kernel the instruction count, an opcode histogram and the lines that
prove a hardware path (HMMA = tensor-core mma.sync, UCGABAR = cluster barrier, LDGSTS = cp.async, ATOMG / RED = global atomics).
usage: python tools/sass_listing.py > profiles/r2_iteration_kernels_sass.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "dynamicfuion_python_b200", "lib", "libnnrt_b200.so")
KERNELS = ["warp_vertices_kernel", "face_raster_kernel", "band_raster_kernel", "assemble_kernel", "system_stem_kernel", "reduced_gather_kernel",
           "corner_solve_regtile_kernel", "corner_cluster_factor_kernel", "corner_cluster_finish_kernel"]
PROOF = ("HMMA", "UCGABAR", "LDGSTS", "ATOMG", "RED.", "ATOM.")

out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
current, body = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        current = m.group(1)
        body[current] = []
    elif current and re.match(r"\s+/\*[0-9a-f]{4}\*/", line):
        body[current].append(line.strip())
print(f"# cuobjdump -sass {os.path.relpath(LIB, ROOT)} (sm_100a), kernels of one GN iteration\n")
for name, lines in body.items():
    if not any(k in name for k in KERNELS):
        continue
    ops = collections.Counter()
    for l in lines:
        parts = l.split("*/", 1)[1].split()
        op = parts[1] if parts[0].startswith("@") else parts[0]
        ops[op.split(".")[0].rstrip(";")] += 1
    print(f"## {demangle(name)}\n{len(lines)} instructions; " + ", ".join(f"{o} {c}" for o, c in ops.most_common(14)))
    shown = collections.Counter()
    for l in lines:
        for p in PROOF:
            if p in l and shown[p] < 4:
                shown[p] += 1
                print("    " + re.sub(r"\s+/\* 0x[0-9a-f]+ \*/$", "", l))
    print()

#!/usr/bin/env python
"""Opcode census of the shipped library: how many tcgen05 / TMEM / TMA / mbarrier instructions each kernel holds.

    python tools/sass_census.py [path/to/libvq_b200.so] > profiles/rNN/sass_census.txt

UTCHMMA = tcgen05.mma (kind::f16), UTCQMMA/UTCOMMA other kinds, LDTM / STTM = tcgen05.ld / st, UTMALDG = TMA tensor load,
UBLKCP = cp.async.bulk, UTCBAR = tcgen05.commit, SYNCS = mbarrier ops, LDGSTS = cp.async, RED / ATOM = global atomics.
Runs cuobjdump -sass (no GPU needed)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "vqvae_b200", "libvq_b200.so")
OPS = ["UTCHMMA", "UTCQMMA", "UTCOMMA", "LDTM", "STTM", "UTCBAR", "UTMALDG", "UBLKCP", "SYNCS", "LDGSTS", "REDG", "RED.",
       "ATOMG", "ATOMS", "FMNMX", "VIMNMX", "HMNMX2", "MATCH", "ACQBULK", "UTCATOMSWS"]
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
name, per, total = None, collections.OrderedDict(), collections.Counter()
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip() or m.group(1)
        name = name.replace("vq::(anonymous namespace)::", "").replace("(anonymous namespace)::", "")
        name = re.sub(r"^void ", "", name)
        depth, cut = 0, len(name)
        for i, ch in enumerate(name):          # cut the argument list: first "(" outside template brackets
            depth += ch == "<"
            depth -= ch == ">"
            if ch == "(" and depth == 0:
                cut = i
                break
        name = name[:cut]
        per[name] = collections.Counter()
        continue
    if name is None or "/*" not in line:
        continue
    m = re.search(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if not m:
        continue
    op = m.group(1)
    per[name]["instructions"] += 1
    for o in OPS:
        if op.startswith(o):
            per[name][o.rstrip(".")] += 1
            total[o.rstrip(".")] += 1
print(f"# {os.path.relpath(lib, ROOT)}: {len(per)} kernels; totals: " + ", ".join(f"{k} {v}" for k, v in total.most_common()))
for k, c in per.items():
    ops = ", ".join(f"{o} {n}" for o, n in c.items() if o != "instructions")
    print(f"{k:90s} {c['instructions']:6d} instr  {ops}")

#!/usr/bin/env python
"""profiles/rN_sass_opcodes.txt: SASS opcode histogram per kernel of libgforb.so (cuobjdump -sass), so that kernel-level
claims (VIMNMX3 arc network, IDP.4A orientation, UTCIMMA / LDTM / UBLKCP in the tensor-core all-pairs kernel, no HMMA
anywhere) can be checked without rebuilding.  usage: python tools/sass_opcodes.py > profiles/r2_sass_opcodes.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
txt = subprocess.run(["cuobjdump", "-sass", os.path.join(ROOT, "gf_orb_slam_b200", "libgforb.so")], capture_output=True, text=True).stdout
SPECIAL = ("UTCIMMA", "UTCQMMA", "UTCHMMA", "LDTM", "STTM", "UTCBAR", "UTCATOMSWS", "UBLKCP", "UTMALDG", "SYNCS", "HMMA", "IMMA", "POPC", "IDP", "VIMNMX3",
           "DFMA", "DMUL", "DADD", "MUFU")
print("# cuobjdump -sass gf_orb_slam_b200/libgforb.so: opcode histogram per kernel (base mnemonic: count)")
print("# tcgen05.mma -> UTCIMMA, tcgen05.ld -> LDTM, tcgen05.commit -> UTCBAR, tcgen05.alloc -> UTCATOMSWS, cp.async.bulk -> UBLKCP, mbarrier -> SYNCS")
for m in sorted(re.finditer(r"Function : (\S+)\n(.*?)(?=\n\s*Function : |\Z)", txt, re.S), key=lambda m: m.group(1)):
    name, body = m.group(1), m.group(2)
    ops = collections.Counter()
    for op in re.findall(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", body, re.M):
        ops[op.split(".")[0]] += 1
    demangled = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip().split("(")[0]
    top = " ".join("%s:%d" % kv for kv in ops.most_common(12))
    sp = " ".join("%s:%d" % (k, ops[k]) for k in SPECIAL if ops.get(k))
    print("%s  total=%d\n    top: %s\n    of note: %s" % (demangled, sum(ops.values()), top, sp or "-"))

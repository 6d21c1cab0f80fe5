#!/usr/bin/env python
"""Static SASS summary of libsphb.so per kernel (cuobjdump -sass): instruction count, FP64 / integer / memory /
control mix, local-memory (spill) instructions, async-copy mnemonics.  Writes a markdown table to stdout."""
import collections
import os
import re
import subprocess
import sys

LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "sph_b200", "libsphb.so")
out = subprocess.run(["cuobjdump", "-sass", LIB], stdout=subprocess.PIPE, text=True).stdout
kern = None
stats = collections.OrderedDict()
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], stdout=subprocess.PIPE, text=True).stdout.strip()
        kern = re.sub(r"\(sphb::\w+(, int)?\)$|\(.*\)$", "", kern).replace("sphb::", "").replace("void ", "")
        stats[kern] = collections.Counter()
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and kern:
        stats[kern][m.group(2)] += 1
CLASSES = [("FP64", ("DADD", "DMUL", "DFMA", "DSETP", "DMNMX")), ("int/ALU", ("IMAD", "IADD3", "LOP3", "SHF", "ISETP", "SEL", "FSEL", "VIADD", "VIMNMX", "LEA", "MOV", "CS2R", "UMOV", "PRMT", "IABS", "FSETP", "I2FP", "F2F", "F2I", "I2F", "MUFU")),
           ("shared", ("LDS", "STS", "LDSM")), ("global", ("LDG", "STG", "LD", "ST", "ATOMG", "RED", "ATOM", "LDC", "ULDC")),
           ("local (spill)", ("LDL", "STL")), ("control", ("BRA", "BSSY", "BSYNC", "EXIT", "WARPSYNC", "BAR", "CALL", "RET", "BREAK", "NANOSLEEP")),
           ("shuffle/vote", ("SHFL", "VOTE", "VOTEU", "REDUX", "MATCH"))]
ASYNC = ("UBLKCP", "UTMALDG", "LDGSTS", "UTMASTG", "SYNCS")
keep = [k for k in stats if re.match(r"k_(force_sweep<true, 8>|force_sweep<true, 16>|force_first<0, true, true, false>|force_refine<true, true>|density<true, true, false>|kick_drift|sweep_verify|developed|count_stats)", k)]
print("| kernel | SASS instructions | " + " | ".join(c for c, _ in CLASSES) + " | async copy (UBLKCP/UTMALDG/LDGSTS) |")
print("|---|---|" + "---|" * (len(CLASSES) + 1))
for k in keep:
    c = stats[k]
    tot = sum(c.values())
    cols = []
    for name, ops in CLASSES:
        n = sum(c[o] for o in ops)
        cols.append("%d (%.0f %%)" % (n, 100.0 * n / tot))
    print("| `%s` | %d | %s | %d |" % (k, tot, " | ".join(cols), sum(c[o] for o in ASYNC)))

#!/usr/bin/env python
"""Opcode / stall breakdown of one kernel from an ncu report's source page.
usage: tools/ncu_ops.py report.ncu-rep kernel_regex [nwarps]"""
import collections, csv, io, re, subprocess, sys

rep, kre = sys.argv[1], sys.argv[2]
nw = float(sys.argv[3]) if len(sys.argv) > 3 else None
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
# the first kernel block whose "Function Name" matches the regex
fi = next(i for i, r in enumerate(rows) if r and r[0] in ("Function Name", "Kernel Name") and re.search(kre, r[1]))
hi = next(i for i, r in enumerate(rows) if i > fi and "Source" in r and "Instructions Executed" in r)
hdr = rows[hi]
ie, isrc, ismp = hdr.index("Instructions Executed"), hdr.index("Source"), hdr.index("# Samples")
agg, smp, tot = collections.Counter(), collections.Counter(), 0
stalls = collections.Counter()
snames = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
n = 0
for r in rows[hi + 1:]:
    if len(r) <= ie or not r[ie].isdigit():
        if n: break
        continue
    n += 1
    s = re.sub(r"^@!?U?P\d+\s+", "", r[isrc].strip())
    op = s.split()[0]
    op = ".".join(op.split(".")[:2]) if op.startswith(("LDS", "STS", "IMAD", "LDG", "STG", "LDL", "STL")) else op.split(".")[0]
    v = int(r[ie]); agg[op] += v; tot += v; smp[op] += int(r[ismp])
    for sn in snames:
        stalls[sn] += int(r[hdr.index(sn)])
print("SASS lines", n, "total warp inst", tot, ("per warp %.0f" % (tot / nw)) if nw else "")
for op, v in agg.most_common(34):
    print(f"{op:14s} {v:14d} {v / tot:.3f}" + (f" {v / nw:9.1f}/warp" if nw else "") + f"  samples {smp[op]}")
ts = sum(stalls.values())
for k, v in stalls.most_common(12):
    print(f"{k:28s} {v:9d} {v / ts:.3f}")

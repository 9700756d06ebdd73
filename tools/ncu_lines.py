#!/usr/bin/env python
"""Per-source-line instruction / sample / shared-wavefront breakdown of one kernel from an ncu report.
usage: tools/ncu_lines.py report.ncu-rep kernel_regex [topN]"""
import csv, io, subprocess, sys
rep, kre = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
import re
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
fi = next(i for i, r in enumerate(rows) if r and r[0] in ("Function Name", "Kernel Name") and re.search(kre, r[1]))
hi = next(i for i, r in enumerate(rows) if i > fi and r and r[0] == "Line No")
hdr = rows[hi]
ie, ismp = hdr.index("Instructions Executed"), hdr.index("# Samples")
iw = hdr.index("L1 Wavefronts Shared") if "L1 Wavefronts Shared" in hdr else None  # absent when the kernel uses no shared memory
iwi = hdr.index("L1 Wavefronts Shared Ideal") if iw is not None else None
lines = []
for r in rows[hi + 1:]:
    if len(r) <= ie: continue
    if r[0] in ("Function Name", "Kernel Name"): break  # next kernel
    if r[0] == "Line No": continue  # next source file of the same kernel
    if r[0].isdigit():
        I = lambda s: int(s) if s.lstrip("-").isdigit() else 0
        lines.append((int(r[0]), r[1].strip(), I(r[ie]), I(r[ismp]), I(r[iw]) if iw is not None else 0, I(r[iwi]) if iw is not None else 0))
ti = sum(l[2] for l in lines); ts = sum(l[3] for l in lines); tw = sum(l[4] for l in lines)
print(f"total inst {ti} samples {ts} shared wavefronts {tw} (ideal {sum(l[5] for l in lines)})")
key = (lambda l: -l[2]) if (len(sys.argv) > 4 and sys.argv[4] == "inst") else (lambda l: -l[3])
for l in sorted(lines, key=key)[:top]:
    print(f"{l[0]:5d} inst {l[2]/ti:6.3f} smp {l[3]/ts:6.3f} wf {l[4]/max(tw,1):6.3f} (x{l[4]/max(l[5],1):4.1f}) | {l[1][:100]}")

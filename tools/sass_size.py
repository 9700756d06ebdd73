#!/usr/bin/env python
"""Static SASS size of one kernel by source line (icache budget).  usage: tools/sass_size.py obj.o kernel_regex [top]"""
import collections, re, subprocess, sys, tempfile, os, glob
obj, kre = os.path.abspath(sys.argv[1]), sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
d = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", obj], cwd=d, capture_output=True)
cub = glob.glob(os.path.join(d, "*.cubin"))[0]
out = subprocess.run(["nvdisasm", "-g", "-c", cub], capture_output=True, text=True).stdout
cnt = collections.Counter(); cur = None; infn = False; total = 0
for ln in out.splitlines():
    m = re.match(r"\s*\.text\.(\S+):", ln)
    if m:
        infn = re.search(kre, m.group(1)) is not None
        continue
    if not infn: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    if re.match(r"\s+/\*[0-9a-f]{4,5}\*/", ln):
        cnt[cur] += 1; total += 1
print("total instructions", total, "=", total * 16 // 1024, "KB")
for k, v in cnt.most_common(top): print(f"{v:6d} {k}")

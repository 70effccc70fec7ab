#!/usr/bin/env python
"""Static SASS instruction count per source line for one kernel of libbphot.so (nvdisasm -g listing).
usage: sass_lines.py <kernel-substring> [topN]"""
import re, subprocess, sys, collections, os, tempfile, glob
pat = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
so = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "pythonphot_b200", "csrc", "libbphot.so")
d = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=d, stdout=subprocess.DEVNULL)
txt = subprocess.run(["nvdisasm", "-g", "-c"] + glob.glob(d + "/*.cubin"), capture_output=True, text=True).stdout
cur = None; loc = None; cnt = collections.Counter(); total = 0; infn = False
for ln in txt.splitlines():
    m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
    if m: infn = pat in m.group(1); continue
    if not infn: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m: loc = (os.path.basename(m.group(1)), int(m.group(2))); continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln):
        cnt[loc] += 1; total += 1
print("kernel ~%s: %d SASS instructions (%.1f KB)" % (pat, total, total * 16 / 1024.))
for (f, l), c in cnt.most_common(top):
    print("%5d  %s:%d" % (c, f, l))

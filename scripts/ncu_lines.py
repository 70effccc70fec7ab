#!/usr/bin/env python
"""Per-source-line summary of an ncu report's source page (samples, warp instructions, top stall reasons).
usage: ncu_lines.py report.ncu-rep [topN]"""
import csv, subprocess, sys, io, collections
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
fname = None; hdr = None; agg = {}; tot_s = 0; tot_i = 0
for r in rows:
    if len(r) >= 2 and r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if len(r) >= 2 and r[0] == "Function Name": continue
    if len(r) > 5 and r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) < len(hdr) or r[0] == "": continue
    d = dict(zip(hdr[4:], r[4:]))
    try: s = int(d["# Samples"]); ins = int(d["Instructions Executed"])
    except Exception: continue
    st = {k[6:]: int(v) for k, v in d.items() if k.startswith("stall_") and "Not Issued" not in k and v.isdigit() and int(v)}
    key = (fname, int(r[0]), r[1].strip()[:90])
    a = agg.setdefault(key, [0, 0, collections.Counter()])
    a[0] += s; a[1] += ins; a[2].update(st); tot_s += s; tot_i += ins
print("total samples %d, warp instructions %d" % (tot_s, tot_i))
allst = collections.Counter()
for a in agg.values(): allst.update(a[2])
print("stalls overall:", ", ".join("%s %.1f%%" % (k, 100. * v / max(tot_s, 1)) for k, v in allst.most_common(10)))
print("%-22s %6s %6s  %s" % ("file:line", "samp%", "inst%", "source | top stalls"))
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%-22s %6.2f %6.2f  %s | %s" % ("%s:%d" % key[:2], 100. * a[0] / max(tot_s, 1), 100. * a[1] / max(tot_i, 1), key[2],
                                        ",".join("%s:%d" % kv for kv in a[2].most_common(3))))

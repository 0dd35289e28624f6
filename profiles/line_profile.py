#!/usr/bin/env python
"""Per-source-line warp-instruction counts of one kernel from an ncu report captured with --import-source on.
usage: line_profile.py report.ncu-rep [top_n]   (reads `ncu -i ... --page source --csv --print-source cuda,sass`;
the CUDA line rows carry the sums over their SASS rows; inlined device functions are charged to their own lines)"""
import csv, io, subprocess, sys, collections

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 60
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
per = {}
cur_file = None
hdr = None
for row in csv.reader(io.StringIO(txt)):
    if not row:
        continue
    if row[0] in ("File Name", "File Path"):
        cur_file = row[1].split("/")[-1]; continue
    if row[0] == "Line No":
        hdr = row; i_inst = hdr.index("Instructions Executed"); i_smp = hdr.index("# Samples"); continue
    if hdr is None or row[0] in ("Kernel Name", "Function Name", ""):
        continue
    try:
        line = int(row[0]); n = int(row[i_inst]); s = int(row[i_smp])
    except ValueError:
        continue
    per[(cur_file, line)] = (n, s, row[1].strip())
tot = sum(v[0] for v in per.values()); tots = sum(v[1] for v in per.values())
print("total warp instructions", tot, "samples", tots)
byfile = collections.Counter(); byfile_s = collections.Counter()
for (f, l), (n, s, src) in per.items():
    byfile[f] += n; byfile_s[f] += s
for f, n in byfile.most_common():
    print("%-28s %12d  %5.1f %% inst  %5.1f %% samples" % (f, n, 100.0 * n / tot, 100.0 * byfile_s[f] / max(tots, 1)))
print()
if len(sys.argv) > 3:     # ranges: file:lo-hi,...
    for spec in sys.argv[3].split(","):
        f, r = spec.split(":"); lo, hi = map(int, r.split("-"))
        n = sum(v[0] for (ff, l), v in per.items() if ff == f and lo <= l <= hi)
        s = sum(v[1] for (ff, l), v in per.items() if ff == f and lo <= l <= hi)
        print("%-40s %12d %5.1f %% inst %5.1f %% samples" % (spec, n, 100.0 * n / tot, 100.0 * s / max(tots, 1)))
    print()
for (f, l), (n, s, src) in sorted(per.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%-20s %5d %11d %5.1f%% smp %5.1f%% | %s" % (f, l, n, 100.0 * n / tot, 100.0 * s / max(tots, 1), src[:100]))

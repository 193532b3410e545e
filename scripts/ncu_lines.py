#!/usr/bin/env python
"""Per-source-line instruction counts of one kernel from an .ncu-rep captured with --import-source on (read here, no GPU):
   python scripts/ncu_lines.py REPORT KERNEL_REGEX [top N]      (first matching launch only)"""
import csv, io, subprocess, sys, collections, os

rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + kern],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
agg = collections.OrderedDict()
cur_file, hdr, seen_files = None, None, set()
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        if r[1] in seen_files:      # second launch starts: stop
            break
        cur_file = r[1]; seen_files.add(cur_file); hdr = None
        continue
    if r[0] == "Function Name":
        continue
    if r[0] == "Line No":
        hdr = r
        iI, iT, iS = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
        iL = hdr.index("stall_long_sb")
        continue
    if hdr is None or len(r) < len(hdr):
        continue
    if r[0] != "":
        try:
            line = int(r[0]); src = r[1].strip()[:90]
        except ValueError:
            pass
        continue                    # the source line's own summary row: skip, sum the SASS rows under it
    n = len(hdr)
    g = lambda i: r[i - n]          # metrics counted from the end: the source text may hold unescaped quotes / commas
    k = (os.path.basename(cur_file), line)
    a = agg.setdefault(k, [0, 0, 0, 0, src])
    try:
        a[0] += int(g(iI) or 0); a[1] += int(g(iT) or 0); a[2] += int(g(iS) or 0); a[3] += int(g(iL) or 0)
    except ValueError:
        pass
totI = sum(a[0] for a in agg.values()); totT = sum(a[1] for a in agg.values()); totS = sum(a[2] for a in agg.values())
print(f"total warp-inst {totI}  thread-inst {totT}  avg threads/inst {totT / max(totI, 1):.2f}  samples {totS}")
byfile = collections.Counter()
for (f, l), a in agg.items():
    byfile[f] += a[0]
print("by file:", {f: f"{100 * v / totI:.1f}%" for f, v in byfile.most_common()})
print(f"{'file:line':24s} {'%inst':>6s} {'thr/inst':>8s} {'%samples':>8s} {'%long_sb':>8s}  source")
key = 2 if "--by-samples" in sys.argv else 0
for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][key])[:top]:
    print(f"{f + ':' + str(l):24s} {100 * a[0] / totI:6.2f} {a[1] / max(a[0], 1):8.1f} {100 * a[2] / max(totS, 1):8.2f} {100 * a[3] / max(totS, 1):8.2f}  {a[4]}")
if "--ranges" in sys.argv:
    # cumulative per file in source order, for section totals
    for f in byfile:
        print("==", f)
        acc = 0
        for (ff, l), a in sorted(agg.items(), key=lambda kv: kv[0][1]):
            if ff == f and a[0]:
                print(f"  {l:5d} {100 * a[0] / totI:6.2f}% thr/inst {a[1] / max(a[0], 1):5.1f} samp {100 * a[2] / max(totS, 1):5.2f}%  {a[4][:70]}")

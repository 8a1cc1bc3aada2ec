"""Aggregate an ncu `--page source --csv --print-source cuda,sass` export into step-kernel phases (developer tool)."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
cur = None; hdr = None; agg = []
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]; continue
    if r and r[0] == "Line No":
        hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0] != "":
        d = dict(zip(hdr[4:], r[4:]))
        agg.append((cur, int(r[0]), int(d["Instructions Executed"]), int(d["# Samples"]), int(d["Thread Instructions Executed"]), int(d.get("stall_barrier", 0) or 0)))
tot = sum(a[2] for a in agg); tots = sum(a[3] for a in agg); totb = sum(a[5] for a in agg)
print("warp instr %d  samples %d  (barrier-stall samples %d)" % (tot, tots, totb))
by = {}
for f, ln, ins, smp, thr, bar in agg:
    k = f
    e = by.setdefault(k, [0, 0, 0, 0]); e[0] += ins; e[1] += smp - bar; e[2] += thr; e[3] += bar
for k, e in sorted(by.items(), key=lambda kv: -kv[1][1]):
    print("%-24s %5.1f%% ins  %5.1f%% non-barrier samples  thr/ins %4.1f  barrier %5.1f%%" % (k, 100 * e[0] / tot, 100 * e[1] / max(tots - totb, 1), e[2] / max(e[0], 1), 100 * e[3] / max(totb, 1)))
if len(sys.argv) > 2:
    # line ranges "file:lo-hi=name"
    for spec in sys.argv[2:]:
        loc, name = spec.split("=")
        f, rg = loc.split(":"); lo, hi = map(int, rg.split("-"))
        ins = sum(a[2] for a in agg if a[0] == f and lo <= a[1] <= hi)
        smp = sum(a[3] - a[5] for a in agg if a[0] == f and lo <= a[1] <= hi)
        thr = sum(a[4] for a in agg if a[0] == f and lo <= a[1] <= hi)
        print("  %-28s %5.1f%% ins  %5.1f%% non-barrier samples  thr/ins %4.1f" % (name, 100 * ins / tot, 100 * smp / max(tots - totb, 1), thr / max(ins, 1)))

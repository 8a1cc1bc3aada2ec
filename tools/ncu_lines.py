"""Top source lines of an ncu `--page source --csv --print-source cuda,sass` export (developer tool)."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur = None; hdr = None; agg = []
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]; continue
    if r and r[0] == "Line No":
        hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0] != "":
        d = dict(zip(hdr[4:], r[4:]))
        agg.append((cur, int(r[0]), r[1].strip()[:110], int(d["Instructions Executed"]), int(d["# Samples"]), int(d["Thread Instructions Executed"])))
tot = sum(a[3] for a in agg); tots = sum(a[4] for a in agg)
print("total warp instr %d samples %d" % (tot, tots))
for a in sorted(agg, key=lambda a: -a[4])[:topn]:
    print("%5.1f%% smp %5.1f%% ins  thr/ins %4.1f  %s:%d  %s" % (100 * a[4] / tots, 100 * a[3] / tot, a[5] / max(a[3], 1), a[0], a[1], a[2]))

"""Aggregate an ncu `--page source --csv --print-source cuda,sass` export by source file and line bucket (developer tool):
static SASS instructions, executed warp instructions, stall samples (all / no_instruction / barrier / short+long scoreboard / wait)."""
import csv
import sys
from collections import defaultdict

rows = csv.reader(open(sys.argv[1]))
bucket = int(sys.argv[2]) if len(sys.argv) > 2 else 50
cur = None; hdr = None
agg = defaultdict(lambda: [0, 0, 0, 0, 0, 0, 0])
mode = None
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]; hdr = None; continue
    if r and r[0] == "Line No":
        hdr = r; continue
    if not hdr or len(r) != len(hdr) or r[0] == "":
        continue
    d = dict(zip(hdr[4:], r[4:]))
    try:
        ln = int(r[0])
    except ValueError:
        continue
    key = (cur, ln // bucket * bucket)
    a = agg[key]
    a[1] += int(d["Instructions Executed"] or 0); a[2] += int(d["# Samples"] or 0)
    a[3] += int(d.get("stall_no_inst", 0) or 0); a[4] += int(d.get("stall_barrier", 0) or 0)
    a[5] += int(d.get("stall_short_sb", 0) or 0) + int(d.get("stall_long_sb", 0) or 0); a[6] += int(d.get("stall_wait", 0) or 0)
tot = [sum(a[i] for a in agg.values()) for i in range(7)]
print("total: executed %d  samples %d  no_inst %d (%.1f%%)  barrier %d (%.1f%%)  scoreboard %d (%.1f%%)  wait %d (%.1f%%)" %
      (tot[1], tot[2], tot[3], 100 * tot[3] / tot[2], tot[4], 100 * tot[4] / tot[2], tot[5], 100 * tot[5] / tot[2], tot[6], 100 * tot[6] / tot[2]))
print("%-28s %9s %8s %8s %8s %8s %8s" % ("file:line", "exec%", "samp%", "noinst%", "barr%", "sb%", "wait%"))
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][2])[:int(sys.argv[3]) if len(sys.argv) > 3 else 40]:
    print("%-28s %8.2f%% %7.2f%% %7.2f%% %7.2f%% %7.2f%% %7.2f%%" % ("%s:%d" % k, 100 * a[1] / tot[1], 100 * a[2] / tot[2], 100 * a[3] / max(tot[3], 1), 100 * a[4] / max(tot[4], 1),
                                                             100 * a[5] / max(tot[5], 1), 100 * a[6] / max(tot[6], 1)))

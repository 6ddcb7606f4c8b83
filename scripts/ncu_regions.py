#!/usr/bin/env python
"""Warp instructions per source line of one kernel from an .ncu-rep (innermost inlined line), per track.
usage: scripts/ncu_regions.py rep.ncu-rep kernel-regex n_tracks [min_per_track]"""
import csv, subprocess, sys
rep, pat, ntr = sys.argv[1], sys.argv[2], float(sys.argv[3])
thr = float(sys.argv[4]) if len(sys.argv) > 4 else 25.0
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + pat],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Line No"][0]
h = rows[hi]; N = len(h); ie = h.index("Instructions Executed") - N; isamp = h.index("# Samples") - N
cur = None; agg = {}; smp = {}; txt = {}
for r in rows[hi + 1:]:
    if r[0].isdigit():
        cur = int(r[0]); txt[cur] = ",".join(r[1:len(r) - N + 2])[:110]
    elif r[0] == "" and cur is not None:
        try:
            agg[cur] = agg.get(cur, 0) + float(r[ie] or 0); smp[cur] = smp.get(cur, 0) + float(r[isamp] or 0)
        except ValueError:
            pass
tot = sum(agg.values()); ts = sum(smp.values())
print("warp instructions %.0f = %.1f per track" % (tot, tot / ntr))
for k in sorted(agg):
    if agg[k] / ntr >= thr:
        print("%4d %8.1f  samp %4.1f%%  %s" % (k, agg[k] / ntr, 100 * smp[k] / ts, txt[k]))

#!/usr/bin/env python
"""Summarise an ncu run brought back in gpurun_out/<tag>/ into profiles/ (text, committed).
usage: scripts/ncu_summarize.py gpurun_out/<tag> profiles/<name>   -> <name>_launches.txt, <name>_kernels.txt
"""
import collections, csv, os, subprocess, sys

src, dst = sys.argv[1], sys.argv[2]

def launches():
    p = os.path.join(src, "launches.csv")
    if not os.path.exists(p):
        return
    rows = [r for r in csv.reader(open(p, errors="replace")) if len(r) > 5]
    hdr, agg = None, collections.defaultdict(lambda: [0, 0.0])
    for r in rows:
        if r[0] == "ID":
            hdr = r
            continue
        if hdr is None:
            continue
        d = dict(zip(hdr, r))
        if d.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(d["Metric Value"].replace(",", ""))
        u = d["Metric Unit"]
        v = v / 1e3 if u in ("ns", "nsecond") else v * 1e3 if u in ("ms", "msecond") else v
        k = d["Kernel Name"][:70]
        agg[k][0] += 1
        agg[k][1] += v
    tot = sum(v[1] for v in agg.values())
    with open(dst + "_launches.txt", "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised: compare SHARES)\n")
        f.write("# source: %s/launches.csv\n" % src)
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write("%-72s n=%4d total_us=%10.1f avg_us=%8.1f share=%.3f\n" % (k, v[0], v[1], v[1] / v[0], v[1] / tot))

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active"]

def kernels():
    rep = os.path.join(src, "prof.ncu-rep")
    if not os.path.exists(rep):
        return
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    # dram bytes per launch of each kernel family -> profiles/ncu_traffic.json (bench.py's roofline.traffic reads it)
    fam = {"klt_kernel": "klt_run_kernel", "fast_kernel": "fast_tile_kernel", "pyrdown_kernel": "pyrdown_kernel",
           "ingest_kernel": "ingest_kernel", "bucket_insert_kernel": "bucket_insert_kernel"}
    tr = {"source": rep, "unit": "bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum)"}
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for key, pat in fam.items():
        vals = [float(r[idx["dram__bytes_read.sum"]]) * mult[units[idx["dram__bytes_read.sum"]]] +
                float(r[idx["dram__bytes_write.sum"]]) * mult[units[idx["dram__bytes_write.sum"]]]
                for r in rows[2:] if pat in r[idx["Kernel Name"]]]
        if vals:
            tr[key] = max(vals)   # the level-0 launch for the per-level kernels
    # warp instructions the LK launch executed and its issue utilisation under ncu -> bench.py's roofline.issue
    for r in rows[2:]:
        if "klt_run_kernel" in r[idx["Kernel Name"]]:
            tr["klt_warp_inst_per_launch"] = float(r[idx["smsp__inst_executed.sum"]])
            tr["klt_issue_active_pct_ncu"] = float(r[idx["smsp__issue_active.avg.pct_of_peak_sustained_active"]])
            tr["klt_workload_note"] = "64 streams x 2000 tracks (bench.py default workload)"
    import hashlib, json
    h = hashlib.sha1()   # bench.py only trusts instruction counts captured from THIS klt.cu (same hash as bench.source_sha)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for fn in ("klt.cu", "klt_common.cuh"):
        h.update(open(os.path.join(root, "svo_b200", "csrc", fn), "rb").read())
    tr["klt_source_sha"] = h.hexdigest()[:16]
    json.dump(tr, open(os.path.join(os.path.dirname(dst), "ncu_traffic.json"), "w"), indent=1)
    with open(dst + "_kernels.txt", "w") as f:
        f.write("# ncu --set full --clock-control none; source: %s/prof.ncu-rep (one row block per captured launch)\n" % src)
        for r in rows[2:]:
            f.write("---- %s\n" % r[idx["Kernel Name"]][:110])
            for w in WANT:
                if w in idx:
                    f.write("  %-72s %s %s\n" % (w, r[idx[w]], units[idx[w]]))

launches()
kernels()

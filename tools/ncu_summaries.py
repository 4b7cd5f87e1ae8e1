"""Turn gpurun_out/launches.csv (ncu launch list) and gpurun_out/prof_f.ncu-rep (ncu --set full) into the text
summaries committed under profiles/, and record the per-launch DRAM traffic of every kernel in
profiles/ncu_traffic.json (bench.py reads roofline.traffic from there -- no literals in bench.py).
Usage: python tools/ncu_summaries.py r02_a [config=cfg4] [overflow=break]"""
import collections, csv, io, json, os, shutil, subprocess, sys
tag = sys.argv[1]
config = sys.argv[2] if len(sys.argv) > 2 else "cfg4"
overflow = sys.argv[3] if len(sys.argv) > 3 else "break"
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
go, pr = os.path.join(root, "gpurun_out"), os.path.join(root, "profiles")


def short(name):
    for k in ("k_insert", "k_mid", "k_gather", "k_overflow", "k_scan_big", "k_bev_map", "k_bev_canvas", "k_upload",
              "k_mean_vfe", "k_fill_zero"):
        if k in name:
            return k
    return name[:40]


# ---- launch list
lp = os.path.join(go, "launches.csv")
if os.path.exists(lp):
    rows = [r for r in csv.reader(l for l in open(lp) if not l.startswith("=="))]
    hdr = rows[0]
    ik, iv, im = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Name")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        if len(r) <= iv or r[im] != "gpu__time_duration.sum":
            continue
        a = agg.setdefault(r[ik], [0, 0.0])
        a[0] += 1; a[1] += float(r[iv].replace(",", ""))
    unit = rows[1][hdr.index("Metric Unit")]
    scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(unit, 1.0)
    tot = sum(v[1] for v in agg.values())
    with open(os.path.join(pr, f"{tag}_launches_summary.txt"), "w") as f:
        f.write(f"ncu --metrics gpu__time_duration.sum --clock-control none, python tools/ncu_step.py 64 {overflow} 4 {config} "
                "(cold-cache, serialised: compare shares, not absolutes)\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{k[:70]:70s} n={v[0]:3d} avg={v[1] / v[0] * scale:9.1f}us share={100 * v[1] / tot:5.1f}%\n")
    shutil.copy(lp, os.path.join(pr, f"{tag}_launches_{config}.csv"))
    print(open(os.path.join(pr, f"{tag}_launches_summary.txt")).read())

# ---- full set
rep = os.path.join(go, "prof_f.ncu-rep")
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
want = ["launch__grid_size", "launch__registers_per_thread", "gpu__time_duration.sum", "dram__bytes_read.sum",
        "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "lts__t_sectors_srcunit_tex_op_atom.sum",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed.sum", "smsp__average_warp_latency_per_inst_issued.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio"]
ik = hdr.index("Kernel Name")
with open(os.path.join(pr, f"{tag}_ncu_full_summary.txt"), "w") as f:
    f.write(f"ncu --set full --clock-control none, {config} ({overflow}), one launch of each kernel of a step; "
            "cold L2 (ncu flushes caches per kernel)\n")
    f.write(f"{'Kernel Name':88s} |" + "|".join(f" {short(d[ik]):>22s} " for d in data) + "\n")
    for w in want:
        if w not in hdr:
            continue
        i = hdr.index(w)
        f.write(f"{w:72s} {units[i]:15s} |" + "|".join(f" {d[i]:>22s} " for d in data) + "\n")
print(open(os.path.join(pr, f"{tag}_ncu_full_summary.txt")).read())


def to_bytes(val, unit):
    v = float(val.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
tp = os.path.join(pr, "ncu_traffic.json")
traffic = json.load(open(tp)) if os.path.exists(tp) else {}
traffic[f"{config}/{overflow}"] = {
    "source": f"profiles/{tag}_ncu_full_summary.txt (ncu --set full, cold L2, dram__bytes_read.sum + dram__bytes_write.sum per launch)",
    "kernels": {short(d[ik]): to_bytes(d[ir], units[ir]) + to_bytes(d[iw], units[iw]) for d in data}}
json.dump(traffic, open(tp, "w"), indent=1, sort_keys=True)
print(json.dumps(traffic[f"{config}/{overflow}"], indent=1))

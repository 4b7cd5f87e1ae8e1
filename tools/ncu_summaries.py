"""Turn gpurun_out/launches.csv (ncu launch list) and gpurun_out/prof_f.ncu-rep (ncu --set full)
into the text summaries committed under profiles/.  Usage: python tools/ncu_summaries.py r01_f"""
import collections, csv, io, os, subprocess, sys
tag = sys.argv[1]
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
go, pr = os.path.join(root, "gpurun_out"), os.path.join(root, "profiles")
# ---- launch list
rows = [r for r in csv.reader(l for l in open(os.path.join(go, "launches.csv")) if not l.startswith("=="))]
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
    f.write("ncu --metrics gpu__time_duration.sum --clock-control none, python bench.py --steps 2 --warmup 3 "
            "--no-e2e --no-cpu-baseline (cold-cache, serialised: compare shares)\n")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"{k[:70]:70s} n={v[0]:3d} avg={v[1] / v[0] * scale:9.1f}us share={100 * v[1] / tot:5.1f}%\n")
os.replace(os.path.join(go, "launches.csv"), os.path.join(pr, f"{tag}_launches_cfg4.csv")) if False else None
import shutil
shutil.copy(os.path.join(go, "launches.csv"), os.path.join(pr, f"{tag}_launches_cfg4.csv"))
# ---- full set
raw = subprocess.run(["ncu", "-i", os.path.join(go, "prof_f.ncu-rep"), "--page", "raw", "--csv"],
                     capture_output=True, text=True).stdout
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
    f.write("ncu --set full --clock-control none, cfg4 (64 x 120k points, break), one launch of each kernel of a step; "
            "cold L2 (ncu flushes caches per kernel)\n")
    f.write(f"{'Kernel Name':88s} |" + "|".join(f" {d[ik][:26]:>26s} " for d in data) + "\n")
    for w in want:
        if w not in hdr:
            continue
        i = hdr.index(w)
        f.write(f"{w:72s} {units[i]:15s} |" + "|".join(f" {d[i]:>26s} " for d in data) + "\n")
print(open(os.path.join(pr, f"{tag}_launches_summary.txt")).read())
print(open(os.path.join(pr, f"{tag}_ncu_full_summary.txt")).read())

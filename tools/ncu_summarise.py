"""Turn ncu output brought back in gpurun_out/ into the committed summaries under profiles/.

    python tools/ncu_summarise.py launches gpurun_out/launches.csv profiles/rN_launches_summary.txt "<command that was profiled>"
    python tools/ncu_summarise.py kernel   gpurun_out/prof.ncu-rep profiles/rN_ncu_<name>.txt [--json profiles/ncu_top_kernel.json --match spif_ffn_kernel]

`launches` aggregates the gpu__time_duration.sum pass per kernel (count, average, share of the listed launches);
`kernel` prints the headline counters of every captured launch (duration, DRAM bytes, throughput %, registers, ...) and,
with --json, writes the mean DRAM bytes per launch of the kernels matching --match (bench.py reads it as
roofline.traffic)."""
import collections
import csv
import json
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio"]

TO_BYTES = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
TO_US = {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "nsecond": 1e-3, "s": 1e6}


def launches(src, dst, what):
    rows = list(csv.reader(open(src)))
    hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    H = rows[hdr]
    ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
    agg = collections.defaultdict(list)
    for r in rows[hdr + 1:]:
        if len(r) > vi:
            v = float(r[vi].replace(",", "")) * TO_US.get(r[ui], 1.0)
            agg[r[ki].split("(")[0][:78]].append(v)
    tot = sum(sum(v) for v in agg.values())
    out = [f"ncu launch list: {what}", "gpu__time_duration.sum, --clock-control none; cold-cache and serialised: compare SHARES, not absolutes",
           f"{'kernel':80s} {'n':>5s} {'avg_us':>8s} {'share':>6s}"]
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        out.append(f"{k:80s} {len(v):5d} {sum(v) / len(v):8.2f} {sum(v) / tot * 100:5.1f}%")
    open(dst, "w").write("\n".join(out) + "\n")
    print("\n".join(out))


def kernel(src, dst, js=None, match=None):
    raw = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    H, U = rows[0], rows[1]
    out = [f"ncu --set full summary of {src}"]
    traffic, durs = [], []
    for r in rows[2:]:
        name = r[H.index("Kernel Name")]
        out.append(name[:110])
        rec = {}
        for w in WANT:
            if w in H:
                i = H.index(w)
                out.append(f"    {w:82s} {r[i]:>16s} {U[i]}")
                rec[w] = (r[i], U[i])
        if match and match in name and "dram__bytes_read.sum" in rec:
            rd = float(rec["dram__bytes_read.sum"][0].replace(",", "")) * TO_BYTES.get(rec["dram__bytes_read.sum"][1], 1.0)
            wr = float(rec["dram__bytes_write.sum"][0].replace(",", "")) * TO_BYTES.get(rec["dram__bytes_write.sum"][1], 1.0)
            traffic.append(rd + wr)
            durs.append(float(rec["gpu__time_duration.sum"][0].replace(",", "")) * TO_US.get(rec["gpu__time_duration.sum"][1], 1.0))
    open(dst, "w").write("\n".join(out) + "\n")
    print("\n".join(out))
    if js and traffic:
        json.dump({"kernel": match, "launches_captured": len(traffic), "dram_bytes_per_launch": sum(traffic) / len(traffic),
                   "us_per_launch_under_ncu": sum(durs) / len(durs), "source": src}, open(js, "w"), indent=1)
        print("wrote", js)


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else "")
    else:
        js = sys.argv[sys.argv.index("--json") + 1] if "--json" in sys.argv else None
        match = sys.argv[sys.argv.index("--match") + 1] if "--match" in sys.argv else None
        kernel(sys.argv[2], sys.argv[3], js, match)

"""Summarise ncu outputs brought back in gpurun_out/ into small text files under profiles/."""
import collections, csv, re, subprocess, sys

def launches(path, out):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", row["Kernel Name"])[:64]
        v = float(row["Metric Value"].replace(",", ""))
        v = v / 1e3 if row["Metric Unit"] == "ns" else (v * 1e3 if row["Metric Unit"] == "ms" else v)
        agg[name][0] += 1
        agg[name][1] += v
    tot = sum(v[1] for v in agg.values())
    with open(out, "w") as f:
        f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised)\n")
        f.write(f"# launches {sum(v[0] for v in agg.values())}  total {tot:.1f} us\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{k:66s} n={v[0]:5d} total_us={v[1]:10.1f} avg_us={v[1]/v[0]:8.1f} share={v[1]/tot:6.1%}\n")

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__waves_per_multiprocessor", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "lts__t_sectors_op_red.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"]

def full(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    with open(out, "w") as f:
        f.write(f"# ncu --set full --clock-control none --import-source on : {rep}\n")
        for r in rows[2:]:
            f.write("---- " + r[idx["Kernel Name"]][:70] + "\n")
            for w in WANT + [h for h in hdr if "issue_stalled" in h and "per_issue_active" in h]:
                if w in idx:
                    f.write(f"  {w:78s} {r[idx[w]]:>18s} {units[idx[w]]}\n")

if __name__ == "__main__":
    tag = sys.argv[1]
    launches("gpurun_out/launches.csv", f"profiles/{tag}_ncu_launches.txt")
    full("gpurun_out/prof_render.ncu-rep", f"profiles/{tag}_ncu_render_full.txt")

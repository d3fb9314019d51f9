"""Turn an `ncu --set full` report into the committed text summary and the per-kernel DRAM traffic table.
    python profiles/summarize_ncu.py gpurun_out/r01e_full.ncu-rep profiles/r01e_ncu_full.txt "<header line>" [--no-traffic]
Writes the summary and refreshes profiles/ncu_traffic.json (dram__bytes_read.sum + dram__bytes_write.sum per
launch, averaged over the captured launches of each kernel id bench.py reports)."""
import csv
import io
import json
import os
import subprocess
import sys

KEEP = [
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "gpu__time_duration.sum", "l1tex__t_sector_hit_rate.pct", "launch__block_size", "launch__grid_size",
    "launch__registers_per_thread", "lts__t_sector_hit_rate.pct", "sm__cycles_active.avg", "sm__cycles_elapsed.avg",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
]
# kernel-name fragment -> kernel id of bench.py's per-kernel table
IDS = [("preprocess_fwd_kernel", "preprocess_fwd"), ("blend_fwd_kernel", "blend_fwd"), ("blend_bwd_kernel", "blend_bwd"), ("blend_bwd2_kernel", "blend_bwd"),
       ("preprocess_bwd_kernel", "preprocess_bwd"), ("fill_spans_kernel", "memset")]
UNIT = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}


def main():
    rep, out, header = sys.argv[1], sys.argv[2], sys.argv[3]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    name_col = col["Kernel Name"]
    traffic = {}
    with open(out, "w") as f:
        f.write("# " + header + "\n")
        f.write(f"# source: {rep} (scratch); values per launch, made by profiles/summarize_ncu.py\n")
        for r in data:
            f.write(r[name_col][:120] + "\n")
            for k in KEEP:
                if k in col:
                    f.write(f"    {k:75s} {r[col[k]]:>16s} {units[col[k]]}\n")
            for frag, kid in IDS:
                if frag in r[name_col]:
                    tot = 0.0
                    for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                        tot += float(r[col[k]].replace(",", "")) * UNIT[units[col[k]]]
                    traffic.setdefault(kid, []).append(tot)
    tj = {k: {"dram_bytes_per_launch": sum(v) / len(v), "launches_captured": len(v)} for k, v in traffic.items()}
    tj["_source"] = (f"{out} ({header}), dram__bytes_read.sum + "
                     "dram__bytes_write.sum averaged over the captured launches")
    if "--no-traffic" in sys.argv[4:]:
        return
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ncu_traffic.json"), "w") as f:
        json.dump(tj, f, indent=1)
    print(json.dumps(tj, indent=1))


if __name__ == "__main__":
    main()

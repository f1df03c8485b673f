"""Summarise an `ncu --set full` capture (read here, no GPU needed) and tie its DRAM traffic to a build.

    python profiles/ncu_traffic.py gpurun_out/prof.ncu-rep --sha $(cat gpurun_out/src_sha16.txt) \
        --workload 50000x5000x8:1024:zipf --bracket stage_a_kernel,score_ --out profiles/r2

Writes <out>/ncu_summary_<tag>.txt (one block per captured launch: duration, DRAM bytes, instructions,
pipe utilisation, occupancy) and, with --bracket, <out>/traffic.json = dram__bytes_read.sum +
dram__bytes_write.sum summed over ONE launch of every kernel whose name contains one of the given
substrings — the figure bench.py reports as roofline.traffic when the source hash still matches.
"""
import argparse
import csv
import json
import os
import subprocess

WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__inst_executed.sum",
    "smsp__inst_executed.sum", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__issue_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__cycles_active.avg",
    "lts__t_sectors_op_write.sum", "lts__t_sectors_op_read.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
]


def to_bytes(val, unit):
    v = float(val.replace(",", ""))
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(unit, 1)
    return v * mult


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("--sha", default="")
    ap.add_argument("--workload", default="")
    ap.add_argument("--bracket", default="")
    ap.add_argument("--out", default="profiles/r2")
    ap.add_argument("--tag", default="")
    a = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", a.rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    name_i = col.get("Kernel Name")
    tag = a.tag or os.path.splitext(os.path.basename(a.rep))[0]
    os.makedirs(a.out, exist_ok=True)
    lines, seen, traffic = [], {}, {}
    for r in data:
        name = r[name_i]
        short = name.split("(")[0]
        lines.append(f"== {short}   [{name[:160]}]")
        for w in WANT:
            if w in col and r[col[w]] != "":
                lines.append(f"   {w:78s} {units[col[w]]:10s} {r[col[w]]}")
        rd = to_bytes(r[col["dram__bytes_read.sum"]], units[col["dram__bytes_read.sum"]])
        wr = to_bytes(r[col["dram__bytes_write.sum"]], units[col["dram__bytes_write.sum"]])
        seen[short] = (rd, wr)  # last captured launch of each kernel
    with open(os.path.join(a.out, f"ncu_summary_{tag}.txt"), "w") as f:
        f.write(f"# ncu --set full --clock-control none; source {a.rep}; sources sha16 {a.sha}\n" + "\n".join(lines) + "\n")
    if a.bracket:
        pats = [p for p in a.bracket.split(",") if p]
        per = {k: v for k, v in seen.items() if any(p in k for p in pats)}
        total = sum(rd + wr for rd, wr in per.values())
        traffic = {"src_sha16": a.sha, "workload": a.workload, "traffic_bytes": int(total),
                   "per_kernel": {k: {"read": int(rd), "write": int(wr)} for k, (rd, wr) in per.items()},
                   "source": f"profiles/r2/ncu_summary_{tag}.txt (ncu --set full, one launch of each bracketed kernel)"}
        with open(os.path.join(a.out, "traffic.json"), "w") as f:
            json.dump(traffic, f, indent=1)
        print(json.dumps(traffic))
    print("\n".join(lines[:60]))


if __name__ == "__main__":
    main()

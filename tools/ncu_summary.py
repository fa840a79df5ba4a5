"""Condense `ncu --set full` reports into the per-launch CSV summaries kept under profiles/.

Usage: python tools/ncu_summary.py gpurun_out/r2_stat.ncu-rep [more.ncu-rep ...] [--out profiles] [--traffic]
Writes <out>/<report name>_ncu_summary.csv (one row per profiled launch, the columns below) and, with --traffic,
rewrites profiles/traffic.json: DRAM bytes (read + write) per launch of every kernel seen, averaged over its launches.
Needs only the `ncu` CLI (no GPU)."""
import csv
import io
import json
import os
import re
import subprocess
import sys

KEEP = [
    r"gpu__time_duration\.sum$",
    r"dram__bytes_read\.sum$",
    r"dram__bytes_write\.sum$",
    r"dram__throughput\.avg\.pct_of_peak_sustained_elapsed$",
    r"lts__throughput\.avg\.pct_of_peak_sustained_elapsed$",
    r"l1tex__throughput\.avg\.pct_of_peak_sustained_elapsed$",
    r"sm__throughput\.avg\.pct_of_peak_sustained_elapsed$",
    r"sm__inst_executed\.avg\.per_cycle_elapsed$",
    r"sm__inst_issued\.avg\.pct_of_peak_sustained_active$",
    r"sm__warps_active\.avg\.pct_of_peak_sustained_active$",
    r"launch__registers_per_thread$",
    r"launch__occupancy_limit_(registers|shared_mem|warps|blocks)$",
    r"launch__shared_mem_per_block_(static|dynamic)$",
    r"smsp__average_warps?_issue_stalled_[a-z_]+_per_issue_active\.ratio$",
    r"smsp__average_warp_latency_issue_stalled_[a-z_]+\.ratio$",
    r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum$",
    r"smsp__inst_executed_op_shared_atom\.sum$",
]
HEAD = ["ID", "Kernel Name", "Block Size", "Grid Size"]


def rows_of(report):
    raw = subprocess.run(["ncu", "-i", report, "--page", "raw", "--csv"], check=True, capture_output=True,
                         text=True).stdout
    rd = list(csv.reader(io.StringIO(raw)))
    names, units, data = rd[0], rd[1], rd[2:]
    cols = [i for i, n in enumerate(names) if n in HEAD]
    for i, n in enumerate(names):
        base = n.split(".", 2)[-1] if n.count(".") >= 2 and n.split(".")[1].startswith("Triage") else n
        if any(re.search(k, n) for k in KEEP) and not n.split(".")[0].isupper():
            cols.append(i)
    header = [names[i] + (f" [{units[i]}]" if units[i] else "") for i in cols]
    return header, [[r[i] for i in cols] for r in data], names, units, data


def to_bytes(val, unit):
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[unit]
    return float(val) * mult


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    out = "profiles"
    if "--out" in sys.argv:
        out = sys.argv[sys.argv.index("--out") + 1]
        args.remove(out)
    traffic = {}
    for rep in args:
        header, rows, names, units, data = rows_of(rep)
        dst = os.path.join(out, os.path.basename(rep).replace(".ncu-rep", "") + "_ncu_summary.csv")
        with open(dst, "w", newline="") as f:
            w = csv.writer(f)
            w.writerow(header)
            w.writerows(rows)
        print("wrote", dst, len(rows), "launches")
        ir, iw, ik, it = (names.index("dram__bytes_read.sum"), names.index("dram__bytes_write.sum"),
                          names.index("Kernel Name"), names.index("gpu__time_duration.sum"))
        for r in data:
            k = re.sub(r"^void |\(.*$", "", r[ik])
            k = re.sub(r"<.*", "", k).replace("tmb::", "")
            t = traffic.setdefault(k + "@" + os.path.basename(rep).replace(".ncu-rep", ""), [])
            t.append((to_bytes(r[ir], units[ir]) + to_bytes(r[iw], units[iw]), float(r[it])))
    if "--traffic" in sys.argv:
        path = os.path.join(out, "traffic.json")
        cur = json.load(open(path)) if os.path.exists(path) else {}
        gen = cur.setdefault("per_kernel", {})
        for k, v in traffic.items():
            gen[k] = {"dram_bytes_per_launch": round(sum(b for b, _ in v) / len(v)), "launches": len(v),
                      "avg_duration_" + "us": round(sum(t for _, t in v) / len(v), 2)}
        cur["per_kernel_source"] = ("tools/ncu_summary.py --traffic over the .ncu-rep files of tools/ncu_kernels.sh: "
                                    "dram__bytes_read.sum + dram__bytes_write.sum, mean over the profiled launches")
        json.dump(cur, open(path, "w"), indent=1)
        print("updated", path)


if __name__ == "__main__":
    main()

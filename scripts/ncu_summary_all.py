"""Condense an .ncu-rep holding several kernels (one launch each) into one metric/unit/value CSV per kernel under
profiles/, plus a compact JSON table.  Usage: python scripts/ncu_summary_all.py gpurun_out/x.ncu-rep profiles/<tag>"""
import csv
import io
import json
import re
import subprocess
import sys

KEEP = ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "launch__grid_size", "launch__block_size", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "lts__t_sectors_op_red.sum", "lts__t_sectors_op_atom.sum", "lts__t_sectors_op_red.sum.per_second",
        "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sectors_srcunit_tex_op_write.sum")


def main(rep, prefix):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    names, units = rows[0], rows[1]
    table = {}
    for vals in rows[2:]:
        d = dict(zip(names, vals))
        kn = d.get("Kernel Name", "kernel")
        short = re.sub(r"[^A-Za-z0-9_]", "", re.sub(r"<.*", "", kn.split("(")[0]).split("::")[-1])
        n = 1
        key = short
        while key in table:
            n += 1
            key = f"{short}_{n}"
        with open(f"{prefix}_{key}_ncu_summary.csv", "w", newline="") as f:
            w = csv.writer(f)
            w.writerow(["metric", "unit", "value"])
            for nme, u, v in zip(names, units, vals):
                if nme in KEEP or (nme.startswith("smsp__average_warps_issue_stalled") and nme.endswith("per_issue_active.ratio")):
                    w.writerow([nme, u, v])

        def num(k):
            try:
                return float(d.get(k, "nan").replace(",", ""))
            except ValueError:
                return float("nan")
        unit = dict(zip(names, units))
        scale = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0}
        dram = num("dram__bytes_read.sum") * scale.get(unit.get("dram__bytes_read.sum"), 1.0) + \
            num("dram__bytes_write.sum") * scale.get(unit.get("dram__bytes_write.sum"), 1.0)
        table[key] = {"kernel": kn[:120], "time_us": num("gpu__time_duration.sum") * ({"ms": 1e3, "us": 1.0, "ns": 1e-3}.get(unit.get("gpu__time_duration.sum"), 1.0)),
                      "dram_bytes": dram, "warp_inst": num("smsp__inst_executed.sum"),
                      "issue_active_pct": num("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                      "warps_active_pct": num("sm__warps_active.avg.pct_of_peak_sustained_active"),
                      "l2_red_sectors": num("lts__t_sectors_op_red.sum"), "l2_atom_sectors": num("lts__t_sectors_op_atom.sum"),
                      "registers": num("launch__registers_per_thread")}
    with open(f"{prefix}_ncu_table.json", "w") as f:
        json.dump(table, f, indent=1)
    print(json.dumps(table, indent=1))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])

#!/usr/bin/env python
"""Condense an ncu report (`ncu --set full` capture) into one small table for profiles/.

    python scripts/ncu_summary.py gpurun_out/prof.ncu-rep [more.ncu-rep ...] > profiles/rNN_ncu_summary.md

Reads the raw page through `ncu -i ... --page raw --csv` and keeps the handful of counters the
roofline discussion in DESIGN.md uses: duration, DRAM bytes, L2/L1/shared throughput, tensor-pipe
activity, issue-slot use, registers and shared memory per CTA.
"""
import csv
import io
import subprocess
import sys

METRICS = [
    ("time_us", "gpu__time_duration.sum"),
    ("sm_clk_mhz", "smsp__cycles_elapsed.avg.per_second"),
    ("dram_rd_MB", "dram__bytes_read.sum"),
    ("dram_wr_MB", "dram__bytes_write.sum"),
    ("dram_pct", "dram__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("l2_pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("l1_pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("sm_pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("tensor_pct", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
    ("tensor_rt_pct", "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed"),
    ("issue_pct", "sm__inst_issued.avg.pct_of_peak_sustained_active"),
    ("ipc", "sm__inst_executed.avg.per_cycle_active"),
    ("regs", "launch__registers_per_thread"),
    ("smem_dyn_KB", "launch__shared_mem_per_block_dynamic"),
    ("grid", "launch__grid_size"),
    ("block", "launch__block_size"),
    ("occ_pct", "sm__warps_active.avg.pct_of_peak_sustained_active"),
]

SCALE = {"Gbyte": 1e3, "Mbyte": 1.0, "Kbyte": 1e-3, "byte": 1e-6, "msecond": 1e3, "usecond": 1.0, "us": 1.0,
         "ms": 1e3, "ns": 1e-3, "nsecond": 1e-3, "second": 1e6, "Ghz": 1e3, "Mhz": 1.0, "hz": 1e-6, "Khz": 1e-3}


def rows_of(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True, check=True)
    rows = list(csv.reader(io.StringIO(out.stdout)))
    return rows[0], rows[1], rows[2:]


def main():
    # --json FILE: also write {kernel: {"dram_bytes", "tensor_pipe_active", "time_us", "report"}} (first launch of each
    # kernel) - the per-launch figures bench.py quotes in its roofline block (profiles/ncu_metrics.json)
    args = sys.argv[1:]
    json_path = None
    if "--json" in args:
        k = args.index("--json")
        json_path = args[k + 1]
        args = args[:k] + args[k + 2:]
    table = {}
    print("| report | kernel | " + " | ".join(k for k, _ in METRICS) + " |")
    print("|---|---|" + "---|" * len(METRICS))
    for path in args:
        hdr, units, rows = rows_of(path)
        idx = {h: i for i, h in enumerate(hdr)}
        for r in rows:
            cells = []
            for key, name in METRICS:
                if name not in idx:
                    cells.append("-")
                    continue
                raw, unit = r[idx[name]].replace(",", ""), units[idx[name]]
                try:
                    v = float(raw)
                except ValueError:
                    cells.append(raw)
                    continue
                if key.endswith("_MB") or key in ("time_us", "sm_clk_mhz"):
                    v *= SCALE.get(unit, 1.0)
                if key == "smem_dyn_KB":
                    v *= {"Kbyte": 1.0, "byte": 1 / 1024, "Mbyte": 1024.0}.get(unit, 1.0)
                cells.append(f"{v:.1f}" if abs(v) < 1e6 and v != int(v) else f"{int(v)}")
            name = r[idx["Kernel Name"]].split("(")[0][-48:]
            print(f"| {path.split('/')[-1]} | {name} | " + " | ".join(cells) + " |")
            full = r[idx["Kernel Name"]]
            if json_path and full not in table:
                def num(i):
                    try:
                        return float(cells[i])
                    except ValueError:
                        return None
                keys = [k for k, _ in METRICS]
                rd, wr, tp = num(keys.index("dram_rd_MB")), num(keys.index("dram_wr_MB")), num(keys.index("tensor_pct"))
                table[full] = {"dram_bytes": (rd + wr) * 1e6 if rd is not None and wr is not None else None,
                               "tensor_pipe_active": tp / 100.0 if tp is not None else None,
                               "time_us": num(keys.index("time_us")), "report": path.split("/")[-1]}
    if json_path:
        import json
        with open(json_path, "w") as f:
            json.dump(table, f, indent=1)


if __name__ == "__main__":
    main()

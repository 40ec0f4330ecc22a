#!/usr/bin/env python
"""Writes profiles/ncu_fields.json: the per-launch figures only an ncu capture can give (DRAM traffic, issue-slot
utilisation, executed instructions ...), one entry per kernel key, each stamped with the sha256 of the kernel's source
file at capture time.  bench.py reads them through ncu_fields(): when the source has changed since the capture it reports
null and says why, instead of describing an older kernel.

    tools/ncu_fields.py <key> <report.ncu-rep> <source file> [summary.csv]

<key>: render_bvh_spp_kernel<false,1> | render_bvh_kernel | conv3x3_tc_rows_kernel | conv3x3_tc_kernel_split.
The first kernel of the report is used.  With [summary.csv] the selected metrics are also written there (committed)."""
import csv
import hashlib
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
WANT = {
    "gpu__time_duration.sum": ("kernel_ms_under_ncu", 1e-6),                 # ns -> ms (cold cache, serialised)
    "dram__bytes_read.sum": ("dram_read_bytes", None),
    "dram__bytes_write.sum": ("dram_write_bytes", None),
    "smsp__inst_executed.sum": ("warp_instructions_per_launch", 1.0),
    "smsp__issue_active.avg.pct_of_peak_sustained_active": ("issue_slots_busy_pct", 1.0),
    "smsp__thread_inst_executed_per_inst_executed.ratio": ("threads_per_instruction", 1.0),
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active": ("fma_pipe_pct", 1.0),
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active": ("alu_pipe_pct", 1.0),
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active": ("lsu_pipe_pct", 1.0),
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active": ("xu_pipe_pct", 1.0),
    "sm__warps_active.avg.pct_of_peak_sustained_active": ("warps_active_pct", 1.0),
    "l1tex__t_sector_hit_rate.pct": ("l1_hit_pct", 1.0),
    "launch__registers_per_thread": ("registers_per_thread", 1.0),
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": ("tensor_pipe_pct", 1.0),
}
UNIT_SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1.0, "us": 1e3, "usecond": 1e3,
              "ms": 1e6, "msecond": 1e6, "nsecond": 1.0, "second": 1e9}


def main():
    key, rep, src = sys.argv[1], sys.argv[2], sys.argv[3]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(out.splitlines()))
    head, units, vals = rows[0], rows[1], rows[2]
    entry, picked = {}, []
    for k, u, v in zip(head, units, vals):
        if k == "Kernel Name":
            entry["kernel_name"] = v
        if k in WANT:
            name, scale = WANT[k]
            try:
                x = float(v.replace(",", ""))
            except ValueError:
                continue
            x *= UNIT_SCALE.get(u, 1.0)
            if scale is not None:
                x *= scale
            entry[name] = x
            picked.append((k, u, v))
    if "dram_read_bytes" in entry and "dram_write_bytes" in entry:
        entry["traffic"] = entry["dram_read_bytes"] + entry["dram_write_bytes"]
    entry["source_file"] = str(Path(src).resolve().relative_to(ROOT))
    entry["source_sha256"] = hashlib.sha256(Path(src).read_bytes()).hexdigest()
    entry["csv"] = str(Path(sys.argv[4]).resolve().relative_to(ROOT)) if len(sys.argv) > 4 else Path(rep).name
    path = ROOT / "profiles" / "ncu_fields.json"
    data = json.loads(path.read_text()) if path.exists() else {}
    data[key] = entry
    path.write_text(json.dumps(data, indent=1, sort_keys=True) + "\n")
    if len(sys.argv) > 4:
        with open(sys.argv[4], "w") as f:
            f.write("metric,unit,value\n")
            for k, u, v in picked:
                f.write(f"{k},{u},{v}\n")
    print(json.dumps(entry, indent=1))


if __name__ == "__main__":
    main()

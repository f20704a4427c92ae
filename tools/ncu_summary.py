#!/usr/bin/env python
"""Summarise an .ncu-rep (read with `ncu -i`) into the JSON kept under profiles/: one record per captured
launch with duration, DRAM bytes, L2 atomic/reduction sectors, throughput percentages, issue utilisation and the
stall mix.  python tools/ncu_summary.py gpurun_out/x.ncu-rep [--last-per-kernel] > profiles/rNN_x.json"""
import csv
import io
import json
import subprocess
import sys

KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sectors_op_red.sum", "lts__t_sectors_op_atom.sum", "lts__t_sectors_op_read.sum", "lts__t_sectors_op_write.sum",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed_op_shared_atom.sum", "smsp__inst_executed_op_global_red.sum",
        "l1tex__t_set_accesses_pipe_lsu_mem_global_op_red.sum", "lts__t_sectors_srcunit_tex_op_red.sum"]


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    recs = []
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        rec = {"kernel": d["Kernel Name"], "id": d["ID"]}
        for k in hdr:
            if k in KEEP or "issue_stalled" in k and k.endswith("per_issue_active.ratio"):
                try:
                    v = float(d[k].replace(",", ""))
                except ValueError:
                    continue
                name = k.replace("smsp__average_warps_issue_stalled_", "stall_").replace("_per_issue_active.ratio", "")
                rec[name] = v
                u = units[hdr.index(k)]
                if u and not name.startswith("stall_"):
                    rec[name + "__unit"] = u
        recs.append(rec)
    if "--last-per-kernel" in sys.argv:
        last = {}
        for r in recs:
            last[r["kernel"]] = r
        recs = list(last.values())
    json.dump(recs, sys.stdout, indent=1)


if __name__ == "__main__":
    main()

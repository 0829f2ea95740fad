#!/usr/bin/env python3
"""profiles/traffic.json from ncu --set full captures of the render kernel (read here, no GPU):

    python tools/ncu_traffic.py key=report.ncu-rep [key=report ...] [--commit SHA] > profiles/traffic.json

key names the workload the capture belongs to, "<scene>_<W>x<H>_w<world>" (bench.py looks its own workload up).  Per launch:
warp instructions, issue-active %, duration, DRAM bytes, the bytes that went through L1 (global and local = spill traffic),
L1 misses to L2, L2 reads from the SMs, L2 misses to DRAM.  Everything bench.py's roofline block divides by a LIVE kernel
time comes from here; the capture's own duration is kept for the share check only."""
import csv
import io
import json
import subprocess
import sys

M = {"inst": "smsp__inst_executed.sum", "issue": "smsp__issue_active.avg.pct_of_peak_sustained_active", "dur": "gpu__time_duration.sum",
     "dr": "dram__bytes_read.sum", "dw": "dram__bytes_write.sum", "gld": "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
     "gldm": "l1tex__t_sectors_pipe_lsu_mem_global_op_ld_lookup_miss.sum", "gst": "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum",
     "lld": "l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum", "lst": "l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum",
     "l2r": "lts__t_sectors_srcunit_tex_op_read.sum", "l2rm": "lts__t_sectors_srcunit_tex_op_read_lookup_miss.sum",
     "lanes": "smsp__thread_inst_executed_per_inst_executed.ratio", "regs": "launch__registers_per_thread", "warps": "sm__warps_active.avg.pct_of_peak_sustained_active",
     "l1hit": "l1tex__t_sector_hit_rate.pct", "grid": "Grid Size", "name": "Kernel Name"}
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}


def load(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    d = {}
    for h, u, v in zip(hdr, units, vals):
        for k, name in M.items():
            if h == name:
                try:
                    d[k] = float(v.replace(",", "")) * SCALE.get(u, 1)
                except ValueError:
                    d[k] = v
    return d


out, commit = {}, None
args = sys.argv[1:]
if "--commit" in args:
    i = args.index("--commit"); commit = args[i + 1]; del args[i:i + 2]
for a in args:
    key, rep = a.split("=", 1)
    d = load(rep)
    out[key] = {
        "source": "%s (ncu --set full --clock-control none, one launch of %s, grid %s)" % (rep, d.get("name"), d.get("grid")),
        "commit": commit,
        "warp_instructions": int(d["inst"]), "issue_active_pct": d["issue"], "ncu_duration_ms": d["dur"],
        "active_lanes_per_instruction": d["lanes"], "registers_per_thread": int(d["regs"]), "warps_active_pct": d["warps"], "l1_hit_pct": d["l1hit"],
        "dram_read_bytes": int(d["dr"]), "dram_write_bytes": int(d["dw"]), "dram_bytes": int(d["dr"] + d["dw"]),
        "l1_global_load_bytes": int(d["gld"]) * 32, "l1_global_store_bytes": int(d["gst"]) * 32,
        "l1_local_load_bytes": int(d["lld"]) * 32, "l1_local_store_bytes": int(d["lst"]) * 32,
        "l1_global_miss_to_l2_bytes": int(d["gldm"]) * 32, "l2_read_bytes_from_sm": int(d["l2r"]) * 32, "l2_read_miss_to_dram_bytes": int(d["l2rm"]) * 32,
    }
json.dump(out, sys.stdout, indent=1)
print()

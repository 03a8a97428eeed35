#!/usr/bin/env python3
"""profiles/r2_ncu_traffic.json from an `ncu --set full` report of one render_kernel launch of the bench command: the DRAM
bytes bench.py quotes as roofline.traffic plus the issue / lane / cache figures, stamped with the hash of the device sources
the capture was taken from (bench.py drops the fields when the sources have changed since).
usage: ncu_traffic_json.py report.ncu-rep WORKLOAD_PREFIX rays_per_launch out.json"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

rep, workload, rays, out = sys.argv[1], sys.argv[2], float(sys.argv[3]), sys.argv[4]
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr, units, vals = rows[0], rows[1], rows[2]
m = {h: (vals[i], units[i]) for i, h in enumerate(hdr)}


def num(name, scale=None):
    v, u = m[name]
    x = float(v.replace(",", ""))
    mult = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}.get(u)
    return x * mult if mult else x


inst = num("smsp__inst_executed.sum")
rec = {
    "workload": workload, "kernel": m["Kernel Name"][0] if "Kernel Name" in m else "render_kernel", "kernel_source_sha16": bench.kernel_source_hash(),
    "duration_ms_under_ncu": num("gpu__time_duration.sum"),
    "dram_bytes_read": int(num("dram__bytes_read.sum")), "dram_bytes_write": int(num("dram__bytes_write.sum")),
    "l1_to_l2_write_bytes": int(num("lts__t_bytes_equiv_l1sectormiss_pipe_lsu_mem_global_op_st.sum")) if "lts__t_bytes_equiv_l1sectormiss_pipe_lsu_mem_global_op_st.sum" in m else None,
    "issue_slot_frac": num("smsp__issue_active.avg.pct_of_peak_sustained_active") / 100.0,
    "lanes_per_warp_instruction": num("smsp__thread_inst_executed_per_inst_executed.ratio"),
    "warp_instructions": inst, "warp_instructions_per_ray": inst / rays if rays else None,
    "l1_hit_pct": num("l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": num("lts__t_sector_hit_rate.pct"),
    "warps_active_pct_of_peak": num("sm__warps_active.avg.pct_of_peak_sustained_active"),
    "sm_active_avg_over_max": num("sm__cycles_active.avg") / num("sm__cycles_active.max"),
    "registers_per_thread": int(num("launch__registers_per_thread")),
    "source": os.path.basename(rep) + " (ncu --set full --clock-control none, one launch of the bench command; times under ncu are not bench values)",
}
for k in ("sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_xu.sum", "sm__inst_executed_pipe_lsu.sum"):
    if k in m:
        rec[k.replace("sm__inst_executed_", "").replace(".sum", "") + "_frac_of_warp_instructions"] = num(k) / inst
with open(out, "w") as f:
    json.dump(rec, f, indent=1)
print(json.dumps(rec))

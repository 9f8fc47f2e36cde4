#!/usr/bin/env python
"""Summarise one kernel of an .ncu-rep as JSON (bench.py reads roofline.traffic from it) and as text.
usage: ncu_to_json.py file.ncu-rep out_prefix [kernel-substring]"""
import csv, io, json, subprocess, sys
rep, prefix = sys.argv[1], sys.argv[2]
sub = sys.argv[3] if len(sys.argv) > 3 else ""
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
row = [r for r in rows[2:] if sub in r[hdr.index("Kernel Name")]][0]
def val(name):
    i = hdr.index(name)
    x = float(row[i].replace(",", ""))
    u = units[i].lower()
    scale = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "tbyte": 1e12, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1)
    return x * scale
d = {"kernel": row[hdr.index("Kernel Name")], "source_report": rep, "duration_ms": val("gpu__time_duration.sum"),
     "dram_bytes_read": int(val("dram__bytes_read.sum")), "dram_bytes_write": int(val("dram__bytes_write.sum")),
     "registers_per_thread": int(val("launch__registers_per_thread")),
     "issue_active_pct": val("smsp__issue_active.avg.pct_of_peak_sustained_active"),
     "l1tex_throughput_pct": val("l1tex__throughput.avg.pct_of_peak_sustained_elapsed"),
     "warp_instructions": int(val("smsp__inst_executed.sum")),
     "shared_bank_conflicts": int(val("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum")),
     "lts_hit_rate_pct": val("lts__t_sector_hit_rate.pct")}
for k in ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor.sum",
          "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
          "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__shared_mem_per_block_dynamic"):
    if k in hdr:
        d[k] = val(k)
json.dump(d, open(prefix + ".json", "w"), indent=1)
print(json.dumps(d, indent=1))

#!/usr/bin/env python
"""Summarise an .ncu-rep (raw page + SASS page) -> stdout / JSON. usage: ncu_summary.py rep [out.json]"""
import csv, io, json, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
keep = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'launch__grid_size', 'launch__block_size', 'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'lts__t_bytes.sum']
out = {"report": rep, "kernels": []}
for r in rows[2:]:
    out["kernels"].append({k: (r[hdr.index(k)] + " " + units[hdr.index(k)]).strip() for k in keep if k in hdr})
sass = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
kern, h, data = None, None, {}
for r in csv.reader(io.StringIO(sass)):
    if r and r[0] == 'Kernel Name': kern = r[1]; data[kern] = []; continue
    if r and r[0] == 'Address': h = r; continue
    if kern and len(r) > 10: data[kern].append(r)
for k, rs in data.items():
    ie = h.index('Instructions Executed')
    tot = sum(int(r[ie]) for r in rs)
    st = {n: sum(int(r[h.index(n)]) for r in rs) for n in h if n.startswith('stall_') and 'Not Issued' not in n}
    s = sum(st.values()) or 1
    ops = {}
    for r in rs:
        t = r[1].split()
        op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
        ops[op] = ops.get(op, 0) + int(r[ie])
    out.setdefault("sass", []).append({"kernel": k[:60], "static_instr": len(rs), "warp_instr": tot,
        "stalls": {a: round(b / s, 3) for a, b in sorted(st.items(), key=lambda x: -x[1])[:8]},
        "opcodes": {a: round(b / tot, 3) for a, b in sorted(ops.items(), key=lambda x: -x[1])[:20]}})
print(json.dumps(out, indent=1))
if len(sys.argv) > 2:
    json.dump(out, open(sys.argv[2], "w"), indent=1)

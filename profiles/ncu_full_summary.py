"""Selected counters of an `ncu --set full` capture, one record per profiled launch:
    ncu -i capture.ncu-rep --page raw --csv | python profiles/ncu_full_summary.py "<what was captured>" > summary.json
Keeps what the roofline discussion needs: duration, DRAM bytes, FP64 tensor (DMMA) pipe activity, issue-slot use,
occupancy, the warp-stall breakdown, registers, shared memory, grid."""
import csv
import json
import re
import sys

KEEP = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "dram_read"),
    ("dram__bytes_write.sum", "dram_write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"),
    ("lts__t_bytes.sum", "l2_bytes"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct"),
    ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "fp64_pipe_pct_active"),
    ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "fp64_cycles_pct_active"),
    ("smsp__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_elapsed", "dmma_pct_elapsed"),
    ("sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_elapsed", "dmma_pct_elapsed_sm"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor_pct_elapsed"),
    ("sm__inst_executed_pipe_tensor.sum", "tensor_inst"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_active_pct"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy_pct"),
    ("launch__registers_per_thread", "registers"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__shared_mem_per_block_dynamic", "smem_dynamic"),
    ("launch__occupancy_limit_registers", "occ_limit_regs"),
    ("launch__occupancy_limit_shared_mem", "occ_limit_smem"),
    ("sm__cycles_elapsed.avg", "cycles"),
    ("smsp__cycles_active.avg", "smsp_cycles_active"),
]
rows = list(csv.reader(line for line in sys.stdin if line.startswith('"')))
if not rows:
    sys.exit("no csv rows on stdin")
hdr = rows[0]
col = {name: i for i, name in enumerate(hdr)}
units = rows[1] if len(rows) > 1 and not rows[1][0].isdigit() else None
data = rows[2:] if units else rows[1:]
out = []
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio")]
dmma_cols = [(i, h) for i, h in enumerate(hdr) if "dmma" in h.lower()]
for r in data:
    name = re.sub(r"\(.*$", "", r[col["Kernel Name"]]).replace("void ", "").replace("b200::", "")
    rec = {"kernel": name, "id": r[col["ID"]]}
    for metric, key in KEEP:
        if metric in col:
            v = r[col[metric]].replace(",", "")
            try:
                rec[key] = float(v)
            except ValueError:
                rec[key] = v
            if units:
                rec[key + "_unit"] = units[col[metric]]
    stalls = {}
    for i, h in stall_cols:
        try:
            stalls[h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]] = float(r[i].replace(",", ""))
        except ValueError:
            pass
    rec["stalls_per_issue"] = dict(sorted(stalls.items(), key=lambda kv: -kv[1])[:6])
    rec["dmma_counters"] = {h: r[i] for i, h in dmma_cols}
    out.append(rec)
print(json.dumps({"what": sys.argv[1] if len(sys.argv) > 1 else "", "launches": out}, indent=1))

"""Summarise an `ncu --set full` report: one launch per distinct kernel, the metrics DESIGN.md quotes.
usage: ncu_summarize.py report.ncu-rep out_summary.txt out_traffic.json"""
import csv, io, json, subprocess, sys
rep, out_txt, out_json = sys.argv[1:4]
METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
           "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
           "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
           "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
           "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "lts__t_sector_hit_rate.pct",
           "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers"]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
names, units = rows[hdr], rows[hdr + 1]
col = {n: i for i, n in enumerate(names)}
seen, txt, traffic = {}, [], {}
def to_bytes(v, u):
    v = float(v.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
for r in rows[hdr + 2:]:
    if len(r) < len(names):
        continue
    k = r[col["Kernel Name"]]
    short = k.split("(")[0].replace("void ", "").replace("r3dgw::", "")
    if short in seen:
        continue
    seen[short] = True
    txt.append(f"== {short}")
    for m in METRICS:
        if m in col:
            txt.append(f"  {m:<82} {r[col[m]]:>14} {units[col[m]]}")
    dur_ms = to_bytes(r[col["gpu__time_duration.sum"]], "") * {"ns": 1e-6, "us": 1e-3, "ms": 1, "s": 1e3}.get(units[col["gpu__time_duration.sum"]], 1)
    traffic[short] = {"dram_traffic_bytes_per_launch": to_bytes(r[col["dram__bytes_read.sum"]], units[col["dram__bytes_read.sum"]]) +
                      to_bytes(r[col["dram__bytes_write.sum"]], units[col["dram__bytes_write.sum"]]), "duration_ms": dur_ms}
open(out_txt, "w").write("\n".join(txt) + "\n")
json.dump(traffic, open(out_json, "w"), indent=1)
print("\n".join(txt[:60]))

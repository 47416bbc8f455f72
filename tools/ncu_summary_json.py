"""Key metrics of the last kernel in an .ncu-rep as the JSON summary bench.py reads for `roofline.traffic`.
usage: ncu_summary_json.py file.ncu-rep out.json <elements> [note]"""
import csv, json, subprocess, sys
rep, out, nelem = sys.argv[1], sys.argv[2], int(sys.argv[3])
note = sys.argv[4] if len(sys.argv) > 4 else ""
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines())); h, units, r = rows[0], rows[1], rows[-1]
want = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__warps_active.avg.per_cycle_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"]
d = {}
for i, x in enumerate(h):
    if x in want or ("issue_stalled" in x and x.endswith("per_issue_active.ratio")):
        d[x] = {"value": r[i], "unit": units[i]}
def to_bytes(m):
    v, u = float(d[m]["value"]), d[m]["unit"].lower()
    return v * {"byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "tbyte": 1e12}[u]
tot = to_bytes("dram__bytes_read.sum") + to_bytes("dram__bytes_write.sum")
d["derived"] = {"dram_bytes_per_launch": tot, "elements": nelem, "dram_bytes_per_element": tot / nelem, "note": note}
json.dump(d, open(out, "w"), indent=1)
print(out, d["Kernel Name"]["value"], d["gpu__time_duration.sum"], "B/element", tot / nelem)

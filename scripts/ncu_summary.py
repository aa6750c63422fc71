"""Summarise `ncu --page raw --csv` exports: one JSON entry + one markdown row per captured kernel launch.
usage: ncu_summary.py out_prefix raw1.csv [raw2.csv ...]"""
import csv, json, sys
KEEP = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "launch__registers_per_thread",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers", "launch__waves_per_multiprocessor",
        "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio", "dram__bytes_read.sum",
        "dram__bytes_write.sum", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "sm__icc_request_hit_rate.pct",
        "gcc__average_cache_request_hit_rate.pct", "sm__cycles_elapsed.max",
        "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum", "smsp__sass_thread_inst_executed_op_dadd_pred_on.sum",
        "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum", "sass__inst_executed_local_loads", "sass__inst_executed_local_stores"]
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6, "usecond": 1, "msecond": 1e3, "nsecond": 1e-3}
def num(v):
    try: return float(v.replace(",", ""))
    except Exception: return v
out, md = {}, ["| kernel | grid x block | time (us) | issue active % | FP64 pipe % | warps active % | lanes/inst | DRAM read+write (bytes) | top stall reasons |", "|---|---|---|---|---|---|---|---|---|"]
for f in sys.argv[2:]:
    rows = list(csv.reader(open(f)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d, stalls = {}, {}
        for i, h in enumerate(hdr):
            if h in KEEP:
                v = num(r[i])
                if units[i] in SCALE and isinstance(v, float): v *= SCALE[units[i]]
                d[h] = v
            elif h.startswith("smsp__pcsamp_warps_issue_stalled_") and "not_issued" not in h:
                stalls[h[len("smsp__pcsamp_warps_issue_stalled_"):]] = num(r[i])
        tot = sum(v for v in stalls.values() if isinstance(v, float)) or 1.0
        d["stall_share"] = {k: round(v / tot, 3) for k, v in sorted(stalls.items(), key=lambda kv: -kv[1])[:6]}
        name = "%s grid %s" % (d["Kernel Name"].split("(")[0].replace("void ", ""), d["Grid Size"])
        k, n = name, 1
        while k in out: n += 1; k = "%s #%d" % (name, n)
        out[k] = d
        md.append("| %s | %s x %s | %.1f | %.1f | %.1f | %.1f | %.1f | %.3g | %s |" % (
            d["Kernel Name"].split("(")[0].replace("void ", ""), d["Grid Size"], d["Block Size"], d["gpu__time_duration.sum"],
            d["smsp__issue_active.avg.pct_of_peak_sustained_active"], d["sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"],
            d["sm__warps_active.avg.pct_of_peak_sustained_active"], d["smsp__thread_inst_executed_per_inst_executed.ratio"],
            d["dram__bytes_read.sum"] + d["dram__bytes_write.sum"], ", ".join("%s %.0f%%" % (a, 100 * b) for a, b in list(d["stall_share"].items())[:4])))
json.dump(out, open(sys.argv[1] + ".json", "w"), indent=1)
open(sys.argv[1] + ".md", "w").write("\n".join(md) + "\n")
print("\n".join(md))

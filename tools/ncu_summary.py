"""Summarise an .ncu-rep (ncu --set full) into profiles/: the raw per-launch CSV and a small JSON of the metrics DESIGN.md cites.
usage: python tools/ncu_summary.py gpurun_out/prof_trace.ncu-rep profiles/r01c_k_trace  "note"
"""
import os as _os, sys as _sys; _sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))
import csv, io, json, subprocess, sys

rep, out, note = sys.argv[1], sys.argv[2], (sys.argv[3] if len(sys.argv) > 3 else "")
rays_per_launch = float(sys.argv[4]) if len(sys.argv) > 4 else None      # rays one captured launch traces (from the same run's stats)
if rep.endswith(".csv"):        # the raw page already exported on the GPU box (reports above gpurun's 64 MB limit)
    raw = open(rep).read()
else:
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
open(out + "_full_raw.csv", "w").write(raw)
rows = list(csv.reader(io.StringIO(raw)))
hdr, data = rows[0], rows[2:]
col = {h: i for i, h in enumerate(hdr)}


def num(r, name):
    i = col.get(name)
    if i is None or r[i] in ("", "n/a"):
        return None
    return float(r[i].replace(",", ""))


units = dict(zip(hdr, rows[1]))
launches = []
for r in data:
    dur = num(r, "gpu__time_duration.sum")
    dunit = units.get("gpu__time_duration.sum", "ns")
    dur_us = dur / 1e3 if dunit in ("ns", "nsecond") else (dur if dunit in ("us", "usecond") else dur * 1e3)
    rd, wr = num(r, "dram__bytes_read.sum"), num(r, "dram__bytes_write.sum")
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    rd_b = rd * scale.get(units.get("dram__bytes_read.sum", "byte"), 1) if rd is not None else None
    wr_b = wr * scale.get(units.get("dram__bytes_write.sum", "byte"), 1) if wr is not None else None
    launches.append({
        "kernel": r[col["Kernel Name"]][:90],
        "grid": r[col["Grid Size"]], "block": r[col["Block Size"]],
        "duration_us": round(dur_us, 3),
        "dram_bytes": (rd_b or 0) + (wr_b or 0),
        "dram_pct_of_peak": num(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
        "l2_hit_pct": num(r, "lts__t_sector_hit_rate.pct"),
        "threads_per_inst": num(r, "smsp__thread_inst_executed_per_inst_executed.ratio"),
        "issue_active_pct": num(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
        "alu_pipe_pct": num(r, "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
        "fma_pipe_pct": num(r, "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
        "xu_pipe_pct": num(r, "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
        "lsu_pipe_pct": num(r, "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
        "regs_per_thread": num(r, "launch__registers_per_thread"),
        "inst_executed": num(r, "smsp__inst_executed.sum"),
        "rays": rays_per_launch,
        "achieved_occupancy_pct": num(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
        "local_ld_st_sectors": (num(r, "l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum") or 0) + (num(r, "l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum") or 0),
    })
big = [l["dram_bytes"] for l in launches if l["dram_bytes"]]
json.dump({"source": rep.split("/")[-1] + " (ncu --set full --clock-control none --import-source on)", "note": note, "launches": launches,
           "traffic_bytes_per_launch": (sum(big) / len(big)) if big else None},
          open(out + "_summary.json", "w"), indent=1)
for l in launches:
    print(l["kernel"][:60].ljust(60), l["duration_us"], "us  dram MB", round(l["dram_bytes"] / 1e6, 1), " thr/inst", l["threads_per_inst"], " issue", l["issue_active_pct"],
          " alu", l["alu_pipe_pct"], " xu", l["xu_pipe_pct"], " occ", l["achieved_occupancy_pct"], " regs", l["regs_per_thread"])

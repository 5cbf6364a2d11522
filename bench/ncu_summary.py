"""Summarise an ncu report into a markdown table for profiles/ (the metrics the profiling
recipe names: duration, DRAM bytes, tensor-pipe activity, warps, registers, top stall reasons).

    python bench/ncu_summary.py gpurun_out/prof.ncu-rep profiles/ncu_xyz.md "title"
"""
import csv
import subprocess
import sys

rep, out, title = sys.argv[1], sys.argv[2], sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL,
                     text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
W = {
    "Kernel Name": "kernel", "gpu__time_duration.sum": "time", "launch__grid_size": "grid",
    "launch__block_size": "block", "launch__registers_per_thread": "regs",
    "launch__shared_mem_per_block_dynamic": "dyn smem", "dram__bytes_read.sum": "DRAM rd",
    "dram__bytes_write.sum": "DRAM wr", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "DRAM %",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor pipe % (active)",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "SM %",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "warps active %",
}
idx = {k: hdr.index(k) for k in W if k in hdr}
stall = [(i, h) for i, h in enumerate(hdr) if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
lines = ["# " + title, "", "Source: `%s` (`ncu --set full --clock-control none`, one B200)." % rep.split("/")[-1],
         "Durations under ncu are serialised/cold-cache: compare shares, not absolutes.", "",
         "| " + " | ".join(W[k] for k in idx) + " | top stalls (warp-cycles per issue) |", "|" + "---|" * (len(idx) + 1)]
for r in rows[2:]:
    cells = []
    for k, i in idx.items():
        v = r[i]
        if k == "Kernel Name":
            v = "`" + v.split("(")[0].replace("psb::", "")[:48] + "`"
        else:
            try:
                f = float(v)
                u = units[i]
                v = ("%.1f %s" % (f, u)) if f < 1e4 else ("%.3g %s" % (f, u))
            except ValueError:
                pass
        cells.append(v)
    st = sorted(((float(r[i] or 0), h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""))
                 for i, h in stall), reverse=True)[:3]
    cells.append(", ".join("%s %.1f" % (n, v) for v, n in st))
    lines.append("| " + " | ".join(cells) + " |")
open(out, "w").write("\n".join(lines) + "\n")
print("\n".join(lines))

"""Print the hottest SASS instructions of one kernel from an ncu report (source page).

    python bench/ncu_hot.py gpurun_out/prof.ncu-rep lstm_fwd [N]
"""
import csv
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern],
                     stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
hdr = rows[1]
c = hdr.index("Warp Stall Sampling (All Samples)")
stalls = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
data = []
for r in rows[2:]:
    if len(r) <= c or r[0] == "Address" or r[0] == "Kernel Name":
        continue
    try:
        v = float(r[c] or 0)
    except ValueError:
        continue
    why = sorted(((float(r[i] or 0), hdr[i][6:]) for i in stalls if i < len(r)), reverse=True)[:2]
    data.append((v, r[1].strip()[:90], why))
tot = sum(d[0] for d in data) or 1
print("total samples", tot, "instructions", len(data))
for v, s, why in sorted(data, key=lambda d: -d[0])[:top]:
    print("%5.1f%%  %-90s %s" % (100 * v / tot, s, ",".join("%s:%d" % (n, x) for x, n in why if x)))

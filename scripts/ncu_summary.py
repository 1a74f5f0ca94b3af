"""Per-kernel table from an ncu report: python scripts/ncu_summary.py report.ncu-rep  (runs `ncu -i ... --page raw --csv`)."""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
ix = {h: i for i, h in enumerate(hdr)}
cols = ["gpu__time_duration.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "smsp__issue_active.avg.pct_of_peak_sustained_active"]
stalls = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
if not stalls:
    stalls = [h for h in hdr if h.startswith("smsp__average_warp_latency_issue_stalled_") or h.startswith("smsp__average_warps_issue_stalled_")]
units = rows[1]
for r in rows[2:]:
    name = r[ix["Kernel Name"]][:70]
    vals = []
    for c in cols:
        if c in ix:
            vals.append(f"{c.split('.')[0].replace('__', ' ')}={r[ix[c]]}{units[ix[c]] if units[ix[c]] not in ('', '%') else ('%' if units[ix[c]] == '%' else '')}")
    st = []
    for s in stalls:
        try:
            st.append((float(r[ix[s]].replace(",", "")), s.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")))
        except ValueError:
            pass
    st.sort(reverse=True)
    print(f"{name}\n    " + "; ".join(vals) + "\n    top stalls (warps stalled per issue): " + ", ".join(f"{n} {v:.2f}" for v, n in st[:4]))

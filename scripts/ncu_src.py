"""Summarise an ncu report's source page: samples per SASS region and the top stalled instructions.
usage: python scripts/ncu_src.py report.ncu-rep kernel_regex [group]"""
import csv, subprocess, sys
rep, rx = sys.argv[1], sys.argv[2]
grp = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = rows[0]
want = ["gpu__time_duration.sum", "sm__cycles_elapsed.max", "smsp__inst_executed.sum", "launch__registers_per_thread",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"]
for r in rows[2:]:
    print({w: r[hdr.index(w)] for w in want if w in hdr})
    st = {h.replace("smsp__pcsamp_warps_issue_stalled_", ""): int(float(r[i])) for i, h in enumerate(hdr)
          if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued") and r[i] not in ("", "0")}
    print(sorted(st.items(), key=lambda kv: -kv[1])[:8])
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", f"::regex:{rx}:1"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
hdr = rows[1]
si, ii = hdr.index("# Samples"), hdr.index("Instructions Executed")
acc, seen = [], set()
for n, r in enumerate(rows[2:]):
    try:
        s, e = int(r[si]), int(r[ii])
    except Exception:
        continue
    if r[0] in seen:
        break
    seen.add(r[0])
    acc.append((n, r[1].strip(), s, e))
tot = sum(a[2] for a in acc)
print("samples", tot, "instructions", len(acc), "executed", sum(a[3] for a in acc))
for i in range(0, len(acc), grp):
    ch = acc[i:i + grp]
    print(f"{i:5d} {sum(c[2] for c in ch):5d} {sum(c[3] for c in ch):9d}  {ch[0][1][:60]}")
print("top:")
for c in sorted(acc, key=lambda c: -c[2])[:20]:
    print(c)

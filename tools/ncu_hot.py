"""Print the hottest SASS instructions (by warp-stall samples) of an `ncu --page source --csv` dump."""
import csv
import sys

path, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25
rows = list(csv.reader(open(path)))
hdr = rows[1]
i_src, i_smp, i_exe = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") or h.startswith("Stall")]
data = rows[2:]
tot = sum(int(r[i_smp]) for r in data)
print("total samples", tot, "instructions", len(data))
print("stall columns:", [hdr[i] for i in stall_cols][:40])
order = sorted(range(len(data)), key=lambda k: -int(data[k][i_smp]))[:top]
for k in sorted(order):
    r = data[k]
    reasons = sorted(((int(r[i]), hdr[i]) for i in stall_cols if r[i] not in ("", "0")), reverse=True)[:2]
    print(f"{k:5d} {100*int(r[i_smp])/tot:5.1f}% exec={r[i_exe]:>9} {r[i_src].strip()[:70]:70s} {reasons}")

"""Per-instruction stall summary of one kernel from `ncu -i X.ncu-rep --page source --csv ...` output.
    python tools/ncu_stalls.py src.csv [top]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = next(r for r in rows if r and r[0] == "Address")
ix = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows if len(r) == len(hdr) and r[0] != "Address"]


def num(r, k):
    try:
        return int(r[ix[k]])
    except (ValueError, KeyError):
        return 0


tot = sum(num(r, "# Samples") for r in data)
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = {s: sum(num(r, s) for r in data) for s in stalls}
print("kernel:", rows[0][1][:100], " total samples", tot)
for s, v in sorted(agg.items(), key=lambda x: -x[1])[:10]:
    print(f"  {s:24s} {v:8d} {100.0 * v / max(tot, 1):5.1f}%")
for r in sorted(data, key=lambda r: -num(r, "# Samples"))[:top_n]:
    st = sorted(((num(r, s), s) for s in stalls), reverse=True)[:2]
    print(str(num(r, "# Samples")).rjust(6), str(num(r, "Instructions Executed")).rjust(9),
          r[ix["Source"]].strip()[:64].ljust(64), [(v, s[6:]) for v, s in st if v])

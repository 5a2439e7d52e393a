"""python tools/ab/choose_scan.py v2 v3 ...  -> the name of the candidate build whose 32-query search (scan + merge, queued
back to back) over the 625k-row shard is the shortest in gpurun_out/f_probe_<name>.txt, or nothing when none beats the
default build's (f_probe_a.txt).  Used by tools/ab/r2_final.sh."""
import re
import sys


def search_us(name):
    try:
        for ln in open(f"gpurun_out/f_probe_{name}.txt"):
            if ln.startswith("rows   625000"):
                return float(re.search(r"back to back\s+([0-9.]+) us", ln).group(1))
    except Exception:
        pass
    return 1e30


cands = sorted((search_us(v), v) for v in sys.argv[1:])
print(cands[0][1] if cands and cands[0][0] < search_us("a") else "")

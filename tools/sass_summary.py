"""SASS / resource summary of libbsc_hot.so for profiles/: per kernel the Blackwell mnemonics (tcgen05 = UTC*MMA / LDTM /
STTM, TMA = UTMALDG / UTMASTG / UBLKCP), legacy HMMA (must be 0), and registers / spills from `nvcc -Xptxas -v`.
    python tools/sass_summary.py > profiles/r2_sass_summary.txt"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "bioscan-clip_b200"))
import build  # noqa: E402

lib = build.build_lib()
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
kern, cur = {}, None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        kern[cur] = []
    elif cur:
        kern[cur].append(line)
demangle = subprocess.run(["c++filt"], input="\n".join(kern), capture_output=True, text=True).stdout.splitlines()
keys = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTCBAR", "SYNCS", "MUFU.EX2", "HMMA"]
print("libbsc_hot.so: %d kernels, sm_100a (cuobjdump -sass | grep -c per kernel)" % len(kern))
print("%-74s" % "kernel" + " ".join("%7s" % k[-7:] for k in keys))
tot = dict.fromkeys(keys, 0)
for name, dm in zip(kern, demangle):
    body = "\n".join(kern[name])
    row = []
    for k in keys:
        if k == "UTCHMMA":
            c = len(re.findall(r"UTCHMMA", body))
        elif k == "HMMA":
            c = len(re.findall(r"(?<!UTC)HMMA", body))
        else:
            c = body.count(k)
        tot[k] += c
        row.append(c)
    short = re.sub(r"\(.*", "", dm.replace("(anonymous namespace)::", "").replace("bsc::", "")).replace("void ", "")[:72]
    if any(row[:8]):
        print("%-74s" % short + " ".join("%7d" % c for c in row))
print("%-74s" % "TOTAL (all %d kernels; those without tensor / TMA ops are not listed)" % len(kern)
      + " ".join("%7d" % tot[k] for k in keys))
print()
print("registers / spills (nvcc -Xptxas -v):")
for src in build.SOURCES:
    cmd = [build.NVCC, *build.FLAGS, "-Xptxas=-v", "-c", os.path.join(build.CSRC, src), "-o", "/dev/null"]
    err = subprocess.run(cmd, capture_output=True, text=True).stderr
    fn = None
    for line in err.splitlines():
        m = re.search(r"Compiling entry function '(\S+)'", line)
        if m:
            fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            fn = re.sub(r"\(.*", "", fn.replace("(anonymous namespace)::", "").replace("bsc::", "")).replace("void ", "")
        m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
        if m:
            spill = (m.group(2), m.group(3))
        m = re.search(r"Used (\d+) registers", line)
        if m and fn:
            print("  %-80s %4s regs  spill st/ld %s/%s B" % (fn[:80], m.group(1), spill[0], spill[1]))

"""Counts the SASS mnemonics that prove what the kernels are built from (DPX min-plus, warp reductions, cp.async, mbarrier; no
TMA / tensor-core instructions on this path), per kernel of the shipped library.  No GPU needed:
    python tests/sass_counts.py > profiles/r02_sass_counts.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "stereo_matching_b200", "libsgm_b200.so")
WATCH = ["VIMNMX3", "VIADDMNMX", "VIMNMX", "CREDUX", "REDUX", "FMNMX3", "FMNMX", "LDGSTS", "LDGDEPBAR", "SYNCS", "SHFL", "VOTE", "POPC", "MATCH",
         "UTMALDG", "UTMASTG", "UBLKCP", "UTCHMMA", "UTCQMMA", "HMMA", "LDTM", "STTM", "FFMA", "FADD", "FMUL", "MUFU", "ATOM", "RED", "NANOSLEEP"]
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
per = collections.OrderedDict()
cur = None
for ln in sass.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        cur = per.setdefault(m.group(1), collections.Counter())
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", ln)
    if m and cur is not None:
        op = m.group(1)
        cur["_total"] += 1
        base = op.split(".")[0]
        if base in WATCH:
            cur[base] += 1
        if base in ("VIMNMX3", "VIADDMNMX", "VIMNMX") and ".S16x2" in op.upper().replace("X2", "x2").upper():
            cur[base + ".S16X2"] += 1
total = collections.Counter()
print(f"# SASS mnemonic counts per kernel of {os.path.relpath(lib, ROOT)} (cuobjdump -sass; sm_100a).  Columns: only non-zero counts.")
for name, c in per.items():
    short = demangle(name).replace("(anonymous namespace)::", "").replace("sgmb200::", "").replace("void ", "")
    short = re.sub(r">\(.*", ">", short) if ">(" in short else re.sub(r"\(.*", "", short)  # drop the parameter list, keep template arguments
    items = " ".join(f"{k}={v}" for k, v in sorted(c.items()) if k != "_total" and v)
    print(f"{short[:110]:<110} instr={c['_total']:<6} {items}")
    total.update(c)
print("#")
print("# whole library: " + " ".join(f"{k}={v}" for k, v in sorted(total.items()) if k != "_total"))
print("# tensor-core / TMA mnemonics (UTC*MMA, HMMA, LDTM, STTM, UTMALDG, UTMASTG, UBLKCP): " +
      str(sum(total[k] for k in ("UTCHMMA", "UTCQMMA", "HMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP"))) + " (none: nothing on this path is a contraction; tiles are staged with cp.async = LDGSTS)")

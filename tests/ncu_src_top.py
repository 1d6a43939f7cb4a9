"""Dev helper: top stalled SASS instructions of one kernel from `ncu --page source --csv` output.
usage: python tests/ncu_src_top.py file.csv [kernel_index] [topN]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 40
# split into kernels
kernels = []
for r in rows:
    if r and r[0] == "Kernel Name":
        kernels.append({"name": r[1], "hdr": None, "data": []})
    elif kernels and kernels[-1]["hdr"] is None:
        kernels[-1]["hdr"] = r
    elif kernels:
        kernels[-1]["data"].append(r)
for i, k in enumerate(kernels):
    print(i, k["name"][:90], len(k["data"]))
k = kernels[which]
hdr, data = k["hdr"], [r for r in k["data"] if len(r) > 5]
col = {h: i for i, h in enumerate(hdr)}
S, X = col["# Samples"], col["Instructions Executed"]
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[S]) for r in data)
print("total samples", tot, "executed", sum(int(r[X]) for r in data))
agg = {s: sum(int(r[col[s]]) for r in data) for s in stalls}
print({s: v for s, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]})
top = sorted(range(len(data)), key=lambda i: -int(data[i][S]))[:topn]
for i in sorted(top):
    r = data[i]
    st = sorted(((int(r[col[s]]), s[6:]) for s in stalls), reverse=True)[:2]
    print(i, r[col["Source"]][:80].ljust(80), r[S], r[X], st)

"""Dev helper: warp-stall samples of one kernel from `ncu --page source --csv --print-source cuda,sass`:
totals per stall reason, and the source lines with the most samples (with their top reasons).
usage: python tests/ncu_stalls.py file.csv <kernel-substring> [topN]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2]
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 30
hdr = None
cur_fn = None
per_reason = collections.Counter()
per_line = collections.defaultdict(lambda: [0, 0, collections.Counter(), ""])
cur_file = None
seen_sections = collections.Counter()
for r in rows:
    if not r:
        continue
    if r[0] == "Function Name":
        cur_fn = r[1]
        seen_sections[cur_fn] += 1
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or cur_fn is None or want not in cur_fn or not r[0].strip().isdigit():
        continue
    # only CUDA-source sections carry a file; take the per-line view (first 'Source' column is the source text)
    try:
        smp = int(r[6])
        ins = int(r[7])
    except (ValueError, IndexError):
        continue
    key = (cur_file, int(r[0]))
    per_line[key][0] += smp
    per_line[key][1] += ins
    per_line[key][3] = r[1].strip()[:90]
    for i, name in enumerate(hdr):
        if name.startswith("stall_") and i < len(r) and r[i]:
            try:
                v = int(r[i])
            except ValueError:
                continue
            per_reason[name] += v
            per_line[key][2][name] += v
tot = sum(per_reason.values())
print("kernel ~ %r: %d stall samples, %d instructions" % (want, tot, sum(v[1] for v in per_line.values())))
for k, v in per_reason.most_common():
    print("  %-24s %7d  %5.1f %%" % (k, v, 100.0 * v / max(tot, 1)))
print("top lines by samples:")
for k, v in sorted(per_line.items(), key=lambda kv: -kv[1][0])[:topn]:
    top = ", ".join("%s %d" % (n.replace("stall_", ""), c) for n, c in v[2].most_common(3))
    print("  %-16s %4d smp %6d ins %9d  [%s]  %s" % (k[0], k[1], v[0], v[1], top, v[3]))

"""Dev helper: executed warp instructions per CUDA source line of one kernel, per pixel.
usage: python tests/ncu_line_top.py file_from_print-source_cuda,sass.csv <kernel-substring> [pixels] [topN]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2]
px = float(sys.argv[3]) if len(sys.argv) > 3 else 375 * 1242 * 2
topn = int(sys.argv[4]) if len(sys.argv) > 4 else 40
cur_file = cur_fn = None
agg = collections.defaultdict(lambda: [0, 0, ""])
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        cur_fn = r[1]
        continue
    if r[0] == "Line No" or cur_fn is None or want not in cur_fn:
        continue
    if r[0].strip().isdigit():
        try:
            s, x = int(r[6]), int(r[7])
        except ValueError:
            continue
        key = (cur_file, int(r[0]))
        agg[key][0] += s
        agg[key][1] += x
        agg[key][2] = r[1].strip()[:100]
tot = sum(v[1] for v in agg.values())
print("attributed instr/pixel %.1f, samples %d" % (tot / px, sum(v[0] for v in agg.values())))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:topn]:
    print("%-18s %5d %6.1f smp %5d  %s" % (k[0], k[1], v[1] / px, v[0], v[2]))

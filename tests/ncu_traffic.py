"""Dev helper: per-kernel DRAM traffic / pipe utilisation table from an `ncu --set full` capture.

usage: python tests/ncu_traffic.py gpurun_out/r01_vX_agg.ncu-rep <tag> [out.json]
Writes profiles/aggregate_traffic.json (or out.json): bench.py reads `dram_bytes_per_frame` for roofline.traffic.
"""
import csv
import io
import json
import subprocess
import sys

WANT = {
    "read": "dram__bytes_read.sum",
    "write": "dram__bytes_write.sum",
    "us": "gpu__time_duration.sum",
    "dram_pct_of_peak": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "issue_active_pct": "sm__inst_issued.avg.pct_of_peak_sustained_active",
    "pipe_alu_pct": "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "pipe_fma_pct": "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "pipe_lsu_pct": "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "registers": "launch__registers_per_thread",
    "warp_instructions": "smsp__inst_executed.sum",
}
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "ms": 1e3}
AGG = ("sgm_hpair", "sgm_sweep", "isgm_")


def short(name):
    name = name.replace("void ", "").replace("unnamed>::", "")
    return name.split("(")[0]


def main():
    rep, tag = sys.argv[1], sys.argv[2]
    out = sys.argv[3] if len(sys.argv) > 3 else "profiles/aggregate_traffic.json"
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    kn = hdr.index("Kernel Name")
    per = {}
    for r in rows[2:]:
        d = {}
        for k, m in WANT.items():
            if m in hdr:
                c = hdr.index(m)
                d[k] = float(r[c].replace(",", "")) * SCALE.get(units[c], 1)
        per.setdefault(short(r[kn]), d)  # first launch of each kernel
    total = sum(v["read"] + v["write"] for k, v in per.items() if k.startswith(AGG))
    doc = {
        "per_kernel": per,
        "dram_bytes_per_frame": total,
        "source": f"ncu --set full --clock-control none (tests/collect_profiles.sh {tag}): dram__bytes_read.sum + "
        "dram__bytes_write.sum of the aggregation kernels (horizontal pair + down sweep + up sweep; one launch each per "
        "frame, both views)",
    }
    with open(out, "w") as f:
        json.dump(doc, f, indent=1)
    for k, v in per.items():
        print(k[:70], {a: round(b, 1) for a, b in v.items()})
    print("aggregation DRAM bytes per frame:", total)


if __name__ == "__main__":
    main()

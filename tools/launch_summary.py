"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list of bench.py: per-kernel count / time / share
for ONE training step (the step boundaries are the launches of the kernel that reads the fp32 NCHW batch).  usage: launch_summary.py file.csv [step]"""
import collections
import csv
import re
import sys


def load(path):
    rows = []
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    for row in csv.DictReader(lines):
        if row.get("Metric Name") == "gpu__time_duration.sum":
            rows.append((row["Kernel Name"], float(row["Metric Value"].replace(",", "")), row.get("Grid Size", ""), row.get("Block Size", "")))
    return rows


def main():
    rows = load(sys.argv[1])
    which = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    # a step starts with the kernel that takes the fp32 NCHW batch in: the fused staging kernel (round 2) or nchw_to_nhwc
    starts = [i for i, r in enumerate(rows) if "stem_stage_nchw" in r[0]]
    if len(starts) < 2:
        starts = [i for i, r in enumerate(rows) if "nchw_to_nhwc" in r[0]]
    step = rows[starts[which]:starts[which + 1]]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for k, v, _, _ in step:
        name = re.sub(r"\(.*", "", k).replace("void ", "")
        m = re.match(r"(b200::\w+)<(.*)>", name)
        if m and ("conv_sm100" in name or "gemm_sm100" in name):
            name = "%s<%s>" % (m.group(1), m.group(2).replace("(int)", "").replace("(bool)", ""))
        else:
            name = re.sub(r"<.*", "", name)
        agg[name][0] += 1
        agg[name][1] += v
    tot = sum(v for _, v in agg.values())
    print("one training step = %d launches, %.2f ms summed (ncu: serialised, cold-cache)" % (len(step), tot / 1e6))
    print("%-46s %6s %12s %7s" % ("kernel", "count", "sum us", "share"))
    for k, (n, v) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print("%-46s %6d %12.1f %6.1f%%" % (k[:46], n, v / 1000, 100 * v / tot))
    if "--detail" in sys.argv:
        for k, v, g, b in step:
            if "conv_sm100" in k or "gemm_sm100" in k:
                print("%10.1f us grid %-18s %s" % (v / 1000, g, re.sub(r"\(.*", "", k)[:90]))


if __name__ == "__main__":
    main()

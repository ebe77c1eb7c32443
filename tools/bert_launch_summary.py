"""Per-kernel count / time / share of ONE BERT training step from an `ncu --metrics gpu__time_duration.sum --csv`
launch list of tools/bert_step.py (steps are delimited by the embed_ln_fwd launches).
usage: bert_launch_summary.py file.csv [step_index]"""
import collections
import re
import sys

sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.abspath(__file__)))
import launch_summary as L  # noqa: E402


def main():
    rows = L.load(sys.argv[1])
    starts = [i for i, r in enumerate(rows) if "embed_ln_fwd" in r[0]]
    which = int(sys.argv[2]) if len(sys.argv) > 2 else len(starts) - 1
    end = starts[which + 1] if which + 1 < len(starts) else len(rows)
    step = rows[starts[which]:end]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for k, v, _, _ in step:
        name = re.sub(r"\(.*", "", k).replace("void ", "")
        m = re.match(r"(b200::\w+)<(.*)>", name)
        if m:
            name = "%s<%s>" % (m.group(1), m.group(2).replace("(int)", "").replace("__nv_bfloat16", "bf16"))
        agg[name][0] += 1
        agg[name][1] += v
    tot = sum(v for _, v in agg.values())
    print("launches %d   sum of kernel time %.3f ms (cold-cache, serialised under ncu)" % (len(step), tot / 1e6))
    for k, (c, v) in sorted(agg.items(), key=lambda x: -x[1][1])[:30]:
        print("%-64s %4d %9.3f ms %5.1f%%" % (k[:64], c, v / 1e6, 100 * v / tot))


if __name__ == "__main__":
    main()

"""Joins the raw CSV page of an `ncu --set full` capture of tools/named_launches.py with its label file and prints,
per named launch: duration, DRAM bytes (read + write) next to the ALGORITHMIC bytes and their ratio, achieved GB/s or
TFLOP/s, tensor-pipe activity, L2 hit rate.  Also writes a small JSON with the figures bench.py quotes.
usage: ncu -i rep.ncu-rep --page raw --csv > raw.csv ; python tools/ncu_named_summary.py raw.csv labels.json [out.json]"""
import csv, json, sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
labels = json.load(open(sys.argv[2]))
col = {n: hdr.index(n) for n in hdr}


def val(r, name):
    if name not in col:
        return float("nan")
    try:
        return float(r[col[name]].replace(",", ""))
    except Exception:
        return float("nan")


def scaled(r, name, table):
    v = val(r, name)
    return v * table.get(units[col[name]], 1.0) if name in col else float("nan")


BYTES = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
TIME = {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0, "nsecond": 1e-9, "usecond": 1e-6, "msecond": 1e-3, "second": 1.0}
out = {}
i = 0
print("%-12s %-34s %9s %10s %10s %6s %10s %7s %7s" % ("launch", "kernel", "us", "dram MB", "algo MB", "ratio", "achieved", "tensor%", "L2hit%"))
for label, nk, by, fl in labels:
    tot_t = tot_b = 0.0
    names, tp, l2 = [], [], []
    for r in data[i:i + nk]:
        t = scaled(r, "gpu__time_duration.sum", TIME)
        b = scaled(r, "dram__bytes_read.sum", BYTES) + scaled(r, "dram__bytes_write.sum", BYTES)
        tot_t += t
        tot_b += b
        names.append(r[col["Kernel Name"]].replace("void ", "").replace("b200::", "").split("(")[0][:34])
        tp.append(val(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"))
        l2.append(val(r, "lts__t_sector_hit_rate.pct"))
    i += nk
    ach = "%.0f TF/s" % (fl / tot_t / 1e12) if fl else "%.0f GB/s" % (by / tot_t / 1e9)
    print("%-12s %-34s %9.1f %10.1f %10.1f %6.2f %10s %7.1f %7.1f" % (label, "+".join(sorted(set(names)))[:34], tot_t * 1e6, tot_b / 1e6, by / 1e6,
                                                                 tot_b / by if by else float("nan"), ach, max(tp) if tp else 0, sum(l2) / len(l2) if l2 else 0))
    out[label] = {"kernels": names, "us": tot_t * 1e6, "dram_bytes": tot_b, "algorithmic_bytes": by, "flops": fl}
if len(sys.argv) > 3:
    json.dump({"conv_c_fprop": out.get("c_fprop"), "all": out}, open(sys.argv[3], "w"), indent=1)

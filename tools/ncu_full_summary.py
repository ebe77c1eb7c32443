"""One line per kernel launch of an `ncu --set full` report: duration, DRAM bytes, SM / DRAM throughput, L2 hit rate,
registers, grid, achieved occupancy, IPC.  usage: ncu -i report.ncu-rep --page raw --csv > raw.csv ;
python tools/ncu_full_summary.py raw.csv"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
want = {"name": "Kernel Name", "dur": "gpu__time_duration.sum", "rd": "dram__bytes_read.sum", "wr": "dram__bytes_write.sum",
        "tensor": "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "smthr": "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "dramthr": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l2hit": "lts__t_sector_hit_rate.pct",
        "regs": "launch__registers_per_thread", "grid": "launch__grid_size",
        "occ": "sm__warps_active.avg.pct_of_peak_sustained_active", "ipc": "sm__inst_executed.avg.per_cycle_elapsed"}
idx = {k: (hdr.index(n) if n in hdr else None) for k, n in want.items()}


def f(r, k):
    i = idx[k]
    try:
        return float(r[i].replace(",", ""))
    except Exception:
        return float("nan")


def to_mb(v, unit):
    return v * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(unit, 1.0)


tot = 0.0
for r in data:
    name = r[idx["name"]].replace("void ", "").replace("b200::", "").split("(")[0][:40]
    rd, wr = to_mb(f(r, "rd"), units[idx["rd"]]), to_mb(f(r, "wr"), units[idx["wr"]])
    tot += rd + wr
    print("%-40s %7.1f %s  dram rd %7.2f MB wr %6.2f MB  tensor-pipe %5.1f%%  SM thr %5.1f%%  DRAM thr %5.1f%%  "
          "L2 hit %5.1f%%  regs %3.0f  grid %5.0f  occupancy %5.1f%%  IPC %4.2f"
          % (name, f(r, "dur"), units[idx["dur"]], rd, wr, f(r, "tensor"), f(r, "smthr"), f(r, "dramthr"), f(r, "l2hit"),
             f(r, "regs"), f(r, "grid"), f(r, "occ"), f(r, "ipc")))
print("launches %d   mean DRAM traffic per launch %.1f MB" % (len(data), tot / max(1, len(data))))

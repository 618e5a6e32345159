"""Turns ncu output into the small CSV summaries kept under profiles/.

  python tools/ncu_summary.py launches gpurun_out/launches.csv  > profiles/rNN_launches_<workload>_summary.csv
  ncu -i x.ncu-rep --page raw --csv > raw.csv ; python tools/ncu_summary.py full raw.csv > profiles/rNN_ncu_full_<workload>_summary.csv
"""
import csv
import sys
from collections import OrderedDict

FULL_METRICS = [
    "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "lts__t_sectors_srcunit_tex_op_red.sum",
]


def short(name):
    return name.split("(")[0].replace("void ", "").replace("htgs::", "").strip()


def launches(path):
    rows = [r for r in csv.reader(open(path, errors="replace")) if len(r) > 10]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = OrderedDict()
    for r in rows[1:]:
        v = float(r[vi].replace(",", ""))
        v = v / 1e3 if r[ui] in ("ns", "nsecond") else v * 1e3 if r[ui] in ("ms", "msecond") else v
        a = agg.setdefault(short(r[ki]), [0, 0.0])
        a[0] += 1
        a[1] += v
    ours = {k: v for k, v in agg.items() if not k.startswith("at::") and "elementwise" not in k and "blend_forward_kernel<1>" not in k
            and "blend_forward_kernel<(bool)1>" not in k}
    total = sum(v[1] for v in ours.values())
    print("kernel,launches,avg_us,share_of_listed_time_pct")
    for k, (n, t) in sorted(ours.items(), key=lambda kv: -kv[1][1]):
        print(f"{k},{n},{t / n:.1f},{100 * t / total:.1f}")


def full(path):
    rows = list(csv.reader(open(path, errors="replace")))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    cols = [(m, hdr.index(m)) for m in FULL_METRICS if m in hdr]
    print("kernel," + ",".join(f"{m} [{units[i]}]" for m, i in cols))
    seen = {}
    for r in rows[2:]:
        k = short(r[ki])
        seen[k] = seen.get(k, 0) + 1
        name = k if k != "onesweep_kernel" else f"onesweep_kernel#{seen[k]}"
        print(name + "," + ",".join(r[i].replace(",", "") for _, i in cols))


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])

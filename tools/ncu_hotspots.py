"""Per-region summary of an `ncu --set full --import-source on` capture (the profiles/rNN_ncu_hotspots_*.csv files).

  python tools/ncu_hotspots.py gpurun_out/prof_full.ncu-rep blend_backward blend_forward onesweep ... > profiles/rNN_ncu_hotspots_<workload>.csv

For every named kernel (first captured launch) the SASS listing of ncu's source page is cut into regions: a new
region starts where the per-instruction execution count changes by more than 2 % (loop entries / exits), and
regions holding under 0.15 % of the kernel's instructions are merged into their predecessor.  Per region: the first
instruction, the number of SASS instructions, how often the region ran, its share of executed warp instructions and
its share of the warp-stall samples.
"""
import csv
import io
import subprocess
import sys


def source_rows(rep, kernel):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "-k", f"regex:{kernel}",
                          "--launch-count", "1"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr_at = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hdr_at]
    body = []
    for r in rows[hdr_at + 1:]:
        if not r or r[0] == "Kernel Name" or r[0] == "Address":
            break  # next launch
        body.append(r)
    return rows[0][1].split("(")[0].replace("htgs::", "").replace("void ", ""), hdr, body


def regions(hdr, body):
    a, s, n, x = hdr.index("Address"), hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
    base = int(body[0][a], 16)
    total_x = sum(int(r[x]) for r in body) or 1
    total_n = sum(int(r[n]) for r in body) or 1
    out = []
    for r in body:
        ex, sm = int(r[x]), int(r[n])
        if out and abs(ex - out[-1]["runs"]) <= 0.02 * max(ex, out[-1]["runs"], 1):
            reg = out[-1]
        else:
            reg = dict(addr=int(r[a], 16) - base, first=r[s].strip(), n=0, runs=ex, inst=0, samples=0)
            out.append(reg)
        reg["n"] += 1
        reg["inst"] += ex
        reg["samples"] += sm
    merged = []
    for reg in out:
        if merged and reg["inst"] < 0.0015 * total_x:
            m = merged[-1]
            m["n"] += reg["n"]
            m["inst"] += reg["inst"]
            m["samples"] += reg["samples"]
        else:
            merged.append(reg)
    return merged, total_x, total_n


def main():
    rep, kernels = sys.argv[1], sys.argv[2:]
    print("# ncu --set full --import-source on: share of warp-stall samples and of executed warp instructions per SASS region"
          " (tools/ncu_hotspots.py)")
    print("kernel,region,first_instruction,n_sass,executions_of_region,share_of_instructions_pct,share_of_stall_samples_pct")
    for k in kernels:
        name, hdr, body = source_rows(rep, k)
        regs, tx, tn = regions(hdr, body)
        print(f"# {name}: {tx} warp instructions, {tn} stall samples")
        for r in regs:
            print(f'{name},0x{r["addr"]:x},"{r["first"]}",{r["n"]},{r["runs"]},{100 * r["inst"] / tx:.1f},{100 * r["samples"] / tn:.1f}')


if __name__ == "__main__":
    main()

"""Prints kernel name and gpu__time_duration (us) from an `ncu --csv --metrics gpu__time_duration.sum` log."""
import csv, sys
rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 10]
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
for r in rows[1:]:
    print(f"{r[ki][:60]:60s} {float(r[vi].replace(',', '')) / 1e3:9.2f} us")

"""Print kernel name / grid / duration from an `ncu --metrics gpu__time_duration.sum --csv` log."""
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
h = rows[0]
ki, vi, gi = h.index("Kernel Name"), h.index("Metric Value"), h.index("Grid Size")
for r in rows[1:]:
    print(f"{r[ki][:44]:44s} {r[gi]:>14s} {float(r[vi]) / 1000:8.2f} us")

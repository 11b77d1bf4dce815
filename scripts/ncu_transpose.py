#!/usr/bin/env python
"""`ncu -i X.ncu-rep --page raw --csv` dump -> "metric,unit,value" lines (one launch), the format of
profiles/ncu_full_*.csv:  python scripts/ncu_transpose.py raw.csv > profiles/ncu_full_....csv"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units, vals = rows[0], rows[1], rows[2]
w = csv.writer(sys.stdout)
w.writerow(["metric", "unit", "value"])
for h, u, v in zip(hdr, units, vals):
    w.writerow([h, u, v])

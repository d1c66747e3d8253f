"""Cuts the raw page of an ncu report (ncu -i x.ncu-rep --page raw --csv > raw.csv) down to the columns of a committed capture.
   usage: python profiles/scripts/ncu_filter.py raw.csv profiles/r02_ncu_packed_full_metrics.csv > new.csv
The second file only supplies the column list (its header line)."""
import csv
import sys

raw = list(csv.reader(open(sys.argv[1])))
start = next(i for i, r in enumerate(raw) if r and r[0] == "ID")
keep = next(csv.reader(open(sys.argv[2])))
head = raw[start]
idx = [head.index(k) for k in keep if k in head]
w = csv.writer(sys.stdout, lineterminator="\n")
for r in raw[start:]:
    if len(r) == len(head):
        w.writerow([r[i] for i in idx])

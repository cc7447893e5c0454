"""Top stall-sample SASS instructions of an `ncu --page source --csv --print-source sass` export.
Usage: python tools/ncu_src_top.py file.csv[.gz] [N]"""
import csv
import gzip
import sys

path = sys.argv[1]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 12
op = gzip.open if path.endswith(".gz") else open
rows = list(csv.reader(op(path, "rt")))
name = rows[0][1] if rows and len(rows[0]) > 1 else "?"
data = [r for r in rows[2:] if len(r) > 5 and r[2].isdigit()]
tot = sum(int(r[2]) for r in data)
print(f"# {path}: {name[:110]}")
print(f"# {len(data)} SASS instructions, {tot} stall samples; top {n}:  samples  executed  [index] instruction")
for i, r in sorted(enumerate(data), key=lambda t: -int(t[1][2]))[:n]:
    print("  %6s  %9s  [%4d] %s" % (r[2], r[5], i, r[1].strip()[:90]))

"""Print the last N rows (kernel, microseconds) of an ncu `--metrics gpu__time_duration.sum --csv` launch list."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
h = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
ki, vi = rows[h].index("Kernel Name"), rows[h].index("Metric Value")
for r in rows[h + 1:][-n:]:
    if len(r) > vi:
        print(f"{float(r[vi].replace(',', '')) / 1000:9.2f} us  {r[ki].split('(')[0][-60:]}")

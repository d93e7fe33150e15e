"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: median microseconds per kernel.
    python tools/launch_summary.py gpurun_out/launches.csv [skip_first_n_launches]"""
import collections
import csv
import statistics
import sys

rows = list(csv.reader(open(sys.argv[1])))
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
hdr = rows[hi]
d = collections.OrderedDict()
for r in rows[hi + 1:]:
    if len(r) < len(hdr):
        continue
    rec = dict(zip(hdr, r))
    if int(rec["ID"]) < skip:
        continue
    name = rec["Kernel Name"].split("(")[0].replace("unnamed>::", "").replace("void ", "")[:60]
    d.setdefault(name, []).append(float(rec["Metric Value"].replace(",", "")))
tot = 0.0
for k, v in d.items():
    med = statistics.median(v) / 1e3
    print(f"{k:62s} n={len(v):3d} med={med:8.1f} us")
print("launches per step are the kernels with equal n; sum of medians of repo kernels with n>=3:",
      round(sum(statistics.median(v) / 1e3 for k, v in d.items() if len(v) >= 3), 1), "us")

"""Per-kernel totals of ONE K8 build from an ncu launch list (--metrics gpu__time_duration.sum --csv):
python tools/launch_breakdown.py launches.csv [build_number]"""
import collections
import csv
import sys

lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
per = []
for row in csv.DictReader(lines):
    v = float(row["Metric Value"].replace(",", ""))
    per.append((row["Kernel Name"].split("(")[0], v / 1e3 if row["Metric Unit"] in ("ns", "nsecond") else v))
idx = [i for i, (n, _) in enumerate(per) if "bld_setup" in n]
k = int(sys.argv[2]) if len(sys.argv) > 2 else 0
one = per[idx[k]:idx[k + 1]] if k + 1 < len(idx) else per[idx[k]:]
tot, cnt = collections.defaultdict(float), collections.Counter()
for n, v in one:
    tot[n] += v; cnt[n] += 1
T = sum(tot.values())
for n in sorted(tot, key=lambda q: -tot[q]):
    print("%-28s %4d launches %9.1f us %5.1f%%" % (n, cnt[n], tot[n], 100 * tot[n] / T))
print("total %.1f us over %d launches" % (T, len(one)))
for key in ("split", "scatter", "classify", "heads", "boxes"):
    print(key, [round(v, 1) for n, v in one if key in n])

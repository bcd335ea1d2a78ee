"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: time per kernel name.
usage: python scripts/launch_summary.py <launches.csv> [first_id]"""
import csv, sys, collections, re
rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 10 and r[0].isdigit()]
first = int(sys.argv[2]) if len(sys.argv) > 2 else 0
tot = collections.OrderedDict()
for r in rows:
    if int(r[0]) < first: continue
    name = re.sub(r"\(.*", "", r[4]); name = re.sub(r"<.*", "", name).replace("void ", "")
    t = tot.setdefault(name, [0, 0.0]); t[0] += 1; t[1] += float(r[-1]) / 1e6
s = sum(v[1] for v in tot.values())
for k, (n, ms) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{ms:10.3f} ms {100*ms/s:5.1f}%  x{n:<3d} {k}")
print(f"{s:10.3f} ms total, {len(rows)} launches")

"""python tools/ncu_summary.py launches.csv [skip_first_n_launches_of_each_kernel]
Condenses an `ncu --metrics gpu__time_duration.sum --csv` launch list into one row per kernel: launches, total and mean
duration, share of the total."""
import csv, sys, re, collections
rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
ui = hdr.index("Metric Unit")
tot = collections.OrderedDict()
for r in rows[1:]:
    name = re.sub(r"\(.*", "", r[ki])[:90]
    v = float(r[vi].replace(",", ""))
    v = v / 1e6 if r[ui] in ("ns", "nsecond") else (v / 1e3 if r[ui] in ("us", "usecond") else v)
    t = tot.setdefault(name, [0, 0.0])
    t[0] += 1; t[1] += v
s = sum(v[1] for v in tot.values())
print("%-92s %6s %10s %9s %6s" % ("kernel", "n", "total ms", "mean ms", "share"))
for k, (n, v) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print("%-92s %6d %10.3f %9.4f %5.1f%%" % (k, n, v, v / n, 100 * v / s))
print("%-92s %6s %10.3f" % ("TOTAL", "", s))

"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list: count, total, average, share."""
import collections, csv, re, sys
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith("==")) if len(r) > 14 and r[0].isdigit()]
tot = collections.Counter(); n = collections.Counter()
for r in rows:
    name = re.sub(r"\(.*", "", r[4]).split("::")[-1]
    name = re.sub(r"<.*", "", name)
    tot[name] += float(r[14]) / 1e3
    n[name] += 1
all_us = sum(tot.values())
print("%-44s %7s %12s %10s %8s" % ("kernel", "n", "total us", "avg us", "share"))
for k, v in tot.most_common():
    print("%-44s %7d %12.1f %10.1f %8.3f" % (k[:44], n[k], v, v / n[k], v / all_us))

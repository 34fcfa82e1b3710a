"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel: launches, total us, share."""
import collections
import csv
import re
import sys


def main(path, last=None):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    for i, r in enumerate(rows):
        if r[0] == "ID":
            hdr, rows = r, rows[i + 1:]
            break
    if last:
        rows = rows[-int(last):]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows:
        n = re.sub(r"\(.*", "", r[ki]).replace("vsp::<unnamed>::", "").replace("vsp::", "").replace("void ", "")
        agg[n][0] += 1
        agg[n][1] += float(r[vi].replace(",", "")) / 1e3
    tot = sum(v[1] for v in agg.values())
    print("# %s: %d launches, %.1f us total" % (path, len(rows), tot))
    print("# kernel, launches, total us, share")
    for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print("%-52s %5d %12.1f %5.1f%%" % (n, c, t, t / tot * 100))


if __name__ == "__main__":
    main(*sys.argv[1:])

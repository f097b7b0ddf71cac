"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (count, total ms, share)."""
import collections, csv, sys
def main(path, title):
    rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
    h = rows[0]; ki = h.index("Kernel Name"); vi = h.index("Metric Value"); ui = h.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        t = float(r[vi].replace(",", "")); u = r[ui]
        t = t / 1e6 if u in ("ns", "nsecond") else (t / 1e3 if u in ("us", "usecond") else t)
        a = agg.setdefault(r[ki], [0, 0.0]); a[0] += 1; a[1] += t
    tot = sum(a[1] for a in agg.values())
    print("# " + title + "\n")
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        print("%-110s n=%4d %11.3f ms %6.1f%%" % (k[:110], a[0], a[1], 100 * a[1] / tot))
if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else sys.argv[1])

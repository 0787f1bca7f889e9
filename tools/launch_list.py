"""Summary of an `ncu --metrics gpu__time_duration.sum --csv` launch list: kernel time per kernel name and its share.
    python tools/launch_list.py launches.csv [skip_fraction]   (skip_fraction: leading part of the list to drop, e.g. warm-up)"""
import collections
import csv
import sys


def main(path, skip=0.0):
    rows = list(csv.reader(open(path, newline="")))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    col = {n: i for i, n in enumerate(rows[hi])}
    rec = []
    for r in rows[hi + 1:]:
        if len(r) <= col["Metric Value"]:
            continue
        v, u = float(r[col["Metric Value"]].replace(",", "")), r[col["Metric Unit"]]
        v *= {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3}.get(u, 1.0)
        rec.append((r[col["Kernel Name"]], v))
    rec = rec[int(len(rec) * skip):]
    agg = collections.OrderedDict()
    for k, v in rec:
        k = k.split("(")[0].replace("void ", "")[:70]
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print(f"{len(rec)} launches, {tot:.1f} us of kernel time (cold-cache, serialised: shares are meaningful, absolutes are not)")
    print(f"| kernel | launches | total us | avg us | share |\n|---|---|---|---|---|")
    for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"| `{k}` | {n} | {t:.1f} | {t / n:.2f} | {100 * t / tot:.1f} % |")


if __name__ == "__main__":
    main(sys.argv[1], float(sys.argv[2]) if len(sys.argv) > 2 else 0.0)

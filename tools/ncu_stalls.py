"""Where the warps of one kernel spend their time: warp-state samples of an `ncu --page source --csv --print-source sass`
export, grouped by the mbarrier wait / instruction they sit on.  Usage: ncu_stalls.py file.source.csv [min samples]"""
import csv
import sys


def main(path, floor=300):
    rows = list(csv.reader(open(path, newline="")))
    hdr = rows[1]
    col = {n: i for i, n in enumerate(hdr)}
    data = rows[2:]
    S = lambda r: int(r[col["# Samples"]] or 0)
    reasons = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    total = sum(S(r) for r in data)
    print("total samples", total)
    agg = {h: sum(int(r[col[h]] or 0) for r in data) for h in reasons}
    print("by reason:", ", ".join(f"{h[6:]} {v}" for h, v in sorted(agg.items(), key=lambda x: -x[1])[:8]))
    for i, r in enumerate(data):
        if S(r) >= floor:
            top = sorted(((int(r[col[h]] or 0), h[6:]) for h in reasons), reverse=True)[:2]
            print(f"{i:5d} {S(r):7d} {100.0 * S(r) / total:5.1f}%  x{r[col['Instructions Executed']]:>9s}  {r[1].strip()[:72]:72s} {top}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 300)

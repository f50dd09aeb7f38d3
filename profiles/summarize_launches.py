"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel and per phase
(phases are delimited by embed_kernel, the first kernel of every forward pass)."""
import collections
import csv
import re
import sys


def main(path, header=""):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    rows = list(csv.DictReader(lines))
    names = [re.sub(r"\(.*", "", x["Kernel Name"]) for x in rows]
    idx = [i for i, n in enumerate(names) if "embed_kernel" in n] + [len(rows)]
    if header:
        print(header)
    print("# per-launch times are cold-cache and serialised: compare SHARES, not absolutes\n")
    for p in range(len(idx) - 1):
        a, b = idx[p], idx[p + 1]
        agg = collections.defaultdict(lambda: [0, 0.0])
        for row, n in zip(rows[a:b], names[a:b]):
            v = float(row["Metric Value"].replace(",", ""))
            u = row["Metric Unit"]
            v = v / 1000 if u == "ns" else v * 1000 if u == "ms" else v
            agg[n][0] += 1
            agg[n][1] += v
        tot = sum(v[1] for v in agg.values())
        print(f"--- forward pass {p}: {b - a} launches, total {tot:.1f} us")
        for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            print(f"{t:10.1f} us {100 * t / tot:5.1f}%  x{c:4d}  avg {t / c:8.2f} us  {n[:100]}")
        print()


if __name__ == "__main__":
    main(sys.argv[1], " ".join(sys.argv[2:]))

"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, mean us, share."""
import csv
import re
import sys
from collections import OrderedDict


def main(path, out=None):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        us = v / 1e3 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1e3)
        name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "").strip()
        rows.append((name, us, r["Grid Size"], r["Block Size"]))
    agg = OrderedDict()
    for n, us, g, b in rows:
        a = agg.setdefault(n, [0, 0.0, g, b])
        a[0] += 1; a[1] += us
    total = sum(a[1] for a in agg.values())
    lines = [f"# {path}: {len(rows)} launches, {total:.1f} us total (cold-cache, serialised: compare SHARES)",
             f"{'kernel':70s} {'n':>5s} {'mean_us':>9s} {'share':>7s}  grid block"]
    for n, (c, t, g, b) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        lines.append(f"{n[:70]:70s} {c:5d} {t / c:9.2f} {t / total:7.1%}  {g} {b}")
    text = "\n".join(lines)
    print(text)
    if out:
        open(out, "w").write(text + "\n")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else None)

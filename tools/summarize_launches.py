"""Per-kernel summary of an `ncu --metrics gpu__time_duration.sum --csv` launch list.
Usage: python tools/summarize_launches.py gpurun_out/launches.csv [--skip-torch]"""
import collections
import csv
import re
import sys


def short(name):
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\(.*$", "", name)
    return name if len(name) < 90 else name[:87] + "..."


def main():
    path = sys.argv[1]
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    head = rows[0]
    k_name, k_val, k_grid, k_block = (head.index(x) for x in ("Kernel Name", "Metric Value", "Grid Size", "Block Size"))
    agg = collections.OrderedDict()
    for r in rows[1:]:
        n = short(r[k_name])
        a = agg.setdefault(n, dict(count=0, ns=0.0, grid=r[k_grid], block=r[k_block]))
        a["count"] += 1
        a["ns"] += float(r[k_val].replace(",", ""))
    ours = {n: a for n, a in agg.items() if n.startswith("jtsm::")}
    total = sum(a["ns"] for a in ours.values())
    print("| kernel | launches | avg us | share of our kernels | grid x block |")
    print("|---|---|---|---|---|")
    for n, a in sorted(ours.items(), key=lambda kv: -kv[1]["ns"]):
        print("| `%s` | %d | %.1f | %.1f %% | %s x %s |" % (n, a["count"], a["ns"] / a["count"] / 1e3,
                                                            100 * a["ns"] / total, a["grid"], a["block"]))
    other = sum(a["ns"] for n, a in agg.items() if n not in ours)
    print("\nother kernels in the capture (torch copies / fills / flush): %d launches, %.1f us in total"
          % (sum(a["count"] for n, a in agg.items() if n not in ours), other / 1e3))


if __name__ == "__main__":
    main()

"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per CUDA kernel (markdown table).

    python tools/launch_list_summary.py gpurun_out/launches.csv > profiles/rNN_launches.md

ncu times are cold-cache and serialised: compare SHARES with bench.py's kernel_breakdown, not absolutes.
"""
import collections
import csv
import re
import sys


def main() -> None:
    path = sys.argv[1]
    rows = list(csv.DictReader(l for l in open(path) if l.startswith('"')))
    agg: dict[str, list] = collections.OrderedDict()
    for r in rows:
        k = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "")[:100]
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += float(r["Metric Value"]) / 1e6
    tot = sum(a[1] for a in agg.values())
    print(f"{len(rows)} launches, {tot:.2f} ms summed device time (under ncu)\n")
    print("| kernel | launches | ms (sum) | share |")
    print("|---|---:|---:|---:|")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:40]:
        print(f"| `{k}` | {a[0]} | {a[1]:.3f} | {a[1] / tot:.3f} |")


if __name__ == "__main__":
    main()

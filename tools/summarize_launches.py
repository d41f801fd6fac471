"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list into a markdown table (dev tool)."""
import collections
import csv
import re
import sys


def main(path, title):
    lines = [l for l in open(path) if not l.startswith("==")]
    tot = collections.defaultdict(lambda: [0.0, 0])
    for row in csv.DictReader(lines):
        try:
            v = float(row["Metric Value"].replace(",", ""))
        except (ValueError, KeyError):
            continue
        unit = row["Metric Unit"]
        ms = v / 1e6 if unit in ("ns", "nsecond") else (v / 1e3 if unit in ("us", "usecond") else v)
        name = re.sub(r"\(.*", "", re.sub(r"<.*", "", row["Kernel Name"]))[:80]
        tot[name][0] += ms
        tot[name][1] += 1
    T = sum(v[0] for v in tot.values())
    print(f"# {title}\n\nTotal {T:.1f} ms over {sum(v[1] for v in tot.values())} launches.\n")
    print("| ms | share | launches | kernel |\n|---:|---:|---:|---|")
    for k, v in sorted(tot.items(), key=lambda kv: -kv[1][0])[:30]:
        print(f"| {v[0]:.3f} | {100 * v[0] / T:.1f}% | {v[1]} | `{k}` |")


main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else sys.argv[1])

#!/usr/bin/env python
"""Developer tool: summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel.

    python tools/launch_summary.py profiles/r02_launch_list_c2_batch16.csv > profiles/r02_launch_list_c2_batch16_summary.md
"""
import collections
import csv
import re
import sys


def main():
    rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
    hdr = rows[0]
    ix = {h: i for i, h in enumerate(hdr)}
    tot = collections.Counter()
    cnt = collections.Counter()
    for r in rows[1:]:
        if r[ix["Metric Name"]] != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", r[ix["Kernel Name"]]).replace("void ", "").replace("(int)", "").replace("(bool)", "")
        unit, v = r[ix["Metric Unit"]], float(r[ix["Metric Value"]].replace(",", ""))
        ms = v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(unit, 1e-6)
        tot[name] += ms
        cnt[name] += 1
    total = sum(tot.values())
    print(f"{sum(cnt.values())} launches captured, {total:.1f} ms of kernel time in total.\n")
    print("| kernel | launches | total ms | share |\n|---|---:|---:|---:|")
    for k, v in tot.most_common():
        print(f"| `{k}` | {cnt[k]} | {v:.3f} | {v / total:.3f} |")


if __name__ == "__main__":
    main()

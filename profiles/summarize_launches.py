"""Kernel shares from an `ncu --metrics gpu__time_duration.sum --csv` launch list.

    python profiles/summarize_launches.py gpurun_out/r01_launches_ur5e.csv "command line" > profiles/r01_launches_ur5e_summary.txt
"""

import collections
import csv
import sys


def main(path, cmd):
    lines = [ln for ln in open(path) if not ln.startswith("==")]
    rows = list(csv.DictReader(lines))
    tot = collections.Counter()
    cnt = collections.Counter()
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        ms = v * {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "nsecond": 1e-6, "ms": 1.0, "msecond": 1.0, "s": 1e3, "second": 1e3}[unit]
        tot[r["Kernel Name"]] += ms
        cnt[r["Kernel Name"]] += 1
    total = sum(tot.values())
    print(f"ncu --metrics gpu__time_duration.sum --clock-control none -c 400, {cmd}")
    print("(per-launch times are cold-cache and serialised: compare shares)")
    for k, ms in tot.most_common():
        print(f"{100 * ms / total:5.1f}%  {cnt[k]:4d} launches  {ms:9.3f} ms  {k[:64]}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "")

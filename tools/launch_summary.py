"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list.
Usage: python tools/launch_summary.py launches.csv out.txt "header line" """
import collections
import csv
import sys


def main():
    rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 5]
    hdr = next(r for r in rows if "Kernel Name" in r)
    col = {h: i for i, h in enumerate(hdr)}
    tot, cnt = collections.Counter(), collections.Counter()
    for r in rows:
        if r is hdr or len(r) != len(hdr) or r[col["Metric Name"]] != "gpu__time_duration.sum":
            continue
        name = r[col["Kernel Name"]].split("(")[0].strip()
        v = float(r[col["Metric Value"]].replace(",", ""))
        unit = r[col["Metric Unit"]]
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1.0)
        tot[name] += v
        cnt[name] += 1
    allt = sum(tot.values())
    lines = [sys.argv[3] if len(sys.argv) > 3 else "", ""]
    for name, t in tot.most_common():
        lines.append("%-60s launches %4d  total %9.1f us  mean %8.2f us  share %5.1f%%" % (name[:60], cnt[name], t, t / cnt[name], 100 * t / allt))
    open(sys.argv[2], "w").write("\n".join(lines) + "\n")
    print("\n".join(lines[:14]))


if __name__ == "__main__":
    main()

"""Per-kernel totals of an ncu launch list (`ncu --metrics gpu__time_duration.sum --csv --log-file X.csv ...`).

    python tools/launches_summary.py gpurun_out/launches_r02.csv "<the command that was profiled>" > profiles/launches_r02_summary.txt
"""
import csv
import re
import sys
from collections import OrderedDict


def main(path, command=""):
    rows = list(csv.reader(open(path)))
    hdr = next(r for r in rows if r and r[0] == "ID")
    agg = OrderedDict()
    for r in rows[rows.index(hdr) + 1:]:
        if len(r) < len(hdr):
            continue
        d = dict(zip(hdr, r))
        if d["Metric Name"] != "gpu__time_duration.sum":
            continue
        v = float(d["Metric Value"].replace(",", ""))
        v *= {"ns": 1e-3, "nsecond": 1e-3, "us": 1, "usecond": 1, "ms": 1e3, "msecond": 1e3}.get(d["Metric Unit"], 1)
        m = re.search(r"kvdp_\w+(<[^>]*>)?", d["Kernel Name"])
        name = m.group(0) if m else "other (torch / library): " + d["Kernel Name"][:40]
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print("# ncu launch list: %s" % command)
    print("# per-launch times under ncu are cold-cache and serialised: compare SHARES, not absolutes.")
    print("%-64s %6s %12s %10s %7s" % ("kernel", "count", "total_us", "avg_us", "share"))
    for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-64s %6d %12.1f %10.1f %6.1f%%" % (k[:64], c, t, t / c, 100 * t / tot))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "")

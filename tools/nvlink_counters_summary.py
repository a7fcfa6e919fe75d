"""Turns the ncu CSV of `tools/nvlink_probe.py launches` (NVLink byte counters per kernel launch) into the table kept
under profiles/: payload vs bytes on the wire, per direction, with the wire rate each launch ran at.

    python tools/nvlink_counters_summary.py gpurun_out/nvlink_counters.csv > profiles/nvlink_counters_r02.txt
"""
import csv
import sys
from collections import OrderedDict


def main(path):
    rows = list(csv.reader(open(path)))
    hdr = next(r for r in rows if r and r[0] == "ID")
    data = rows[rows.index(hdr) + 1:]
    by, names = OrderedDict(), {}
    for r in data:
        if len(r) < len(hdr):
            continue
        d = dict(zip(hdr, r))
        by.setdefault(d["ID"], {})[d["Metric Name"]] = float(d["Metric Value"].replace(",", ""))
        names[d["ID"]] = d["Kernel Name"]
    print("NVLink byte counters per launch (ncu --metrics nvltx__bytes*, nvlrx__bytes*; GPU 0 runs the kernel, GPU 1 is the peer)")
    print("user = payload bytes, protocol = packet headers / requests / acks, wire = user + protocol; rates = bytes / kernel duration under ncu")
    print()
    print("%-3s %-28s %9s | %12s %12s %7s %8s | %12s %12s %7s %8s" %
          ("id", "kernel", "dur_us", "tx_user_MB", "tx_proto_MB", "tx_w/u", "tx_GB/s", "rx_user_MB", "rx_proto_MB", "rx_w/u",
           "rx_GB/s"))
    for k, m in by.items():
        name = names[k].split("::")[-1].split("(")[0][:28]
        dur = m["gpu__time_duration.sum"] / 1e9
        txu, txp = m["nvltx__bytes_data_user.sum"], m["nvltx__bytes_data_protocol.sum"]
        rxu, rxp = m["nvlrx__bytes_data_user.sum"], m["nvlrx__bytes_data_protocol.sum"]
        print("%-3s %-28s %9.1f | %12.1f %12.1f %7s %8.1f | %12.1f %12.1f %7s %8.1f" %
              (k, name, dur * 1e6, txu / 1e6, txp / 1e6, ("%.4f" % ((txu + txp) / txu)) if txu else "-",
               (txu + txp) / dur / 1e9, rxu / 1e6, rxp / 1e6, ("%.4f" % ((rxu + rxp) / rxu)) if rxu else "-",
               (rxu + rxp) / dur / 1e9))


if __name__ == "__main__":
    main(sys.argv[1])

#!/usr/bin/env python3
"""Transpose `ncu -i X.ncu-rep --page raw --csv` (one very wide row per kernel) into `metric,unit,value` lines.

    ncu -i gpurun_out/k.ncu-rep --page raw --csv | python tools/ncu_keyvals.py [regex] > profiles/k_raw.csv
"""
import csv
import re
import sys


def main():
    pat = re.compile(sys.argv[1]) if len(sys.argv) > 1 else None
    rows = list(csv.reader(sys.stdin))
    if len(rows) < 3:
        sys.exit("expected a header row, a unit row and at least one kernel row")
    hdr, units = rows[0], rows[1]
    w = csv.writer(sys.stdout)
    for k, vals in enumerate(rows[2:]):
        for h, u, v in zip(hdr, units, vals):
            if pat is None or pat.search(h):
                w.writerow([k, h, u, v])


if __name__ == "__main__":
    main()

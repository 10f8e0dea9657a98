#!/usr/bin/env python
"""Summarise an ncu launch list (--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
--csv) per kernel: launches, duration of the LAST launch (warm allocator, steady state), DRAM bytes and
the DRAM GB/s they imply.  usage: tools/launch_summary.py launches.csv [title] > profiles/xxx.txt"""
import collections
import csv
import sys

UNIT_T = {"ns": 1e-9, "us": 1e-6, "usecond": 1e-6, "ms": 1e-3, "msecond": 1e-3, "s": 1.0, "second": 1.0, "nsecond": 1e-9}
UNIT_B = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    path = sys.argv[1]
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = rows[0]
    ki, mi, vi, ii, ui = (hdr.index(x) for x in ("Kernel Name", "Metric Name", "Metric Value", "ID", "Metric Unit"))
    per = collections.OrderedDict()
    for r in rows[1:]:
        per.setdefault(r[ii], {"k": r[ki]})[r[mi]] = (float(r[vi].replace(",", "")), r[ui])
    agg = collections.OrderedDict()
    for v in per.values():
        agg.setdefault(v["k"].split("(")[0][:64], []).append(v)
    if len(sys.argv) > 2:
        print("# " + " ".join(sys.argv[2:]))
    print("# per kernel: launches, last launch's duration, DRAM read / write bytes, implied DRAM GB/s")
    print("# (ncu --clock-control none; launches are serialised and cold-cache: compare shares, not absolutes)")
    total = 0.0
    out = []
    for k, launches in agg.items():
        v = launches[-1]
        t = v["gpu__time_duration.sum"]
        sec = t[0] * UNIT_T.get(t[1], 1e-9)
        rd = v.get("dram__bytes_read.sum", (0.0, "byte"))
        wr = v.get("dram__bytes_write.sum", (0.0, "byte"))
        rb, wb = rd[0] * UNIT_B.get(rd[1], 1.0), wr[0] * UNIT_B.get(wr[1], 1.0)
        out.append((k, len(launches), sec, rb, wb))
        if not k.startswith("void at::"):
            total += sec
    for k, n, sec, rb, wb in out:
        share = 100.0 * sec / total if total and not k.startswith("void at::") else 0.0
        print(f"{k:64s} n={n:3d} t={sec * 1e6:10.1f} us ({share:5.1f} %)  rd={rb / 1e6:9.1f} MB  wr={wb / 1e6:9.1f} MB"
              f"  {(rb + wb) / sec / 1e9:8.1f} GB/s")


if __name__ == "__main__":
    main()

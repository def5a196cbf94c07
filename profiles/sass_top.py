#!/usr/bin/env python
"""Prints the instructions of an .ncu-rep capture (--set full --import-source on) that hold the most
warp-stall samples.  usage: python profiles/sass_top.py report.ncu-rep [rows=30] ["title"]"""
import csv
import io
import subprocess
import sys

path = sys.argv[1]
rows_wanted = int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
kernel, head, data = rows[0][1], rows[1], rows[2:]
ia, isrc = head.index("Address"), head.index("Source")
isamp, iex = head.index("Warp Stall Sampling (All Samples)"), head.index("Instructions Executed")
total = sum(int(r[isamp]) for r in data)
executed = sum(int(r[iex]) for r in data)
base = int(data[0][ia], 16)
if len(sys.argv) > 3:
    print(sys.argv[3])
print(kernel)
print(f"total stall samples {total}, warp instructions executed {executed}")
print("offset   samples  share  executed  instruction")
for r in sorted(data, key=lambda r: -int(r[isamp]))[:rows_wanted]:
    print(f"{int(r[ia], 16) - base:05x} {int(r[isamp]):7d} {100 * int(r[isamp]) / total:5.1f}% {r[iex]:>9s}  {r[isrc].strip()}")

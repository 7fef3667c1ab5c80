"""Debug tool: launches longer than a threshold (us) from an ncu gpu__time_duration CSV log. usage: ncu_big_launches.py log.csv [min_us]"""
import csv, io, sys
txt = [l for l in open(sys.argv[1]).read().splitlines() if not l.startswith("==")]
rows = list(csv.DictReader(io.StringIO("\n".join(txt))))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 300.0
for r in rows:
    us = float(r["Metric Value"]) / 1e3
    if us >= thr:
        print("%-60s grid %-14s %8.1f us" % (r["Kernel Name"][:60], r["Grid Size"], us))

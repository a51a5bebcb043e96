"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: one line per launch."""
import csv, re, sys
rows = list(csv.DictReader(l for l in open(sys.argv[1]) if l.startswith('"')))
tot = 0
for r in rows:
    n = re.sub(r'\(.*', '', r['Kernel Name']).replace('void ', '').replace('isg::', '')[:48]
    us = float(r['Metric Value']) / 1000
    tot += us
    print(f"{r['ID']:>4} {n:<48} grid {r['Grid Size']:<16} blk {r['Block Size']:<13} {us:9.1f} us")
print("total", tot, "us over", len(rows), "launches")

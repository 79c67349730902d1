"""Join `nvdisasm -g -c` output (SASS with '//## File "...", line N' markers) with `ncu --page source --csv` rows
(same instruction order) and print warp-instructions executed per source line / per line range, divided by `per`
(e.g. the number of dialogues of the launch).  Usage: sass_by_line.py SASS NCU_CSV PER [FILE]"""
import collections
import csv
import re
import sys

sass_file, ncu_csv, per = sys.argv[1], sys.argv[2], float(sys.argv[3])
only = sys.argv[4] if len(sys.argv) > 4 else None
cur = None
ins = []
for l in open(sass_file):
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split('/')[-1], int(m.group(2)))
        continue
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m:
        ins.append((int(m.group(1), 16), cur, m.group(2)))
rows = list(csv.reader(open(ncu_csv)))
hdr = rows[1]
body = rows[2:]
ie = hdr.index("Instructions Executed")
assert len(body) == len(ins), (len(body), len(ins))
by = collections.Counter()
tot = 0.0
for (off, line, txt), r in zip(ins, body):
    e = int(r[ie]) / per
    tot += e
    by[line] += e
print("total per unit %.1f" % tot)
for line, e in sorted(by.items(), key=lambda kv: (kv[0] or ("", 0))):
    if e >= 0.5 and (only is None or (line and line[0] == only)):
        print("%-18s %5d %8.1f" % (line[0] if line else "?", line[1] if line else 0, e))

import csv, re, sys, collections
sass_file, ncu_csv, lo, hi, per = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), float(sys.argv[5])
cur=None; ins=[]
for l in open(sass_file):
    m=re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur=(m.group(1).split('/')[-1], int(m.group(2))); continue
    m=re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m: ins.append((int(m.group(1),16), cur, m.group(2)))
rows=list(csv.reader(open(ncu_csv)))
hdr=rows[1]; body=rows[2:]
ie=hdr.index("Instructions Executed")
tot=0
for (off,line,txt),r in zip(ins,body):
    e=int(r[ie])
    if line and line[0]=='fd_rows.cu' and lo<=line[1]<=hi or (len(sys.argv)>6 and off>=int(sys.argv[6],16) and off<=int(sys.argv[7],16)):
        print(f"{off:05x} {e/per:7.2f} L{line[1] if line else 0:4d} {line[0][:14] if line else '':14s} {txt[:90]}")
        tot+=e
print("sum per unit", tot/per)

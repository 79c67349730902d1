import csv, re, sys, collections
sass_file, ncu_csv, per = sys.argv[1], sys.argv[2], float(sys.argv[3])
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
# aggregate by address ranges: print cumulative by offset buckets with dominant line
bucket=collections.OrderedDict()
for (off,line,txt),r in zip(ins,body):
    e=int(r[ie])/per
    key=off//0x200
    b=bucket.setdefault(key,[0,collections.Counter()])
    b[0]+=e; 
    if line: b[1][line]+=e
for k,(e,c) in bucket.items():
    top=", ".join(f"{l[0][:8]}:{l[1]}({v:.0f})" for l,v in c.most_common(4))
    print(f"{k*0x200:05x} {e:7.1f}  {top}")

"""List, per dialogue-kernel instantiation in the built library, the SASS instructions that show the Blackwell/Hopper-class
data path (bulk TMA copies UBLKCP, mbarrier transaction ops SYNCS.*TRANS64, async-proxy fences) and print one excerpt.
Usage: sass_evidence.py > profiles/r02_sass_k_main.txt   (runs cuobjdump on csrc/libfdb200.so)"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "fraud-detection-spark-kafka-llm_b200", "csrc", "libfdb200.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
arch = re.findall(r"arch = (sm_\w+)", sass)
print("library:", os.path.relpath(lib, ROOT), "| code objects:", sorted(set(arch)))
want = ["UBLKCP", "SYNCS.ARRIVE.TRANS64", "SYNCS.PHASECHK.TRANS64", "SYNCS.EXCH", "FENCE.VIEW.ASYNC", "ATOMS", "REDUX", "LDG.E.ENL2.256", "LDG.E.256"]
cur, counts, excerpt = None, collections.OrderedDict(), {}
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    for w in want:
        if w in line:
            counts[cur][w] += 1
            if w == "UBLKCP" and cur not in excerpt:
                excerpt[cur] = line.strip()
names = subprocess.run(["c++filt"] + list(counts), capture_output=True, text=True).stdout.splitlines()
for (fn, c), dn in zip(counts.items(), names):
    if "k_main" in fn or "k_rows" in fn or "k_wide" in fn:
        print(f"{dn[:90]:90s} " + " ".join(f"{w}={c[w]}" for w in want if c[w]))
print()
print("excerpt (first bulk copy of the first kernel that has one):")
for fn, l in list(excerpt.items())[:1]:
    print(" ", fn)
    print("   ", l)

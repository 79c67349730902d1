"""Regenerate profiles/traffic.json from an ncu capture of k_main<DT, clean> (one launch over ROWS dialogues).
The file records the hash of the kernel sources it was taken from; bench.py quotes it only while that hash matches.
Usage: update_traffic.py REPORT.ncu-rep ROWS CAPTURE_NAME"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

rep, rows, name = sys.argv[1], float(sys.argv[2]), sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(raw.splitlines()))
hdr, units, vals = r[0], r[1], r[2]


def get(metric):
    i = hdr.index(metric)
    v = float(vals[i].replace(",", ""))
    u = units[i].lower()
    scale = {"byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1.0)
    return v * scale


out = {"capture": name, "kernel": vals[hdr.index("Kernel Name")], "rows": rows, "kernel_source_sha": bench.kernel_source_hash(),
       "k_main_dt_clean_bytes_per_dialogue": (get("dram__bytes_read.sum") + get("dram__bytes_write.sum")) / rows,
       "k_main_dt_clean_warp_instructions_per_dialogue": get("smsp__inst_executed.sum") / rows}
with open(os.path.join(ROOT, "profiles", "traffic.json"), "w") as f:
    json.dump(out, f, indent=1)
print(out)

"""profiles/traffic.json from an `ncu --page raw --csv` export of one field_tile_kernel launch:
dram__bytes_read.sum + dram__bytes_write.sum (bench.py's roofline.traffic reads this file).
usage: python tools/ncu_traffic.py <raw.csv> <fields in that launch> <tag of the profiles/ files>"""
import csv
import json
import os
import sys

raw, fields, tag = sys.argv[1], int(sys.argv[2]), sys.argv[3]
rows = list(csv.reader(open(raw)))
hdr = rows[0]
units = rows[1]
best = None
for r in rows[2:]:
    if len(r) == len(hdr) and "field_tile_kernel" in r[hdr.index("Kernel Name")]:
        best = r
assert best is not None, "no field_tile_kernel launch in " + raw


def val(name):
    i = hdr.index(name)
    v = float(best[i].replace(",", ""))
    u = units[i].lower()
    return v * {"byte": 1., "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1.)


rd, wr = val("dram__bytes_read.sum"), val("dram__bytes_write.sum")
out = {"kernel": "field_tile_kernel", "dram_bytes_read": rd, "dram_bytes_write": wr,
       "dram_bytes_per_launch": rd + wr, "fields_per_launch": fields,
       "source": "profiles/%s (ncu --set full, one timed launch)" % tag}
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
json.dump(out, open(os.path.join(root, "profiles", "traffic.json"), "w"), indent=1)
print(out)
